// Shared parameters: the public key-value API of the parameter server (CPU plumbing path).
//   SharedParameter<K>        push / pull / sync + the request handler        (shared_parameter.h)
//   FreqencyFilter<K,V>       CountMin tail-feature filter                    (frequency_filter.h)
//   KVVector<K,V>             sorted keys + aligned values per channel        (kv_vector.h)
//   KVBufferedVector<K,V>     received values buffered per timestamp          (kv_buffered_vector.h)
//   KVMap<K,V>                hash map with per-key operators                 (kv_map.h)
//   KVStore<K,V,Entry,State>  hash map of optimizer entries                   (kv_store.h)
// ref src/parameter/*.h.  The device counterparts live in parameter_server_b200/parallel/kv.py
// (sharded device arrays, peer-memory push / pull kernels).
//
// Replication (setReplica / getReplica / recoverFrom) is a no-op in the reference
// (shared_parameter.h:68-79, kv_map.h:38-40); KVMap and KVVector implement it here: with
// -num_replicas r every applied push is forwarded to the r servers of kReplicaGroup, which
// keep a per-owner copy that a replacement node can pull back with recover().
#pragma once
#include <fstream>

#include "system/heartbeat.h"
#include "system/runtime.h"
#include "util/localizer.h"
#include "util/sketch.h"

namespace PS {

DECLARE_int32(num_workers);
DECLARE_int32(num_replicas);
DECLARE_string(app_name);

template <typename K, typename V>
class FreqencyFilter {
 public:
  void insertKeys(const SArray<K>& key, const SArray<V>& count) {
    CHECK_EQ(key.size(), count.size());
    for (size_t i = 0; i < key.size(); ++i) count_.insert(key[i], count[i]);
  }
  // keys whose estimated frequency exceeds `freqency`
  SArray<K> queryKeys(const SArray<K>& key, int freqency) {
    CHECK_LT(freqency, 255) << "the sketch counters are 8 bit";
    SArray<K> kept;
    for (K k : key)
      if ((int)count_.query(k) > freqency) kept.pushBack(k);
    return kept;
  }
  bool empty() { return count_.empty(); }
  void resize(int n, int k) { count_.resize(n, k, (V)254); }
  void clear() { count_.clear(); }

 private:
  CountMin<K, V> count_;
};

template <typename V>
void compAssOp(V& right, V left, Operator op) {
  switch (op) {
    case PLUS: right += left; break;
    case MINUS: right -= left; break;
    case TIMES: right *= left; break;
    case DIVIDE: right /= left; break;
    default: break;
  }
}
template <typename V>
void compAssOpInt(V& right, V left, Operator op) {
  switch (op) {
    case AND: right &= left; break;
    case OR: right |= left; break;
    case XOR: right ^= left; break;
    default: compAssOp(right, left, op); break;
  }
}
inline void compAssOp(int32& r, int32 l, Operator op) { compAssOpInt(r, l, op); }
inline void compAssOp(int64& r, int64 l, Operator op) { compAssOpInt(r, l, op); }
inline void compAssOp(uint32& r, uint32 l, Operator op) { compAssOpInt(r, l, op); }
inline void compAssOp(uint64& r, uint64 l, Operator op) { compAssOpInt(r, l, op); }

template <typename K>
class SharedParameter : public Customer {
 public:
  SharedParameter(const string& my_name, const string& parent_name)
      : Customer(my_name, parent_name) {}
  virtual ~SharedParameter() {}

  int sync(MessagePtr msg) {
    CHECK(msg->task.shared_para().has_cmd()) << msg->debugString();
    if (!msg->task.has_key_range()) msg->task.mutable_key_range().from(Range<K>::all());
    return port(msg->recver)->submit(msg);
  }
  int push(MessagePtr msg) {
    set(msg).set_cmd(CallSharedParaView::PUSH);
    return sync(msg);
  }
  int pull(MessagePtr msg) {
    set(msg).set_cmd(CallSharedParaView::PULL);
    return sync(msg);
  }
  void waitInMsg(const NodeID& node, int time) { port(node)->waitIncomingTask(time); }
  void waitOutMsg(const NodeID& node, int time) { port(node)->waitOutgoingTask(time); }
  void finish(const NodeID& node, int time) { port(node)->finishIncomingTask(time); }

  void setTailFilterIgnoreChl(bool flag) { key_filter_ignore_chl_ = flag; }
  void setTailFilterSize(int chl, int n, int k) { key_filter_[chl].resize(n, k); }
  void clearTailFilter(int chl) { key_filter_[key_filter_ignore_chl_ ? 0 : chl].clear(); }

  static CallSharedParaView set(MessagePtr msg) {
    msg->task.set_type(TaskView::CALL_CUSTOMER);
    return msg->task.mutable_shared_para();
  }
  static CallSharedParaView get(const MessagePtr& msg) {
    CHECK_EQ((int)msg->task.type(), (int)TaskView::CALL_CUSTOMER);
    CHECK(msg->task.has_shared_para());
    return msg->task.shared_para();
  }

  // the request / reply handler run by the executor thread (ref shared_parameter.h:97-162)
  void process(const MessagePtr& msg) override {
    const bool req = msg->task.request();
    int chl = msg->task.key_channel();
    CallSharedParaView call = get(msg);
    const bool push = call.cmd() == CallSharedParaView::PUSH;
    const bool pull = call.cmd() == CallSharedParaView::PULL;
    MessagePtr reply;
    if (pull && req) {
      reply.reset(new Message(*msg));
      reply->task.set_request(false);
      std::swap(reply->sender, reply->recver);
      reply->task.clear_filter();  // the reply is encoded afresh for the way back
      for (int i = 0; i < msg->task.filter_size(); ++i) {
        FilterConfigView f = msg->task.filter(i);
        if (f.type() == FilterConfigView::COMPRESSING) continue;
        FilterConfigView g = reply->addFilter(f.type());
        if (f.has_clear_cache_if_done()) g.set_clear_cache_if_done(f.clear_cache_if_done());
        for (int j = 0; j < f.fixed_point_size(); ++j) g.add_fixed_point().set_num_bytes(f.fixed_point(j).num_bytes());
      }
    }
    sys_.heartbeatInfo().startBusy();
    if (call.replica()) {
      if (pull && !req && msg->task.key_range().template to<K>() == myKeyRange()) recoverFrom(msg);
      else if ((push && req) || (pull && !req)) setReplica(msg);
      else if (pull && req) getReplica(reply);
    } else if (call.has_tail_filter()) {
      if (key_filter_ignore_chl_) chl = 0;
      TailKeyFilterView cmd = call.tail_filter();
      if (cmd.insert_count() && req && !msg->key.empty()) {
        CHECK(!msg->value.empty());
        auto& filter = key_filter_[chl];
        if (filter.empty()) {
          double w = std::max(1, sys_.yp().num_workers());
          filter.resize((int)(w * cmd.countmin_n() / log(w + 1)), cmd.countmin_k());
        }
        filter.insertKeys(SArray<K>(msg->key), SArray<uint8>(msg->value[0]));
      }
      if (cmd.has_query_key() && pull) {
        if (req) {
          reply->clearData();
          reply->setKey(key_filter_[chl].queryKeys(SArray<K>(msg->key), cmd.query_key()));
          if (cmd.has_query_value()) getValue(reply);
        } else {
          setValue(msg);
        }
      }
    } else {
      if ((push && req) || (pull && !req)) {
        setValue(msg);
        if (push && req && FLAGS_num_replicas > 0 && isServer()) backup(msg);
      } else if (pull && req) {
        getValue(reply);
      }
    }
    sys_.heartbeatInfo().stopBusy();
    if (pull && req) {
      port(reply->recver)->encodeFilter(reply);
      sys_.queue(reply);
      msg->replied = true;
    }
  }

  // ask the replica holders for my range and load it (used by a replacement server)
  void recover(Range<K> range) {
    MessagePtr msg(new Message(kReplicaGroup));
    msg->task.mutable_key_range().from(range);
    CallSharedParaView c = set(msg);
    c.set_cmd(CallSharedParaView::PULL);
    c.set_replica(true);
    int t = port(kReplicaGroup)->submit(msg);
    port(kReplicaGroup)->waitOutgoingTask(t);
  }

 protected:
  virtual void getValue(const MessagePtr& msg) = 0;
  virtual void setValue(const MessagePtr& msg) = 0;
  virtual void setReplica(const MessagePtr& msg) {}
  virtual void getReplica(const MessagePtr& msg) {}
  virtual void recoverFrom(const MessagePtr& msg) {}

  // forward an applied push to the replica group
  void backup(const MessagePtr& msg) {
    if (port(kReplicaGroup)->size() == 0) return;
    MessagePtr b(new Message(kReplicaGroup));
    b->task.set_key_channel(msg->task.key_channel());
    b->task.mutable_key_range().from(myKeyRange());
    b->task.set_has_key(msg->task.has_key());
    b->task.set_key_type(msg->task.key_type());
    for (int i = 0; i < msg->task.value_type_size(); ++i) b->task.add_value_type(msg->task.value_type(i));
    b->key = msg->key;
    b->value = msg->value;
    CallSharedParaView c = set(b);
    c.set_cmd(CallSharedParaView::PUSH);
    c.set_op(get(msg).op());
    c.set_replica(true);
    port(kReplicaGroup)->submit(b);
  }

  Range<K> myKeyRange() { return keyRange(Customer::myNodeID()); }
  Range<K> keyRange(const NodeID& id) { return Range<K>(exec_->rnode(id)->keyRange()); }

  std::unordered_map<int, FreqencyFilter<K, uint8>> key_filter_;
  bool key_filter_ignore_chl_ = true;
};

#define USING_SHARED_PARAMETER           \
  using Customer::port;                  \
  using Customer::myNodeID;              \
  using SharedParameter<K>::get;         \
  using SharedParameter<K>::set;         \
  using SharedParameter<K>::myKeyRange;  \
  using SharedParameter<K>::keyRange;    \
  using SharedParameter<K>::sync

// =================================================================================================
template <typename K, typename V>
class KVVector : public SharedParameter<K> {
 public:
  KVVector(const string& my_name, const string& parent_name = FLAGS_app_name, int k = 1)
      : SharedParameter<K>(my_name, parent_name), val_entry_size_(k) {}
  virtual ~KVVector() { this->stopExecutor(); }

  SArray<K>& key(int channel = 0) { Lock l(mu_); return key_[channel]; }
  SArray<V>& value(int channel = 0) { Lock l(mu_); return val_[channel]; }
  void clear(int channel = 0) { Lock l(mu_); key_.erase(channel); val_.erase(channel); }
  // local positions of a global key range
  SizeR find(int channel, const Range<K>& key_range) { return key(channel).findRange(key_range); }
  int valueEntrySize() const { return val_entry_size_; }

  MessagePtrList slice(const MessagePtr& msg, const KeyRangeList& sep) override {
    if (get(msg).replica()) return Customer::slice(msg, sep);
    return sliceKeyOrderedMsg<K>(msg, sep);
  }

  void getValue(const MessagePtr& msg) override {
    SArray<K> recv_key(msg->key);
    if (recv_key.empty()) return;
    int chl = msg->task.key_channel();
    const int k = val_entry_size_;
    CHECK_EQ(key(chl).size() * k, value(chl).size());
    SArray<V> val;
    parallelOrderedMatch(key(chl), value(chl), recv_key, &val, k);
    msg->clearValue();
    msg->addValue(val);
  }

  void setValue(const MessagePtr& msg) override {
    SArray<K> recv_key(msg->key);
    if (recv_key.empty()) return;
    int chl = msg->task.key_channel();
    SArray<K>& my_key = key(chl);
    SArray<V>& my_val = value(chl);
    const int k = val_entry_size_;
    if (get(msg).has_tail_filter() || get(msg).gather()) {
      // join the received keys (and values) with mine
      if (msg->value.empty() || msg->value[0].empty()) {
        CHECK(my_val.empty());
        my_key = my_key.setUnion(recv_key);
      } else {
        SArray<V> recv_val(msg->value[0]);
        CHECK_EQ(recv_key.size() * k, recv_val.size());
        SArray<K> new_key;
        SArray<V> new_val;
        parallelUnion(my_key, my_val, recv_key, recv_val, &new_key, &new_val, k);
        my_key = new_key;
        my_val = new_val;
      }
    } else {
      CHECK_EQ(msg->value.size(), (size_t)1);
      SArray<V> recv_val(msg->value[0]);
      CHECK_EQ(recv_key.size() * k, recv_val.size());
      size_t n = parallelOrderedMatch<K, V, OpPlus<V>>(recv_key, recv_val, my_key, &my_val, k);
      CHECK_EQ(n, recv_key.size()) << "pushed keys must exist in the receiver";
    }
  }

  // per-owner copies on replica holders
  void setReplica(const MessagePtr& msg) override {
    Lock l(rep_mu_);
    // keyed by the owner's key range, not its node id: a replacement node asks for the range
    auto& rep = replica_[Range<K>(msg->task.key_range().template to<K>()).begin()][msg->task.key_channel()];
    SArray<K> k(msg->key);
    if (msg->value.empty()) return;
    SArray<V> v(msg->value[0]);
    for (size_t i = 0; i < k.size(); ++i) rep[k[i]] += v[i];
  }
  void getReplica(const MessagePtr& msg) override {
    Lock l(rep_mu_);
    // the request names the (lost) range
    auto it = replica_.find(Range<K>(msg->task.key_range().template to<K>()).begin());
    SArray<K> k;
    SArray<V> v;
    if (it != replica_.end())
      for (auto& kv : it->second[msg->task.key_channel()]) { k.pushBack(kv.first); v.pushBack(kv.second); }
    msg->clearData();
    msg->setKey(k);
    msg->addValue(v);
  }
  void recoverFrom(const MessagePtr& msg) override {
    int chl = msg->task.key_channel();
    key(chl) = SArray<K>(msg->key);
    if (!msg->value.empty()) value(chl) = SArray<V>(msg->value[0]);
  }

  USING_SHARED_PARAMETER;

 protected:
  std::mutex mu_, rep_mu_;
  std::unordered_map<int, SArray<K>> key_;
  std::unordered_map<int, SArray<V>> val_;
  std::map<K, std::map<int, std::map<K, V>>> replica_;
  int val_entry_size_;
};

// =================================================================================================
template <typename K, typename V>
class KVBufferedVector : public KVVector<K, V> {
 public:
  KVBufferedVector(const string& my_name, const string& parent_name)
      : KVVector<K, V>(my_name, parent_name) {}
  virtual ~KVBufferedVector() { this->stopExecutor(); }

  // merged data received at time t (first = positions in my key array), consumed by the call
  typedef std::pair<SizeR, std::vector<SArray<V>>> MergedData;
  MergedData received(int t) {
    Lock l(this->mu_);
    auto it = recved_val_.find(t);
    CHECK(it != recved_val_.end()) << this->myNodeID() << " hasn't received data at time " << t;
    MergedData ret = it->second;
    recved_val_.erase(it);
    return ret;
  }
  bool hasReceived(int t) {
    Lock l(this->mu_);
    return recved_val_.count(t) > 0;
  }

  void setValue(const MessagePtr& msg) override {
    SArray<K> recv_key(msg->key);
    if (recv_key.empty()) return;
    int chl = msg->task.key_channel();
    SArray<K>& my_key = this->key(chl);
    if (msg->value.empty()) {  // keys only: merge them and drop the values
      my_key = my_key.setUnion(recv_key);
      this->value(chl).clear();
      return;
    }
    int t = msg->task.time();
    Range<K> key_range = msg->task.key_range().template to<K>();
    SizeR idx_range = my_key.findRange(key_range);
    std::unique_lock<std::mutex> l(merge_mu_);  // several workers' pushes for one t serialise here
    this->mu_.lock();
    MergedData& matched = recved_val_[t];
    this->mu_.unlock();
    for (size_t i = 0; i < msg->value.size(); ++i) {
      SArray<V> recv_data(msg->value[i]);
      CHECK_EQ(recv_data.size(), recv_key.size());
      if (matched.second.size() <= i) {  // first sender allocates
        matched.first = idx_range;
        matched.second.push_back(SArray<V>());
        size_t n = parallelOrderedMatch(recv_key, recv_data, my_key.segment(idx_range), &matched.second[i]);
        CHECK_EQ(n, recv_key.size());
      } else {
        CHECK_EQ(matched.first, idx_range);
        size_t n = parallelOrderedMatch<K, V, OpPlus<V>>(recv_key, recv_data, my_key.segment(idx_range),
                                                         &matched.second[i]);
        CHECK_EQ(n, recv_key.size());
      }
    }
  }

 protected:
  std::mutex merge_mu_;
  std::unordered_map<int, MergedData> recved_val_;
};

// =================================================================================================
template <typename K, typename V>
class KVMap : public SharedParameter<K> {
 public:
  KVMap(const string& my_name, const string& parent_name = FLAGS_app_name)
      : SharedParameter<K>(my_name, parent_name) {}
  virtual ~KVMap() { this->stopExecutor(); }

  MessagePtrList slice(const MessagePtr& msg, const KeyRangeList& sep) override {
    if (get(msg).replica()) return Customer::slice(msg, sep);
    return sliceKeyOrderedMsg<K>(msg, sep);
  }
  void getValue(const MessagePtr& msg) override {
    SArray<K> key(msg->key);
    SArray<V> val(key.size());
    Lock l(mu_);
    for (size_t i = 0; i < key.size(); ++i) {
      auto it = data_.find(key[i]);
      val[i] = it == data_.end() ? V() : it->second;
    }
    msg->clearValue();
    msg->addValue(val);
  }
  void setValue(const MessagePtr& msg) override {
    SArray<K> key(msg->key);
    if (key.empty()) return;
    CHECK_GT(msg->value.size(), (size_t)0);
    SArray<V> value(msg->value[0]);
    CHECK_EQ(key.size(), value.size());
    Operator op = msg->task.shared_para().op();
    Lock l(mu_);
    if (msg->task.request()) {
      for (size_t i = 0; i < key.size(); ++i) compAssOp(data_[key[i]], value[i], op);
    } else {  // a pull reply overwrites the local cache
      for (size_t i = 0; i < key.size(); ++i) data_[key[i]] = value[i];
    }
  }
  void setReplica(const MessagePtr& msg) override {
    SArray<K> key(msg->key);
    if (msg->value.empty()) return;
    SArray<V> value(msg->value[0]);
    Operator op = msg->task.shared_para().op();
    Lock l(mu_);
    // keyed by the owner's key range (a replacement node asks for the range, not for a node id)
    auto& rep = replica_[Range<K>(msg->task.key_range().template to<K>()).begin()];
    for (size_t i = 0; i < key.size(); ++i) compAssOp(rep[key[i]], value[i], op);
  }
  void getReplica(const MessagePtr& msg) override {
    Lock l(mu_);
    std::map<K, V> sorted;
    auto it = replica_.find(Range<K>(msg->task.key_range().template to<K>()).begin());
    if (it != replica_.end()) sorted.insert(it->second.begin(), it->second.end());
    SArray<K> k;
    SArray<V> v;
    for (auto& kv : sorted) { k.pushBack(kv.first); v.pushBack(kv.second); }
    msg->clearData();
    msg->setKey(k);
    msg->addValue(v);
  }
  void recoverFrom(const MessagePtr& msg) override {
    SArray<K> key(msg->key);
    if (msg->value.empty()) return;
    SArray<V> value(msg->value[0]);
    Lock l(mu_);
    for (size_t i = 0; i < key.size(); ++i) data_[key[i]] = value[i];
  }
  // local access
  V& operator[](const K& k) { Lock l(mu_); return data_[k]; }
  size_t size() { Lock l(mu_); return data_.size(); }
  std::unordered_map<K, V>& data() { return data_; }

  USING_SHARED_PARAMETER;

 protected:
  std::mutex mu_;
  std::unordered_map<K, V> data_;
  std::map<K, std::unordered_map<K, V>> replica_;
};

// =================================================================================================
template <typename K, typename S>
class KVState : public SharedParameter<K> {
 public:
  KVState(const string& my_name, const string& parent_name) : SharedParameter<K>(my_name, parent_name) {}
  virtual ~KVState() {}
  void setState(const S& s) { state_ = s; }
  S& state() { return state_; }
  virtual void writeToFile(const string& file) = 0;

 protected:
  S state_;
};

// Entry::get(const V* grad, State*) applies an update, Entry::put(V* out, State*) emits the weight
template <typename K, typename V, typename E, typename S>
class KVStore : public KVState<K, S> {
 public:
  KVStore(const string& my_name, const string& parent_name) : KVState<K, S>(my_name, parent_name) {}
  virtual ~KVStore() { this->stopExecutor(); }
  using KVState<K, S>::state_;

  void getValue(const MessagePtr& msg) override {
    SArray<K> key(msg->key);
    SArray<V> val(key.size());
    for (size_t i = 0; i < key.size(); ++i) data_[key[i]].put(val.data() + i, &state_);
    msg->clearValue();
    msg->addValue(val);
  }
  void setValue(const MessagePtr& msg) override {
    SArray<K> key(msg->key);
    if (key.empty()) return;
    CHECK_GT(msg->value.size(), (size_t)0);
    SArray<V> val(msg->value[0]);
    CHECK_EQ(key.size(), val.size());
    for (size_t i = 0; i < key.size(); ++i) data_[key[i]].get(val.data() + i, &state_);
    state_.update();
  }
  MessagePtrList slice(const MessagePtr& msg, const KeyRangeList& krs) override {
    return sliceKeyOrderedMsg<K>(msg, krs);
  }
  // "key\tvalue" per non-zero weight; keys ascending (the reference iterates the hash map)
  void writeToFile(const string& file) override {
    if (!dirExists(getPath(file))) createDir(getPath(file));
    std::ofstream out(file);
    CHECK(out.good()) << "cannot write " << file;
    std::vector<K> keys;
    keys.reserve(data_.size());
    for (auto& e : data_) keys.push_back(e.first);
    std::sort(keys.begin(), keys.end());
    V v;
    for (K k : keys) {
      data_[k].put(&v, &state_);
      if (v != 0 && v == v) out << k << "\t" << v << "\n";
    }
  }
  size_t size() const { return data_.size(); }

 protected:
  std::unordered_map<K, E> data_;
};

}  // namespace PS
