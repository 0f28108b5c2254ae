// Shared parameters: the public key-value API of the parameter server.
//
//   SharedParameter<K>        push / pull / sync, the timestamped request handler
//   KVVector<K,V>             sorted keys + aligned values per channel
//   KVBufferedVector<K,V>     pushes parked per timestamp until the app collects them
//   KVMap<K,V>                key -> value with per-key operators (PLUS ... XOR)
//   KVStore<K,V,Entry,State>  key -> optimizer entry (host only: entries are C++ functors)
//   FreqencyFilter<K,V>       CountMin tail-feature filter
//
// The API is the reference's (ref src/parameter/{shared_parameter,kv_vector,kv_buffered_vector,
// kv_map,kv_store,frequency_filter}.h); the construction is not.  Every container keeps its
// values in a *column store* with two implementations chosen at run time by -device:
//
//   cpu   values in SArrays, sorted-merge matching on host threads;
//   gpu   values in HBM (DevArray), key matching / operators / gathers are kernels of
//         libpsb200_cuda.so (devkv.cu), received value frames are views of the SENDER's staging
//         arena delivered by the PeerVan (system/peer_van.h) and read over NVLink by the kernel
//         that applies them; pull replies are gathered straight into this node's arena and
//         leave without a copy.  Timestamps, wait_time dependencies, waitInMsg / waitOutMsg /
//         finish and received(t) are the executor's and are identical in both modes.
//
// Host accessors key() / value() keep working in gpu mode: they hand out a host mirror
// (downloaded on demand, uploaded again before the next device operation), so a program written
// against the reference's ps.h - e.g. bin/hello.cc - runs unchanged with its shard in HBM.
// Code that wants to stay on the device uses deviceValue() / deviceKey().
//
// Replication (setReplica / getReplica / recoverFrom are no-ops in the reference,
// shared_parameter.h:68-79): with -num_replicas r every applied push is forwarded to the r nodes
// of kReplicaGroup, which keep a host copy per owner key range that a replacement node pulls
// back with recover().
#pragma once
#include <fstream>
#include <type_traits>

#include "device/cuda_api.h"
#include "system/heartbeat.h"
#include "system/peer_van.h"
#include "system/runtime.h"
#include "util/hash.h"
#include "util/localizer.h"
#include "util/sketch.h"

namespace PS {

DECLARE_int32(num_workers);
DECLARE_int32(num_replicas);
DECLARE_string(app_name);
DECLARE_string(device);

// ---- operators (ref shared_parameter.h:173-193) --------------------------------------------------
template <typename V>
inline void compAssOp(V& right, V left, Operator op) {
  if constexpr (std::is_integral<V>::value) {
    if (op == AND) { right &= left; return; }
    if (op == OR) { right |= left; return; }
    if (op == XOR) { right ^= left; return; }
  }
  if (op == PLUS) right += left;
  else if (op == MINUS) right -= left;
  else if (op == TIMES) right *= left;
  else if (op == DIVIDE) right /= left;
}

// value type tags of the device kernels (include/psb200.h PSB_DT_*), -1: no device support
template <typename V> struct DevType { static const int id = -1; };
template <> struct DevType<float> { static const int id = 0; };
template <> struct DevType<double> { static const int id = 1; };
template <> struct DevType<int32> { static const int id = 2; };
template <> struct DevType<int64> { static const int id = 3; };
template <> struct DevType<uint8> { static const int id = 4; };

// ---- tail-feature filter -------------------------------------------------------------------------
template <typename K, typename V>
class FreqencyFilter {
 public:
  void resize(int n, int k) { sketch_.resize(n, k, (V)254); }
  void clear() { sketch_.clear(); }
  bool empty() { return sketch_.empty(); }
  void insertKeys(const SArray<K>& key, const SArray<V>& count) {
    CHECK_EQ(key.size(), count.size());
    const K* k = key.begin();
    for (const V* c = count.begin(); c != count.end(); ++c, ++k) sketch_.insert(*k, *c);
  }
  // the keys seen more than `freqency` times (estimate; never undercounts)
  SArray<K> queryKeys(const SArray<K>& key, int freqency) {
    CHECK_LT(freqency, 255) << "the sketch counters are 8 bit";
    SArray<K> out;
    for (const K* k = key.begin(); k != key.end(); ++k)
      if ((int)sketch_.query(*k) > freqency) out.pushBack(*k);
    return out;
  }

 private:
  CountMin<K, V> sketch_;
};

// =================================================================================================
// SharedParameter: the RPC surface shared by all containers
// =================================================================================================
template <typename K>
class SharedParameter : public Customer {
 public:
  SharedParameter(const string& my_name, const string& parent_name) : Customer(my_name, parent_name) {}
  virtual ~SharedParameter() {}

  // ---- client side: returns the timestamp of the request ----
  int push(MessagePtr msg) { return call(msg, CallSharedParaView::PUSH); }
  int pull(MessagePtr msg) { return call(msg, CallSharedParaView::PULL); }
  int sync(MessagePtr msg) {
    CHECK(msg->task.shared_para().has_cmd()) << msg->debugString();
    if (!msg->task.has_key_range()) msg->task.mutable_key_range().from(Range<K>::all());
    return port(msg->recver)->submit(msg);
  }
  void waitInMsg(const NodeID& node, int time) { port(node)->waitIncomingTask(time); }
  void waitOutMsg(const NodeID& node, int time) { port(node)->waitOutgoingTask(time); }
  void finish(const NodeID& node, int time) { port(node)->finishIncomingTask(time); }

  void setTailFilterIgnoreChl(bool flag) { one_filter_for_all_ = flag; }
  void setTailFilterSize(int chl, int n, int k) { tail_filter_[chl].resize(n, k); }
  // Called by a server when the keys round of group `chl` is through.  With one sketch per channel that
  // sketch is done.  With the shared sketch (the default, as in the reference) the count / query rounds
  // of LATER groups may still be in flight at that moment, so only the last group's call clears it:
  // by then every worker has pushed the keys of every group, i.e. all its queries have returned.
  // (The reference clears key_filter_[chl] while its shared sketch lives at key 0 - shared_parameter.h:49,122 -
  // so there the shared sketch is never cleared at all.)
  void clearTailFilter(int chl, bool last_group) {
    if (!one_filter_for_all_) tail_filter_[chl].clear();
    else if (last_group) tail_filter_[0].clear();
  }

  static CallSharedParaView set(MessagePtr msg) {
    msg->task.set_type(TaskView::CALL_CUSTOMER);
    return msg->task.mutable_shared_para();
  }
  static CallSharedParaView get(const MessagePtr& msg) {
    CHECK_EQ((int)msg->task.type(), (int)TaskView::CALL_CUSTOMER);
    CHECK(msg->task.has_shared_para());
    return msg->task.shared_para();
  }

  // ---- server / reply side, on the executor thread (ref shared_parameter.h:97-162) ----
  // A message is one of five things; everything else follows from that.
  enum class Kind { STORE, ANSWER, REPLICA_STORE, REPLICA_ANSWER, RECOVER };
  void process(const MessagePtr& msg) override {
    const CallSharedParaView call = get(msg);
    const bool request = msg->task.request();
    const bool is_pull = call.cmd() == CallSharedParaView::PULL;
    CHECK(is_pull || call.cmd() == CallSharedParaView::PUSH);
    // data flows towards me on a push request and on a pull reply; a pull request asks for data
    const bool incoming = is_pull != request;
    Kind kind = incoming ? Kind::STORE : Kind::ANSWER;
    if (call.replica()) {
      if (!incoming) kind = Kind::REPLICA_ANSWER;
      else if (is_pull && msg->task.key_range().template to<K>() == myKeyRange()) kind = Kind::RECOVER;
      else kind = Kind::REPLICA_STORE;
    }
    MessagePtr reply = (is_pull && request) ? makeReply(msg) : MessagePtr();

    sys_.heartbeatInfo().startBusy();
    if (!call.replica() && call.has_tail_filter()) {
      tailFilterStep(msg, reply, call, request, is_pull);
    } else {
      switch (kind) {
        case Kind::STORE:
          setValue(msg);
          if (request && FLAGS_num_replicas > 0 && isServer()) backup(msg);
          break;
        case Kind::ANSWER: getValue(reply); break;
        case Kind::REPLICA_STORE: setReplica(msg); break;
        case Kind::REPLICA_ANSWER: getReplica(reply); break;
        case Kind::RECOVER: recoverFrom(msg); break;
      }
    }
    sys_.heartbeatInfo().stopBusy();

    if (reply) {
      port(reply->recver)->encodeFilter(reply);
      sys_.queue(reply);
      msg->replied = true;
    }
  }

  // a replacement server asks the replica holders for its range and loads the answer
  void recover(Range<K> range) {
    MessagePtr msg(new Message(kReplicaGroup));
    msg->task.mutable_key_range().from(range);
    CallSharedParaView c = set(msg);
    c.set_cmd(CallSharedParaView::PULL);
    c.set_replica(true);
    const int t = port(kReplicaGroup)->submit(msg);
    port(kReplicaGroup)->waitOutgoingTask(t);
  }

 protected:
  virtual void getValue(const MessagePtr& msg) = 0;
  virtual void setValue(const MessagePtr& msg) = 0;
  virtual void setReplica(const MessagePtr& msg) {}
  virtual void getReplica(const MessagePtr& msg) {}
  virtual void recoverFrom(const MessagePtr& msg) {}

  Range<K> myKeyRange() { return keyRange(Customer::myNodeID()); }
  Range<K> keyRange(const NodeID& id) { return Range<K>(exec_->rnode(id)->keyRange()); }

  // the reply to a pull request: same header turned around; every filter of the request except
  // compression is requested for the way back too (compression is added per hop)
  static MessagePtr makeReply(const MessagePtr& msg) {
    MessagePtr r(new Message(*msg));
    r->task.set_request(false);
    r->sender.swap(r->recver);
    r->task.clear_filter();
    for (int i = 0; i < msg->task.filter_size(); ++i) {
      const FilterConfigView want = msg->task.filter(i);
      if (want.type() == FilterConfigView::COMPRESSING) continue;
      FilterConfigView f = r->addFilter(want.type());
      if (want.has_clear_cache_if_done()) f.set_clear_cache_if_done(want.clear_cache_if_done());
      for (int j = 0; j < want.fixed_point_size(); ++j)
        f.add_fixed_point().set_num_bytes(want.fixed_point(j).num_bytes());
    }
    return r;
  }

  // counts go into the sketch on their way in; a query pull is answered with the frequent keys
  // (and, if asked, their values); the reply of such a pull is stored like any other
  void tailFilterStep(const MessagePtr& msg, const MessagePtr& reply, const CallSharedParaView& call,
                      bool request, bool is_pull) {
    const TailKeyFilterView tf = call.tail_filter();
    auto& filter = tail_filter_[one_filter_for_all_ ? 0 : msg->task.key_channel()];
    if (request && tf.insert_count() && !msg->key.empty()) {
      CHECK(!msg->value.empty()) << "insert_count needs the counts as first value";
      if (filter.empty()) {
        const double w = std::max(1, sys_.yp().num_workers());
        filter.resize((int)(w * tf.countmin_n() / log(w + 1)), tf.countmin_k());
      }
      filter.insertKeys(SArray<K>(msg->key), SArray<uint8>(msg->value[0]));
    }
    if (!is_pull || !tf.has_query_key()) return;
    if (!request) { setValue(msg); return; }
    reply->clearData();
    reply->setKey(filter.queryKeys(SArray<K>(msg->key), tf.query_key()));
    if (tf.has_query_value()) getValue(reply);
  }

  // forward an applied push to the nodes that replicate my range
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

 private:
  int call(MessagePtr msg, CallSharedParaView::Command cmd) {
    set(msg).set_cmd(cmd);
    return sync(msg);
  }
  std::unordered_map<int, FreqencyFilter<K, uint8>> tail_filter_;
  bool one_filter_for_all_ = true;
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
// Device plumbing shared by the containers
// =================================================================================================
// What a container needs from the GPU side: the API table, the node's PeerVan (for staging replies
// without a copy) and a stream.  Null when the node runs with -device cpu.
struct DeviceContext {
  CudaApi* cuda = nullptr;
  PeerVan* van = nullptr;
  void* stream = nullptr;
  explicit operator bool() const { return cuda != nullptr; }

  static DeviceContext of(Postoffice& sys) {
    DeviceContext d;
    if (FLAGS_device != "gpu") return d;
    d.van = dynamic_cast<PeerVan*>(&sys.yp().van());
    CHECK(d.van) << "-device gpu needs the PeerVan (started by Postoffice::start)";
    d.cuda = CudaApi::get();
    d.stream = d.van->stream();
    return d;
  }
  void bind() const { cuda->check(cuda->set_device(van->gpu()), "cudaSetDevice"); }
  void sync() const { cuda->check(cuda->stream_sync(stream), "stream sync"); }
  // a value frame as device memory: frames that came over the host van are uploaded into `tmp`
  template <typename V>
  const V* onDevice(const SArray<char>& frame, DevArray<V>* tmp) const {
    if (frame.onDevice()) return reinterpret_cast<const V*>(frame.data());
    tmp->fromHost(reinterpret_cast<const V*>(frame.data()), frame.size() / sizeof(V), stream);
    return tmp->data();
  }
  // download (views of peer memory included)
  template <typename V>
  SArray<V> toHost(const SArray<char>& frame) const {
    if (!frame.onDevice()) return SArray<V>(frame);
    SArray<V> h(frame.size() / sizeof(V));
    if (h.size()) {
      cuda->check(cuda->memcpy_async(h.data(), frame.data(), (int64)frame.size(), stream), "D2H");
      sync();
    }
    return h;
  }
};

// =================================================================================================
// KVVector
// =================================================================================================
template <typename K, typename V>
class KVVector : public SharedParameter<K> {
 public:
  KVVector(const string& my_name, const string& parent_name = FLAGS_app_name, int k = 1)
      : SharedParameter<K>(my_name, parent_name), val_entry_size_(k) {
    if (sizeof(K) == 8 && DevType<V>::id >= 0) dev_ = DeviceContext::of(this->sys_);
  }
  virtual ~KVVector() { this->stopExecutor(); }

  // host views.  In gpu mode value() first brings the mirror up to date; both accessors mark the
  // channel as possibly edited, so it is uploaded again before the next device operation.
  SArray<K>& key(int channel = 0) {
    Lock l(mu_);
    Column& c = col_[channel];
    c.host_edited = true;
    return c.key;
  }
  SArray<V>& value(int channel = 0) {
    Lock l(mu_);
    Column& c = col_[channel];
    if (dev_ && c.device_newer) download(c);
    c.host_edited = true;
    return c.val;
  }
  void clear(int channel = 0) { Lock l(mu_); col_.erase(channel); }
  SizeR find(int channel, const Range<K>& key_range) { return key(channel).findRange(key_range); }
  int valueEntrySize() const { return val_entry_size_; }
  bool onDevice() const { return (bool)dev_; }
  // gpu mode: the shard as it lives in HBM (valid until the keys of the channel change)
  const V* deviceValue(int channel = 0) { Lock l(mu_); Column& c = col_[channel]; upload(c); return c.dval.data(); }
  const uint64* deviceKey(int channel = 0) { Lock l(mu_); Column& c = col_[channel]; upload(c); return c.dkey.data(); }

  // gpu mode: a copy of `host` in this node's staging arena, as a device view that can be put
  // into a message (addValue): it then reaches the receiver through peer memory, never through
  // the host van.  cpu mode: returns `host` itself.
  SArray<V> stage(const SArray<V>& host) {
    if (!dev_ || host.empty()) return host;
    dev_.bind();
    SArray<char> st = dev_.van->allocStaging(host.memSize());
    dev_.cuda->check(dev_.cuda->memcpy_async(st.data(), host.data(), (int64)host.memSize(), dev_.stream), "H2D");
    dev_.sync();
    return SArray<V>(st);
  }

  MessagePtrList slice(const MessagePtr& msg, const KeyRangeList& sep) override {
    if (get(msg).replica()) return Customer::slice(msg, sep);
    return sliceKeyOrderedMsg<K>(msg, sep);
  }

  // answer a pull: my values at the requested keys (zeros for keys I do not hold)
  void getValue(const MessagePtr& msg) override {
    SArray<K> want(msg->key);
    if (want.empty()) return;
    const int chl = msg->task.key_channel();
    const int k = val_entry_size_;
    Lock l(mu_);
    Column& c = col_[chl];
    msg->clearValue();
    if (!dev_) {
      CHECK_EQ(c.key.size() * k, c.val.size());
      SArray<V> out;
      parallelOrderedMatch(c.key, c.val, want, &out, k);
      msg->addValue(out);
      return;
    }
    upload(c);
    const int64* pos = positions(c, want);
    // gathered straight into the staging arena: the PeerVan sends it without another copy
    SArray<char> out = dev_.van->allocStaging(want.size() * k * sizeof(V));
    dev_.cuda->check(dev_.cuda->dkv_gather(out.data(), c.dval.data(), pos, (int64)want.size(),
                                           (int32)(k * sizeof(V)), dev_.stream), "gather");
    dev_.sync();
    msg->addValue(SArray<V>(out));
  }

  // store what arrived (a push, or the reply to my pull): added into my values at the matching
  // keys; with tail_filter / gather set the received keys are joined with mine instead
  void setValue(const MessagePtr& msg) override {
    SArray<K> got(msg->key);
    if (got.empty()) return;
    const int chl = msg->task.key_channel();
    const int k = val_entry_size_;
    const bool join = get(msg).has_tail_filter() || get(msg).gather();
    Lock l(mu_);
    Column& c = col_[chl];
    if (join) {
      if (dev_ && c.device_newer) download(c);
      SArray<V> vals;
      if (!msg->value.empty()) vals = dev_ ? dev_.template toHost<V>(msg->value[0]) : SArray<V>(msg->value[0]);
      joinKeys(c, got, vals, k);
      return;
    }
    CHECK_EQ(msg->value.size(), (size_t)1);
    CHECK_EQ(got.size() * k * sizeof(V), msg->value[0].size());
    if (!dev_) {
      const size_t n = parallelOrderedMatch<K, V, OpPlus<V>>(got, SArray<V>(msg->value[0]), c.key, &c.val, k);
      CHECK_EQ(n, got.size()) << "pushed keys must exist in the receiver";
      return;
    }
    upload(c);
    const int64* pos = positions(c, got);
    DevArray<V> tmp;
    const V* src = dev_.template onDevice<V>(msg->value[0], &tmp);
    const int op = get(msg).has_op() ? (int)get(msg).op() + 1 : 1 /* PLUS */;
    dev_.cuda->check(dev_.cuda->dkv_apply(c.dval.data(), pos, src, (int64)got.size(), k, DevType<V>::id, op,
                                          dev_.stream), "apply");
    dev_.sync();  // the frame may be released (and reused by its sender) once we return
    c.device_newer = true;
  }

  // ---- replicas: host copies per owner range ----
  void setReplica(const MessagePtr& msg) override {
    if (msg->value.empty()) return;
    SArray<K> k(msg->key);
    SArray<V> v = dev_ ? dev_.template toHost<V>(msg->value[0]) : SArray<V>(msg->value[0]);
    Lock l(rep_mu_);
    auto& rep = replica_[ownerOf(msg)][msg->task.key_channel()];
    for (size_t i = 0; i < k.size(); ++i) rep[k[i]] += v[i];
  }
  void getReplica(const MessagePtr& msg) override {
    SArray<K> k;
    SArray<V> v;
    {
      Lock l(rep_mu_);
      auto it = replica_.find(ownerOf(msg));
      if (it != replica_.end())
        for (const auto& kv : it->second[msg->task.key_channel()]) { k.pushBack(kv.first); v.pushBack(kv.second); }
    }
    msg->clearData();
    msg->setKey(k);
    msg->addValue(v);
  }
  void recoverFrom(const MessagePtr& msg) override {
    Lock l(mu_);
    Column& c = col_[msg->task.key_channel()];
    c.key = SArray<K>(msg->key);
    if (!msg->value.empty()) c.val = dev_ ? dev_.template toHost<V>(msg->value[0]) : SArray<V>(msg->value[0]);
    c.host_edited = true;
    c.device_newer = false;
  }

  USING_SHARED_PARAMETER;

 protected:
  // one channel: host arrays (the truth in cpu mode, a mirror in gpu mode) and the HBM copy
  struct Column {
    SArray<K> key;
    SArray<V> val;
    DevArray<uint64> dkey;
    DevArray<V> dval;
    bool host_edited = true;    // key()/value() were handed out since the last upload
    bool device_newer = false;  // kernels changed dval since the last download
    // positions of recently seen key lists in dkey, by signature (the device successor of the
    // KEY_CACHING filter: a repeated key list costs no search at all)
    struct Match { size_t n; DevArray<int64> pos; };
    std::unordered_map<uint64, Match> matches;
  };

  // a replica is filed under the first key of its owner's range: a replacement node asks for a
  // range, not for a node id
  K ownerOf(const MessagePtr& msg) { return Range<K>(msg->task.key_range().template to<K>()).begin(); }

  void joinKeys(Column& c, const SArray<K>& got, const SArray<V>& vals, int k) {
    if (vals.empty()) {
      CHECK(c.val.empty()) << "keys without values joined into a channel that holds values";
      c.key = c.key.setUnion(got);
    } else {
      CHECK_EQ(got.size() * k, vals.size());
      SArray<K> nk;
      SArray<V> nv;
      parallelUnion(c.key, c.val, got, vals, &nk, &nv, k);
      c.key = nk;
      c.val = nv;
    }
    c.host_edited = true;
    c.device_newer = false;
  }

  // ---- gpu mode helpers (mu_ held) ----
  void upload(Column& c) {
    if (!c.host_edited) return;
    dev_.bind();
    if (c.device_newer) download(c);
    const size_t want = c.key.size() * val_entry_size_;
    if (c.val.size() != want) {  // same rule as the host match: a mis-sized value array starts at zero
      c.val.resize(want);
      c.val.setZero();
    }
    c.dkey.fromHost(reinterpret_cast<const uint64*>(c.key.data()), c.key.size(), dev_.stream);
    c.dval.fromHost(c.val.data(), c.val.size(), dev_.stream);
    c.matches.clear();
    c.host_edited = false;
    c.device_newer = false;
  }
  void download(Column& c) {
    dev_.bind();
    c.val.resize(c.dval.size());
    c.dval.toHost(c.val.data(), dev_.stream);
    c.device_newer = false;
  }
  const int64* positions(Column& c, const SArray<K>& keys) {
    const char* bytes = reinterpret_cast<const char*>(keys.data());
    const uint64 sig = ((uint64)crc32c::Value(bytes, keys.memSize()) << 32) ^ keys.size() ^
                       ((uint64)crc32c::Value(bytes, std::min<size_t>(keys.memSize(), 64)) << 7);
    auto it = c.matches.find(sig);
    if (it != c.matches.end() && it->second.n == keys.size()) return it->second.pos.data();
    if (c.matches.size() > 64) c.matches.clear();
    typename Column::Match& m = c.matches[sig];
    m.n = keys.size();
    DevArray<uint64> dk;
    dk.fromHost(reinterpret_cast<const uint64*>(keys.data()), keys.size(), dev_.stream);
    m.pos.resize(keys.size(), false, dev_.stream);
    dev_.cuda->check(dev_.cuda->dkv_positions(c.dkey.data(), (int64)c.dkey.size(), dk.data(), (int64)keys.size(),
                                              m.pos.data(), dev_.stream), "positions");
    dev_.sync();
    return m.pos.data();
  }

  std::mutex mu_, rep_mu_;
  std::unordered_map<int, Column> col_;
  std::map<K, std::map<int, std::map<K, V>>> replica_;
  int val_entry_size_;
  DeviceContext dev_;
};

// =================================================================================================
// KVBufferedVector: a push is parked under its timestamp, summed over senders, until received(t)
// =================================================================================================
template <typename K, typename V>
class KVBufferedVector : public KVVector<K, V> {
 public:
  KVBufferedVector(const string& my_name, const string& parent_name) : KVVector<K, V>(my_name, parent_name) {}
  virtual ~KVBufferedVector() { this->stopExecutor(); }

  // what arrived at time t, aligned with my keys: first = positions [lo, hi) in my key array,
  // second = one array per value of the message.  The call consumes the entry.  In gpu mode the
  // arrays are device views (SArray::onDevice()); receivedOnHost() downloads them.
  typedef std::pair<SizeR, std::vector<SArray<V>>> MergedData;
  MergedData received(int t) {
    Lock l(this->mu_);
    auto it = parked_.find(t);
    CHECK(it != parked_.end()) << this->myNodeID() << " hasn't received data at time " << t;
    MergedData out = std::move(it->second);
    parked_.erase(it);
    return out;
  }
  MergedData receivedOnHost(int t) {
    MergedData m = received(t);
    if (this->dev_)
      for (auto& a : m.second) a = this->dev_.template toHost<V>(SArray<char>(a));
    return m;
  }
  bool hasReceived(int t) { Lock l(this->mu_); return parked_.count(t) > 0; }

  void setValue(const MessagePtr& msg) override {
    SArray<K> got(msg->key);
    if (got.empty()) return;
    const int chl = msg->task.key_channel();
    if (msg->value.empty()) {  // bare keys extend my key set; values (if any) no longer line up
      SArray<K>& mine = this->key(chl);
      mine = mine.setUnion(got);
      this->value(chl).clear();
      return;
    }
    const int t = msg->task.time();
    SArray<K>& mine = this->key(chl);
    const SizeR where = mine.findRange(msg->task.key_range().template to<K>());
    const SArray<K> window = mine.segment(where);
    Lock merge(merge_mu_);  // the pushes of several workers for one t are summed one at a time
    MergedData* slot;
    { Lock l(this->mu_); slot = &parked_[t]; }
    const bool first = slot->second.empty();
    if (first) slot->first = where;
    else CHECK_EQ(slot->first, where) << "pushes for one timestamp must cover the same key range";
    for (size_t i = 0; i < msg->value.size(); ++i) {
      CHECK_EQ(msg->value[i].size(), got.size() * sizeof(V));
      if (slot->second.size() <= i) slot->second.push_back(SArray<V>());
      if (this->dev_) parkOnDevice(&slot->second[i], window, got, msg->value[i]);
      else {
        const size_t n = parallelOrderedMatch<K, V, OpPlus<V>>(got, SArray<V>(msg->value[i]), window, &slot->second[i]);
        CHECK_EQ(n, got.size());
      }
    }
  }

 protected:
  // the parked array lives in HBM (owned by the SArray's deleter); positions are searched on
  // the host (the window is small: one feature block) and uploaded
  void parkOnDevice(SArray<V>* acc, const SArray<K>& window, const SArray<K>& got, const SArray<char>& frame) {
    DeviceContext& d = this->dev_;
    d.bind();
    if (acc->size() != window.size()) {
      void* p = nullptr;
      d.cuda->check(d.cuda->malloc_(&p, (int64)std::max<size_t>(window.size() * sizeof(V), 8)), "parked array");
      d.cuda->check(d.cuda->memset_async(p, 0, (int64)(window.size() * sizeof(V)), d.stream), "memset");
      CudaApi* cuda = d.cuda;
      shared_ptr<char> owner((char*)p, [cuda](char* q) { cuda->free_(q); });
      *acc = SArray<V>::wrapDevice(owner, (V*)p, window.size());
    }
    SArray<int64> pos(got.size());
    const K* w = window.begin();
    const K* const w_end = window.end();
    for (size_t i = 0; i < got.size(); ++i) {
      w = std::lower_bound(w, w_end, got[i]);
      CHECK(w != w_end && *w == got[i]) << "pushed key " << got[i] << " is not one of mine";
      pos[i] = w - window.begin();
    }
    DevArray<int64> dpos;
    dpos.fromHost(pos.data(), pos.size(), d.stream);
    DevArray<V> tmp;
    const V* src = d.template onDevice<V>(frame, &tmp);
    d.cuda->check(d.cuda->dkv_apply(acc->data(), dpos.data(), src, (int64)got.size(), 1, DevType<V>::id,
                                    1 /* PLUS */, d.stream), "park");
    d.sync();
  }

  std::mutex merge_mu_;
  std::unordered_map<int, MergedData> parked_;
};

// =================================================================================================
// KVMap: unordered keys, a push applies the message's operator per key
// =================================================================================================
DECLARE_int32(kvmap_capacity);

template <typename K, typename V>
class KVMap : public SharedParameter<K> {
 public:
  KVMap(const string& my_name, const string& parent_name = FLAGS_app_name)
      : SharedParameter<K>(my_name, parent_name) {
    if (sizeof(K) == 8 && DevType<V>::id >= 0) dev_ = DeviceContext::of(this->sys_);
  }
  virtual ~KVMap() { this->stopExecutor(); }
  bool onDevice() const { return (bool)dev_; }

  MessagePtrList slice(const MessagePtr& msg, const KeyRangeList& sep) override {
    if (get(msg).replica()) return Customer::slice(msg, sep);
    return sliceKeyOrderedMsg<K>(msg, sep);
  }

  void getValue(const MessagePtr& msg) override {
    SArray<K> want(msg->key);
    msg->clearValue();
    Lock l(mu_);
    if (dev_) {
      if (want.empty()) { msg->addValue(SArray<V>()); return; }
      const int64* slot = slots(want, false);
      SArray<char> out = dev_.van->allocStaging(want.size() * sizeof(V));
      dev_.cuda->check(dev_.cuda->dkv_gather(out.data(), dval_.data(), slot, (int64)want.size(), sizeof(V),
                                             dev_.stream), "map gather");
      dev_.sync();
      msg->addValue(SArray<V>(out));
      return;
    }
    SArray<V> out(want.size());
    for (size_t i = 0; i < want.size(); ++i) {
      auto it = data_.find(want[i]);
      out[i] = it == data_.end() ? V() : it->second;
    }
    msg->addValue(out);
  }

  // a request combines (operator of the message), a pull reply overwrites my cached entries
  void setValue(const MessagePtr& msg) override {
    SArray<K> got(msg->key);
    if (got.empty()) return;
    CHECK(!msg->value.empty());
    CHECK_EQ(got.size() * sizeof(V), msg->value[0].size());
    const Operator op = msg->task.shared_para().op();
    const bool combine = msg->task.request();
    Lock l(mu_);
    if (dev_) {
      const int64* slot = slots(got, true);
      DevArray<V> tmp;
      const V* src = dev_.template onDevice<V>(msg->value[0], &tmp);
      dev_.cuda->check(dev_.cuda->dkv_apply(dval_.data(), slot, src, (int64)got.size(), 1, DevType<V>::id,
                                            combine ? (int)op + 1 : 0 /* ASSIGN */, dev_.stream), "map apply");
      dev_.sync();
      return;
    }
    SArray<V> vals(msg->value[0]);
    for (size_t i = 0; i < got.size(); ++i) {
      if (combine) compAssOp(data_[got[i]], vals[i], op);
      else data_[got[i]] = vals[i];
    }
  }

  void setReplica(const MessagePtr& msg) override {
    if (msg->value.empty()) return;
    SArray<K> key(msg->key);
    SArray<V> vals = dev_ ? dev_.template toHost<V>(msg->value[0]) : SArray<V>(msg->value[0]);
    const Operator op = msg->task.shared_para().op();
    Lock l(mu_);
    auto& rep = replica_[Range<K>(msg->task.key_range().template to<K>()).begin()];
    for (size_t i = 0; i < key.size(); ++i) compAssOp(rep[key[i]], vals[i], op);
  }
  void getReplica(const MessagePtr& msg) override {
    std::map<K, V> sorted;
    {
      Lock l(mu_);
      auto it = replica_.find(Range<K>(msg->task.key_range().template to<K>()).begin());
      if (it != replica_.end()) sorted.insert(it->second.begin(), it->second.end());
    }
    SArray<K> k;
    SArray<V> v;
    for (const auto& kv : sorted) { k.pushBack(kv.first); v.pushBack(kv.second); }
    msg->clearData();
    msg->setKey(k);
    msg->addValue(v);
  }
  void recoverFrom(const MessagePtr& msg) override {
    if (msg->value.empty()) return;
    MessagePtr as_reply(new Message(*msg));
    as_reply->task.set_request(false);
    setValue(as_reply);
  }

  // local access.  gpu mode: snapshot() downloads the table; operator[] / data() are host only.
  V& operator[](const K& k) { CHECK(!dev_) << "use snapshot() with -device gpu"; Lock l(mu_); return data_[k]; }
  size_t size() {
    if (dev_) return snapshot().size();
    Lock l(mu_);
    return data_.size();
  }
  std::unordered_map<K, V>& data() { CHECK(!dev_) << "use snapshot() with -device gpu"; return data_; }
  std::unordered_map<K, V> snapshot() {
    Lock l(mu_);
    if (!dev_) return data_;
    std::unordered_map<K, V> out;
    if (dkey_.empty()) return out;
    dev_.bind();
    const size_t cap = dkey_.size();
    DevArray<uint64> ok, cnt;
    DevArray<int64> os;
    ok.resize(cap, false, dev_.stream);
    os.resize(cap, false, dev_.stream);
    cnt.resize(1, true, dev_.stream);
    dev_.cuda->check(dev_.cuda->dmap_export(dkey_.data(), (int64)cap, ok.data(), os.data(), cnt.data(), dev_.stream), "export");
    uint64 n = 0;
    cnt.toHost(&n, dev_.stream);
    std::vector<uint64> hk(cap);
    std::vector<int64> hs(cap);
    std::vector<V> hv(cap);
    ok.toHost(hk.data(), dev_.stream);
    os.toHost(hs.data(), dev_.stream);
    dval_.toHost(hv.data(), dev_.stream);
    for (uint64 i = 0; i < n; ++i) out[(K)hk[i]] = hv[hs[i]];
    return out;
  }

  USING_SHARED_PARAMETER;

 protected:
  // gpu mode: slots of `keys` in the open-addressing table (inserted when `insert`), mu_ held.
  // The returned array lives until the next call.
  const int64* slots(const SArray<K>& keys, bool insert) {
    dev_.bind();
    if (dkey_.empty()) {
      size_t cap = 1;
      while (cap < (size_t)std::max(FLAGS_kvmap_capacity, 16)) cap <<= 1;
      dkey_.resize(cap, false, dev_.stream);
      dval_.resize(cap, true, dev_.stream);
      dev_.cuda->check(dev_.cuda->dmap_clear(dkey_.data(), (int64)cap, dev_.stream), "map clear");
      dflag_.resize(1, true, dev_.stream);
    }
    dprobe_.fromHost(reinterpret_cast<const uint64*>(keys.data()), keys.size(), dev_.stream);
    dslot_.resize(keys.size(), false, dev_.stream);
    dev_.cuda->check(dev_.cuda->dmap_slots(dkey_.data(), (int64)dkey_.size(), dprobe_.data(), (int64)keys.size(),
                                           insert ? 1 : 0, dslot_.data(), dflag_.data(), dev_.stream), "map slots");
    if (insert) {
      int32 full = 0;
      dflag_.toHost(&full, dev_.stream);
      CHECK_EQ(full, 0) << "KVMap table full: raise -kvmap_capacity (now " << dkey_.size() << ")";
    }
    return dslot_.data();
  }

  std::mutex mu_;
  std::unordered_map<K, V> data_;
  std::map<K, std::unordered_map<K, V>> replica_;
  DeviceContext dev_;
  DevArray<uint64> dkey_, dprobe_;
  DevArray<V> dval_;
  DevArray<int64> dslot_;
  DevArray<int32> dflag_;
};

// =================================================================================================
// KVStore: key -> entry with Entry::get(grad, state) / Entry::put(out, state).  Host only: the
// entries are arbitrary C++; the GPU counterpart of the async-SGD store is csrc/cuda/sgd.cu.
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

template <typename K, typename V, typename E, typename S>
class KVStore : public KVState<K, S> {
 public:
  KVStore(const string& my_name, const string& parent_name) : KVState<K, S>(my_name, parent_name) {}
  virtual ~KVStore() { this->stopExecutor(); }
  using KVState<K, S>::state_;

  MessagePtrList slice(const MessagePtr& msg, const KeyRangeList& krs) override {
    return sliceKeyOrderedMsg<K>(msg, krs);
  }
  void getValue(const MessagePtr& msg) override {
    SArray<K> want(msg->key);
    SArray<V> out(want.size());
    V* o = out.data();
    for (K k : want) table_[k].put(o++, &state_);
    msg->clearValue();
    msg->addValue(out);
  }
  void setValue(const MessagePtr& msg) override {
    SArray<K> got(msg->key);
    if (got.empty()) return;
    CHECK(!msg->value.empty());
    SArray<V> grad(msg->value[0]);
    CHECK_EQ(got.size(), grad.size());
    const V* g = grad.data();
    for (K k : got) table_[k].get(g++, &state_);
    state_.update();
  }
  // "key\tvalue" per weight that is neither zero nor NaN, keys ascending (the reference walks its
  // hash map, so its order is arbitrary)
  void writeToFile(const string& file) override {
    if (!dirExists(getPath(file))) createDir(getPath(file));
    std::ofstream out(file);
    CHECK(out.good()) << "cannot write " << file;
    std::map<K, V> sorted;
    for (auto& e : table_) {
      V v;
      e.second.put(&v, &state_);
      if (v != 0 && v == v) sorted[e.first] = v;
    }
    for (const auto& kv : sorted) out << kv.first << "\t" << kv.second << "\n";
  }
  size_t size() const { return table_.size(); }

 protected:
  std::unordered_map<K, E> table_;
};

}  // namespace PS
