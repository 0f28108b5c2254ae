// Message: a task header plus zero-copy key / value payloads, and the key-range slicer.
// ref src/system/message.h:18-165, src/system/message.cc:5-61
#pragma once
#include "proto/messages.h"
#include "util/shared_array.h"

namespace PS {

struct Message;
typedef shared_ptr<Message> MessagePtr;
typedef std::vector<MessagePtr> MessagePtrList;

struct Message {
  static const int kInvalidTime = -1;
  typedef std::function<void()> Callback;

  Message() {}
  explicit Message(const TaskView& tk) : task(tk) {}
  explicit Message(const NodeID& dest, int time = kInvalidTime, int wait_time = kInvalidTime)
      : recver(dest) {
    task.set_time(time);
    if (wait_time != kInvalidTime) task.add_wait_time(wait_time);
  }

  // ---- wire content ----
  Task task;
  SArray<char> key;
  std::vector<SArray<char>> value;

  // ---- local control (never sent) ----
  NodeID sender, recver, original_recver;
  bool replied = false;    // a reply was already sent for this request
  bool finished = true;    // false: the receiver will call finishIncomingTask itself
  bool valid = true;       // false: do not send, fabricate the reply locally (empty slice)
  bool terminate = false;  // tells the van's sending loop to exit
  bool wait = false;       // submit() blocks until the task finished
  Callback recv_handle;    // run on every reply
  Callback fin_handle;     // run once all members of the receiver group replied

  // ---- payload helpers ----
  template <typename T>
  void setKey(const SArray<T>& k) {
    task.set_key_type(encodeType<T>());
    task.set_has_key(true);
    key = SArray<char>(k);
  }
  void clearKey() {
    task.set_has_key(false);
    key.clear();
  }
  template <typename T>
  void addValue(const SArray<T>& v) {
    task.add_value_type(encodeType<T>());
    value.push_back(SArray<char>(v));
  }
  template <typename T>
  void addValue(const std::initializer_list<SArray<T>>& vs) {
    for (const auto& v : vs) addValue(v);
  }
  template <typename T>
  void addValue(const std::vector<SArray<T>>& vs) {
    for (const auto& v : vs) addValue(v);
  }
  void clearValue() {
    task.clear_value_type();
    value.clear();
  }
  void clearData() { clearKey(); clearValue(); }

  FilterConfigView addFilter(FilterConfigView::Type type) {
    FilterConfigView f = task.add_filter();
    f.set_type(type);
    return f;
  }

  size_t memSize() const {
    size_t n = key.size();
    for (const auto& v : value) n += v.size();
    return n;
  }

  string shortDebugString() const {
    std::ostringstream ss;
    if (key.size()) ss << "key [" << key.size() << "] ";
    if (value.size()) {
      ss << "value [";
      for (size_t i = 0; i < value.size(); ++i) ss << (i ? "," : "") << value[i].size();
      ss << "] ";
    }
    const int t = task.time();
    ss << "T: " << t << ", ";
    if (task.wait_time_size()) {
      ss << "wait_time: {";
      for (int i = 0; i < task.wait_time_size(); ++i) ss << (i ? "," : "") << task.wait_time(i);
      ss << "} ";
    }
    ss << sender << "=>" << recver << " ";
    if (!original_recver.empty()) ss << "(" << original_recver << ") ";
    ss << "task: " << task.ShortDebugString();
    return ss.str();
  }
  string debugString() const { return shortDebugString(); }
};

// Split a message whose keys are sorted into one piece per key range (usually the server
// shards).  Keys and every value array are segmented without copying; value arrays may hold
// k entries per key.  A piece that receives no key is marked valid=false so that no bytes
// travel but the group bookkeeping still sees an (instant) reply.
// ref src/system/message.h:120-159
template <typename K>
MessagePtrList sliceKeyOrderedMsg(const MessagePtr& msg, const std::vector<Range<Key>>& ranges) {
  const size_t n = ranges.size();
  MessagePtrList ret(n);
  std::vector<size_t> pos(n + 1, 0);
  SArray<K> key(msg->key);
  Range<Key> msg_range = msg->task.has_key_range() ? msg->task.key_range().template to<Key>()
                                                  : Range<Key>::all();
  for (size_t i = 0; i < n; ++i) {
    if (i == 0) {
      K lo = (K)ranges[0].begin();
      pos[0] = std::lower_bound(key.begin(), key.end(), lo) - key.begin();
    } else {
      CHECK_EQ(ranges[i - 1].end(), ranges[i].begin());
    }
    K hi = (K)ranges[i].end();
    // Range::all() ends at 2^64-1 exclusive by construction; the last shard also owns that key
    pos[i + 1] = (i + 1 == n && ranges[i].end() == kMaxKey)
                     ? key.size()
                     : std::lower_bound(key.begin(), key.end(), hi) - key.begin();
  }
  for (size_t i = 0; i < n; ++i) {
    MessagePtr piece(new Message());
    piece->task = msg->task;
    piece->sender = msg->sender;
    piece->recver = msg->recver;
    piece->original_recver = msg->original_recver;
    piece->wait = msg->wait;
    piece->finished = msg->finished;
    piece->recv_handle = msg->recv_handle;
    piece->fin_handle = msg->fin_handle;
    Range<Key> inter = ranges[i].setIntersection(msg_range);
    if (inter.empty()) {
      piece->valid = false;  // the task's key range does not touch this shard at all
    } else {
      piece->valid = true;
      if (msg->task.has_key()) {
        SizeR seg(pos[i], pos[i + 1]);
        piece->key = SArray<char>(key.segment(seg));
        if (key.size()) {
          for (const auto& v : msg->value) {
            const size_t kb = v.size() / key.size();  // bytes of this value array per key
            piece->value.push_back(v.segment(SizeR(seg.begin() * kb, seg.end() * kb)));
          }
        } else {
          for (const auto& v : msg->value) piece->value.push_back(v.segment(SizeR(0, 0)));
        }
      } else {
        piece->value = msg->value;
      }
      piece->task.mutable_key_range().from(inter);
    }
    ret[i] = piece;
  }
  return ret;
}

}  // namespace PS
