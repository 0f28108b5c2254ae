#include "system/runtime.h"
#include "system/peer_van.h"
#include "util/local_machine.h"

#include <list>

#include "data/common.h"
#include "system/heartbeat.h"

namespace PS {

DEFINE_int32(num_servers, 0, "number of servers");
DEFINE_int32(num_workers, 0, "number of clients");
DEFINE_int32(num_unused, 0, "number of unused nodes kept as hot spares");
DEFINE_int32(num_threads, 2, "number of computational threads");
DEFINE_string(app_name, "app", "the name of the root customer");
DEFINE_string(app_conf, "", "the string configuration of app");
DEFINE_string(app_file, "", "the configuration file of app");
DEFINE_bool(traffic_statistics, false, "print bytes sent / received at exit");
DEFINE_int32(report_interval, 0, "seconds between heartbeat reports to the scheduler; 0 disables");
DEFINE_bool(verbose, false, "chatty executor / data logs");
DEFINE_bool(log_to_file, false, "redirect log messages to ./log_<node id>");
DEFINE_bool(enable_fault_tolerance, false, "replace nodes whose heartbeat is overdue by unused nodes");
DEFINE_int32(num_replicas, 0, "number of replicas per server shard");
DEFINE_string(my_node, "role:SCHEDULER,hostname:'127.0.0.1',port:8000,id:'H'", "my node");
DEFINE_string(scheduler, "role:SCHEDULER,hostname:'127.0.0.1',port:8000,id:'H'", "the scheduler node");
DEFINE_string(server_master, "", "the master of servers (unused, kept for script compatibility)");
DEFINE_bool(print_van, false, "log every message to ./van_<node id>");
DEFINE_int32(bind_to, 0, "binding port (0: the port of -my_node)");
DEFINE_int32(my_rank, -1, "my rank among MPI peers (mpi launcher)");
DEFINE_string(interface, "", "network interface (mpi launcher)");
DEFINE_bool(key_cache, true, "enable caching keys during communication");
DEFINE_bool(message_compression, true, "snappy-compress every message payload");
// B200 data plane (no counterpart in the reference)
DEFINE_string(device, "cpu", "cpu: values in host memory, payloads over the host van; gpu: KV containers "
                             "keep their values in HBM and value frames travel through CUDA-IPC peer memory");
DEFINE_int32(gpu, -1, "CUDA device of this node (-1: the number in the node id modulo the device count)");
DEFINE_int32(peer_arena_mb, 256, "size of this node's staging arena for device value frames");
DEFINE_string(peer_dir, "", "rendezvous directory of the IPC handles (default /tmp/psb_peer_<scheduler port>)");
DEFINE_int32(kvmap_capacity, 1 << 20, "slots of a KVMap shard kept in HBM (-device gpu)");

// ---- natural node-id order ---------------------------------------------------------------------
bool NaturalLess::operator()(const string& a, const string& b) const {
  size_t i = 0, j = 0;
  while (i < a.size() && j < b.size()) {
    if (isdigit((unsigned char)a[i]) && isdigit((unsigned char)b[j])) {
      size_t i2 = i, j2 = j;
      while (i2 < a.size() && isdigit((unsigned char)a[i2])) ++i2;
      while (j2 < b.size() && isdigit((unsigned char)b[j2])) ++j2;
      string na = a.substr(i, i2 - i), nb = b.substr(j, j2 - j);
      na.erase(0, std::min(na.find_first_not_of('0'), na.size() - 1));
      nb.erase(0, std::min(nb.find_first_not_of('0'), nb.size() - 1));
      if (na.size() != nb.size()) return na.size() < nb.size();
      if (na != nb) return na < nb;
      i = i2;
      j = j2;
    } else {
      if (a[i] != b[j]) return a[i] < b[j];
      ++i;
      ++j;
    }
  }
  return a.size() - i < b.size() - j;
}

// =================================================================================================
// RNode
// =================================================================================================
RNode::RNode(const NodeView& node, Executor& exec) : id_(node.id()), rnode_(node), exec_(exec) {}
RNode::~RNode() {}

int RNode::size() { return role() == NodeView::GROUP ? (int)exec_.group(id_).size() : 1; }

int RNode::submit(const MessagePtr& msg) {
  CHECK(msg);
  MessagePtrList msgs = exec_.obj().slice(msg, exec_.keyRanges(id_));
  return submit(msgs);
}
int RNode::submit(const TaskView& task, const Message::Callback& recv_handle) {
  MessagePtr msg(new Message(task));
  if (recv_handle) msg->recv_handle = recv_handle;
  return submit(msg);
}
int RNode::submitAndWait(const TaskView& task, const Message::Callback& recv_handle) {
  MessagePtr msg(new Message(task));
  if (recv_handle) msg->recv_handle = recv_handle;
  msg->wait = true;
  return submit(msg);
}
int RNode::submit(const std::vector<Task>& tasks, const Message::Callback& recv_handle) {
  MessagePtrList msgs;
  for (const auto& t : tasks) {
    msgs.push_back(MessagePtr(new Message(t)));
    if (recv_handle) msgs.back()->recv_handle = recv_handle;
  }
  return submit(msgs);
}
int RNode::submitAndWait(const std::vector<Task>& tasks, const Message::Callback& recv_handle) {
  MessagePtrList msgs;
  for (const auto& t : tasks) {
    msgs.push_back(MessagePtr(new Message(t)));
    msgs.back()->wait = true;
    if (recv_handle) msgs.back()->recv_handle = recv_handle;
  }
  return submit(msgs);
}

int RNode::submit(MessagePtrList& msgs) {
  std::vector<RNodePtr> members = exec_.group(id_);
  CHECK_EQ(members.size(), msgs.size()) << "one message per member of " << id_ << " expected";
  if (msgs.empty()) return time();
  // an explicit timestamp must be shared by all pieces
  int t = Message::kInvalidTime;
  for (auto& m : msgs) {
    CHECK(m);
    if (m->task.has_time() && m->task.time() > Message::kInvalidTime) {
      if (t != Message::kInvalidTime) CHECK_EQ(t, m->task.time());
      t = m->task.time();
    }
  }
  {
    Lock l(mu_);
    if (t > Message::kInvalidTime) {
      time_ = std::max(t, time_);
    } else {
      if (role() == NodeView::GROUP)
        for (auto& w : members)
          if (w.get() != this) time_ = std::max(w->time(), time_);
      t = ++time_;
    }
    if (msgs[0]->fin_handle) msg_finish_handle_[t] = msgs[0]->fin_handle;
  }
  for (auto& m : msgs) {
    m->task.set_request(true);
    m->task.set_customer(exec_.obj().name());
    m->task.set_time(t);
    m->original_recver = id_;
    m->sender = exec_.myNode().id();
  }
  outgoing_task_.start(t);
  for (size_t i = 0; i < msgs.size(); ++i) {
    RNodePtr w = members[i];
    if (w.get() != this) w->outgoing_task_.start(t);
    // the member may have been swapped out (Executor::replace) after the group was listed:
    // data requests go to its replacement, control requests are not waited for
    RNodePtr successor;
    bool retired = false;
    {
      Lock l(w->mu_);
      retired = w->retired_;
      successor = w->replaced_by_;
      if (!retired) {
        msgs[i]->recver = w->id();
        w->time_ = std::max(t, w->time_);
        w->pending_msgs_[t] = msgs[i];
        if (msgs[i]->recv_handle) w->msg_receive_handle_[t] = msgs[i]->recv_handle;
      }
    }
    if (retired) {
      if (msgs[i]->task.type() != TaskView::CALL_CUSTOMER || !successor) {
        w->outgoing_task_.finish(t);
        continue;
      }
      {
        Lock l(successor->mu_);
        msgs[i]->recver = successor->id();
        successor->time_ = std::max(t, successor->time_);
        successor->pending_msgs_[t] = msgs[i];
        if (msgs[i]->recv_handle) successor->msg_receive_handle_[t] = msgs[i]->recv_handle;
      }
      successor->outgoing_task_.start(t);
      w = successor;
    }
    if (msgs[i]->valid) w->encodeFilter(msgs[i]);
    exec_.sys().queue(msgs[i]);
  }
  for (size_t i = 0; i < msgs.size(); ++i)
    if (msgs[i]->wait) members[i]->outgoing_task_.wait(t);
  return t;
}

void RNode::waitOutgoingTask(int time) {
  for (auto& w : exec_.group(id_)) w->outgoing_task_.wait(time);
}
bool RNode::tryWaitOutgoingTask(int time) {
  for (auto& w : exec_.group(id_))
    if (!w->outgoing_task_.tryWait(time)) return false;
  return true;
}
void RNode::finishOutgoingTask(int time) {
  for (auto& w : exec_.group(id_)) w->outgoing_task_.finish(time);
}
void RNode::waitIncomingTask(int time) {
  for (auto& w : exec_.group(id_)) w->incoming_task_.wait(time);
}
bool RNode::tryWaitIncomingTask(int time) {
  for (auto& w : exec_.group(id_))
    if (!w->incoming_task_.tryWait(time)) return false;
  return true;
}
void RNode::finishIncomingTask(int time) {
  for (auto& w : exec_.group(id_)) w->incoming_task_.finish(time);
  exec_.notify();
}

Filter* RNode::findFilter(const FilterConfigView& conf) {
  Lock l(filter_mu_);
  int id = conf.type();
  auto it = filter_.find(id);
  if (it == filter_.end()) it = filter_.emplace(id, unique_ptr<Filter>(Filter::create(conf))).first;
  return it->second.get();
}
void RNode::encodeFilter(const MessagePtr& msg) {
  TaskView tk = msg->task;
  // frames in device memory are never touched by the host: only key filters apply to them
  bool device_frames = false;
  for (const auto& v : msg->value) device_frames |= v.onDevice();
  if (exec_.sys().messageCompression() && msg->memSize() > 0 && !device_frames &&
      !Filter::find(FilterConfigView::COMPRESSING, msg).valid())
    msg->addFilter(FilterConfigView::COMPRESSING);
  for (int i = 0; i < tk.filter_size(); ++i) {
    if (device_frames && tk.filter(i).type() != FilterConfigView::KEY_CACHING) continue;
    findFilter(tk.filter(i))->encode(msg);
  }
}
void RNode::decodeFilter(const MessagePtr& msg) {
  TaskView tk = msg->task;
  bool device_frames = false;
  for (const auto& v : msg->value) device_frames |= v.onDevice();
  for (int i = tk.filter_size() - 1; i >= 0; --i) {
    if (device_frames && tk.filter(i).type() != FilterConfigView::KEY_CACHING) continue;
    findFilter(tk.filter(i))->decode(msg);
  }
}
MessagePtrList RNode::pendingMsgs() {
  Lock l(mu_);
  MessagePtrList r;
  for (auto& kv : pending_msgs_) r.push_back(kv.second);
  return r;
}

// =================================================================================================
// Executor
// =================================================================================================
static const std::vector<NodeID>& groupIDs() {
  static const std::vector<NodeID> ids = {kServerGroup, kWorkerGroup, kCompGroup,
                                          kReplicaGroup, kOwnerGroup, kLiveGroup};
  return ids;
}

void Executor::NodeInfo::addSubNode(const RNodePtr& s) {
  size_t pos = 0;
  while (pos < sub_nodes.size() && sub_nodes[pos]->keyRange().inLeft(s->keyRange())) ++pos;
  sub_nodes.insert(sub_nodes.begin() + pos, s);
  key_ranges.insert(key_ranges.begin() + pos, s->keyRange());
}
void Executor::NodeInfo::removeSubNode(const RNodePtr& s) {
  for (size_t i = 0; i < sub_nodes.size();) {
    if (sub_nodes[i] == s) {
      sub_nodes.erase(sub_nodes.begin() + i);
      key_ranges.erase(key_ranges.begin() + i);
    } else {
      ++i;
    }
  }
}

Executor::Executor(Customer& obj) : obj_(obj), sys_(obj.sys()) {
  my_node_ = sys_.myNode();
  for (const auto& id : groupIDs()) {
    Node n;
    n.set_role(NodeView::GROUP);
    n.set_id(id);
    add(n);
  }
  thread_ = std::thread(&Executor::run, this);
}
Executor::~Executor() { stop(); }

void Executor::stop() {
  if (!done_.exchange(true)) {
    notify();
    if (thread_.joinable()) thread_.join();
  }
}

void Executor::init(const std::vector<Node>& nodes) {
  for (const auto& n : nodes) add(n);
}

RNodePtr Executor::rnode(const NodeID& k) {
  Lock l(node_mu_);
  auto it = nodes_.find(k);
  return it == nodes_.end() ? RNodePtr() : it->second.node;
}
std::vector<RNodePtr> Executor::group(const NodeID& k) {
  Lock l(node_mu_);
  auto it = nodes_.find(k);
  CHECK(it != nodes_.end()) << obj_.name() << "@" << my_node_.id() << ": unknown node or group " << k;
  return it->second.sub_nodes;
}
KeyRangeList Executor::keyRanges(const NodeID& k) {
  Lock l(node_mu_);
  auto it = nodes_.find(k);
  CHECK(it != nodes_.end()) << "unknown node " << k;
  return it->second.key_ranges;
}

void Executor::add(const NodeView& node) {
  Lock l(node_mu_);
  if (node.id() == my_node_.id()) my_node_ = node;
  const NodeID id = node.id();
  RNodePtr w;
  auto it = nodes_.find(id);
  if (it != nodes_.end()) {  // update: re-insert so that group order follows the new range
    w = it->second.node;
    w->setNode(node);
    it->second.removeSubNode(w);
    for (const auto& gid : groupIDs()) nodes_[gid].removeSubNode(w);
  } else {
    w.reset(new RNode(node, *this));
    nodes_[id].node = w;
  }
  auto role = node.role();
  if (role != NodeView::GROUP) {
    nodes_[id].addSubNode(w);
    if (role != NodeView::UNUSED) nodes_[kLiveGroup].addSubNode(w);
  }
  if (role == NodeView::SERVER) {
    nodes_[kServerGroup].addSubNode(w);
    nodes_[kCompGroup].addSubNode(w);
  }
  if (role == NodeView::WORKER) {
    nodes_[kWorkerGroup].addSubNode(w);
    nodes_[kCompGroup].addSubNode(w);
  }
  // replica wiring (chain replication layout the reference sketches in executor.cc:112-138):
  // with r replicas, server i keeps copies of the r shards before it and is replicated on the
  // r servers after it
  if (my_node_.role() == NodeView::SERVER && role == NodeView::SERVER) {
    auto& servers = nodes_[kServerGroup].sub_nodes;
    int n = (int)servers.size(), nrep = std::min(FLAGS_num_replicas, n - 1), me = -1;
    for (int i = 0; i < n; ++i)
      if (servers[i]->id() == my_node_.id()) me = i;
    NodeInfo& rep = nodes_[kReplicaGroup];
    NodeInfo& own = nodes_[kOwnerGroup];
    rep.sub_nodes.clear(); rep.key_ranges.clear();
    own.sub_nodes.clear(); own.key_ranges.clear();
    for (int i = 1; me >= 0 && i <= nrep; ++i) {
      rep.addSubNode(servers[(me + i) % n]);
      own.addSubNode(servers[(me - i + n) % n]);
    }
  }
}

void Executor::remove(const NodeView& node) {
  Lock l(node_mu_);
  auto it = nodes_.find(node.id());
  if (it == nodes_.end()) return;
  RNodePtr w = it->second.node;
  for (const auto& gid : groupIDs()) nodes_[gid].removeSubNode(w);
  nodes_.erase(it);
}

void Executor::replace(const NodeView& dead, const NodeView& live) {
  // the live node takes over role, rank and key range of the dead one.  Data requests the dead
  // node never answered are re-sent to it with their original timestamps, and whoever waits on
  // the dead node's tracker is released when the replacement answers.  Control requests
  // (MANAGE: INIT / RUN ...) are not replayed - Postoffice::replaceNode drives the spare through
  // that sequence itself - they just stop being waited for.
  RNodePtr old = rnode(dead.id());
  remove(dead);
  add(live);
  RNodePtr nw = rnode(live.id());
  nw->replaces_ = old;
  MessagePtrList pending;
  if (old) {  // from here on RNode::submit redirects instead of registering on the dead node
    Lock l(old->mu_);
    old->retired_ = true;
    old->replaced_by_ = nw;
    for (auto& kv : old->pending_msgs_) pending.push_back(kv.second);
  }
  for (auto& m : pending) {
    const int t = m->task.time();
    if (m->task.type() != TaskView::CALL_CUSTOMER) {
      old->outgoing_task_.finish(t);
      continue;
    }
    MessagePtr again(new Message(*m));
    again->recver = live.id();
    {
      Lock l(nw->mu_);
      nw->pending_msgs_[t] = again;
      nw->time_ = std::max(nw->time_, t);
    }
    nw->outgoing_task_.start(t);
    sys_.queue(again);
  }
}

void Executor::updateKeyRange(const NodeView& node) { add(node); }

string Executor::lastRecvReply() {
  CHECK(active_msg_ && !active_msg_->task.request());
  return active_msg_->task.msg();
}

void Executor::accept(const MessagePtr& msg) {
  RNodePtr sender = rnode(msg->sender);
  if (!sender) {
    LL << my_node_.id() << ": message from unknown node " << msg->sender << " dropped: "
       << msg->shortDebugString();
    return;
  }
  sender->decodeFilter(msg);
  {
    Lock l(recved_msg_mu_);
    recved_msgs_.push_back(msg);
  }
  dag_cond_.notify_all();
}

// first message that is a reply, or a request whose wait_time dependencies (earlier tasks of
// the same sender) have all finished
bool Executor::pickMessage() {
  for (auto it = recved_msgs_.begin(); it != recved_msgs_.end(); ++it) {
    MessagePtr& msg = *it;
    RNodePtr sender = rnode(msg->sender);
    if (!sender) {
      recved_msgs_.erase(it);
      return false;
    }
    bool ok = !msg->task.request();
    if (!ok) {
      ok = true;
      for (int i = 0; i < msg->task.wait_time_size(); ++i) {
        int wt = msg->task.wait_time(i);
        if (wt > Message::kInvalidTime && !sender->tryWaitIncomingTask(wt)) ok = false;
      }
    }
    if (ok) {
      active_msg_ = msg;
      recved_msgs_.erase(it);
      return true;
    }
  }
  return false;
}

void Executor::run() {
  Postoffice::setCurrent(&sys_);
  while (!done_) {
    {
      std::unique_lock<std::mutex> lk(recved_msg_mu_);
      if (!pickMessage()) {
        if (done_) break;
        dag_cond_.wait_for(lk, std::chrono::milliseconds(200));
        continue;
      }
    }
    const bool req = active_msg_->task.request();
    const int t = active_msg_->task.time();
    RNodePtr sender = rnode(active_msg_->sender);
    if (!sender) continue;
    if (FLAGS_verbose) LI << obj_.name() << "@" << my_node_.id() << " runs " << active_msg_->shortDebugString();
    if (req) {
      sender->incoming_task_.start(t);
      obj_.process(active_msg_);
      if (active_msg_->finished) {
        sender->incoming_task_.finish(t);
        notify();
        if (!active_msg_->replied) sys_.reply(active_msg_->sender, active_msg_->task);
      }
    } else {
      // a pull reply travels as CALL_CUSTOMER with request=false and carries data for the
      // customer; a plain acknowledgement has type REPLY and only drives the bookkeeping
      if (active_msg_->task.type() != TaskView::REPLY) obj_.process(active_msg_);
      Message::Callback h;
      {
        Lock l(sender->mu_);
        auto b = sender->msg_receive_handle_.find(t);
        if (b != sender->msg_receive_handle_.end()) {
          h = b->second;
          sender->msg_receive_handle_.erase(b);
        }
      }
      if (h) h();
      NodeID original;
      {
        Lock l(sender->mu_);
        auto it = sender->pending_msgs_.find(t);
        if (it == sender->pending_msgs_.end()) {
          LL << my_node_.id() << ": unexpected reply from " << sender->id() << " at time " << t;
          continue;
        }
        original = it->second->original_recver;
        sender->pending_msgs_.erase(it);
      }
      sender->outgoing_task_.finish(t);
      if (sender->replaces_) sender->replaces_->outgoing_task_.finish(t);
      RNodePtr o = rnode(original);
      if (!o) continue;
      if (o->tryWaitOutgoingTask(t)) {
        if (o.get() != sender.get()) o->outgoing_task_.finish(t);  // the group's own tracker
        Message::Callback f;
        {
          Lock l(o->mu_);
          auto a = o->msg_finish_handle_.find(t);
          if (a != o->msg_finish_handle_.end()) {
            f = a->second;
            o->msg_finish_handle_.erase(a);
          }
        }
        if (f) f();
      }
    }
  }
}

// =================================================================================================
// Customer / App
// =================================================================================================
Customer::Customer(const string& name) : sys_(Postoffice::current()), name_(name) {
  exec_.reset(new Executor(*this));
  sys_.yp().add(this);
}
Customer::Customer(const string& name, const string& parent)
    : sys_(Postoffice::current()), name_(name) {
  exec_.reset(new Executor(*this));
  sys_.yp().add(this);
  // a child sees the same nodes as its parent (ref customer.cc:9-14)
  Customer* p = sys_.yp().customer(parent);
  if (p) {
    sys_.yp().depends(name, parent);
    for (const auto& n : sys_.yp().nodes()) exec_->add(n);
  }
}
Customer::~Customer() {
  stopExecutor();
  sys_.yp().remove(name_);
}
void Customer::stopExecutor() { if (exec_) exec_->stop(); }
NodeID Customer::schedulerID() { return sys_.scheduler().id(); }

MessagePtrList Customer::slice(const MessagePtr& msg, const KeyRangeList& krs) {
  MessagePtrList r(krs.size());
  for (size_t i = 0; i < krs.size(); ++i) r[i] = MessagePtr(new Message(*msg));
  return r;
}

App::App(const string& name) : Customer(name) {}

static std::mutex g_factory_mu;
static App::Factory g_factory;
void App::setFactory(const Factory& f) {
  Lock l(g_factory_mu);
  g_factory = f;
}
App* App::create(const string& name, const string& conf) {
  const App::Factory& local = Postoffice::current().options().factory;
  if (local) return local(name, conf);
  Lock l(g_factory_mu);
  CHECK(g_factory) << "no App factory registered (App::setFactory)";
  return g_factory(name, conf);
}

// =================================================================================================
// YellowPages
// =================================================================================================
void YellowPages::add(Customer* obj) {
  {
    Lock l(mu_);
    CHECK_EQ(customers_.count(obj->name()), (size_t)0) << "customer " << obj->name() << " exists";
    customers_[obj->name()] = obj;
  }
  cv_.notify_all();
}
void YellowPages::remove(const string& name) {
  Lock l(mu_);
  customers_.erase(name);
}
Customer* YellowPages::customer(const string& name) {
  // a message for a customer that this node is still constructing waits a little
  std::unique_lock<std::mutex> l(mu_);
  cv_.wait_for(l, std::chrono::seconds(20), [&] { return customers_.count(name) > 0; });
  auto it = customers_.find(name);
  return it == customers_.end() ? nullptr : it->second;
}
Customer* YellowPages::customerIfPresent(const string& name) {
  Lock l(mu_);
  auto it = customers_.find(name);
  return it == customers_.end() ? nullptr : it->second;
}
void YellowPages::depends(const string& child, const string& parent) {
  Lock l(mu_);
  relations_[parent].push_back(child);
}
std::vector<string> YellowPages::childern(const string& parent) {
  Lock l(mu_);
  auto it = relations_.find(parent);
  return it == relations_.end() ? std::vector<string>() : it->second;
}
void YellowPages::addNode(const NodeView& node) {
  Lock l(mu_);
  auto it = nodes_.find(node.id());
  if (it == nodes_.end()) {
    if (node.role() == NodeView::WORKER) ++num_workers_;
    if (node.role() == NodeView::SERVER) ++num_servers_;
    van_->connect(node);
  } else if (it->second.role() != node.role()) {
    if (it->second.role() == NodeView::WORKER) --num_workers_;
    if (it->second.role() == NodeView::SERVER) --num_servers_;
    if (node.role() == NodeView::WORKER) ++num_workers_;
    if (node.role() == NodeView::SERVER) ++num_servers_;
  }
  nodes_[node.id()] = node;
}
void YellowPages::removeNode(const NodeID& id) {
  Lock l(mu_);
  auto it = nodes_.find(id);
  if (it == nodes_.end()) return;
  if (it->second.role() == NodeView::WORKER) --num_workers_;
  if (it->second.role() == NodeView::SERVER) --num_servers_;
  nodes_.erase(it);
}
std::vector<Node> YellowPages::nodes() {
  Lock l(mu_);
  std::vector<Node> r;
  for (const auto& kv : nodes_) r.push_back(kv.second);
  return r;
}
bool YellowPages::hasNode(const NodeID& id) {
  Lock l(mu_);
  return nodes_.count(id) > 0;
}

// =================================================================================================
// Postmaster
// =================================================================================================
std::vector<DataConfig> Postmaster::partitionData(const DataConfigView& conf, int num_workers) {
  DataConfig data = searchFiles(conf);
  LI << "Found " << data.file_size() << " files";
  std::vector<DataConfig> ret = divideFiles(data, num_workers);
  int n = 0;
  for (const auto& p : ret) n += p.file_size();
  LI << "Assign " << n << " files to " << num_workers << " workers";
  return ret;
}
std::vector<Node> Postmaster::partitionServerKeyRange(const std::vector<Node>& nodes,
                                                      Range<Key> range) {
  int num_servers = 0;
  std::vector<Node> ret = nodes;
  for (const auto& o : ret) num_servers += o.role() == NodeView::SERVER;
  int s = 0;
  for (auto& o : ret) {
    Range<Key> kr = o.role() != NodeView::SERVER ? range : range.evenDivide(num_servers, s++);
    o.mutable_key().from(kr);
  }
  return ret;
}
std::vector<Node> Postmaster::assignNodeRank(const std::vector<Node>& nodes) {
  int server_r = 0, worker_r = 0, other_r = 0;
  std::vector<Node> ret = nodes;
  for (auto& o : ret) {
    int k = 0;
    switch (o.role()) {
      case NodeView::SERVER: k = server_r++; break;
      case NodeView::WORKER: k = worker_r++; break;
      case NodeView::SCHEDULER: k = 0; break;
      default: k = other_r++; break;
    }
    o.set_rank(k);
  }
  return ret;
}

// =================================================================================================
// Postoffice
// =================================================================================================
static thread_local Postoffice* tls_current = nullptr;

Postoffice::Postoffice() : hb_info_(new HeartbeatInfo()), dashboard_(new Dashboard()) {}
Postoffice::~Postoffice() {
  done_ = true;
  if (started_) {
    yp_.van_->shutdown();
    MessagePtr stopmsg(new Message());
    stopmsg->terminate = true;
    sending_queue_.push(stopmsg);
  }
  if (recv_thread_.joinable()) recv_thread_.join();
  if (send_thread_.joinable()) send_thread_.join();
  if (hb_thread_.joinable()) hb_thread_.join();
  app_.reset();
}
Postoffice& Postoffice::instance() {
  static Postoffice p;
  return p;
}
Postoffice& Postoffice::current() { return tls_current ? *tls_current : instance(); }
void Postoffice::setCurrent(Postoffice* p) { tls_current = p; }

static bool parseNode(const string& text, Node* node) {
  string err;
  if (!node->raw()->parse(text, &err)) {
    LL << "cannot parse node '" << text << "': " << err;
    return false;
  }
  return true;
}

// MPI-style launch: every process gets the same command line plus -my_rank.  Rank 0 is the
// scheduler, ranks 1..W the workers, then the servers, then the hot spares; each one finds
// its own address and a free port (ref van.cc:235-278, script/mpi_node.sh).
static Node assembleMyNode(const Node& scheduler) {
  if (FLAGS_my_rank == 0) return scheduler;
  Node node;
  const int r = FLAGS_my_rank;
  if (r <= FLAGS_num_workers) {
    node.set_role(NodeView::WORKER);
    node.set_id("W" + std::to_string(r - 1));
  } else if (r <= FLAGS_num_workers + FLAGS_num_servers) {
    node.set_role(NodeView::SERVER);
    node.set_id("S" + std::to_string(r - FLAGS_num_workers - 1));
  } else {
    node.set_role(NodeView::UNUSED);
    node.set_id("U" + std::to_string(r - FLAGS_num_workers - FLAGS_num_servers - 1));
  }
  string interface = FLAGS_interface, ip;
  if (interface.empty()) {
    CHECK(LocalMachine::pickupAvailableInterfaceAndIP(&interface, &ip)) << "no usable network interface";
  } else {
    ip = LocalMachine::IP(interface);
    CHECK(!ip.empty()) << "interface [" << interface << "] has no IPv4 address";
  }
  const unsigned short port = LocalMachine::pickupAvailablePort();
  CHECK_NE(port, 0) << "no free port on " << ip;
  node.set_hostname(ip);
  node.set_port(port);
  return node;
}

void Postoffice::start(int* argc, char*** argv) {
  string err;
  CHECK(FlagRegistry::get().parse(argc, argv, true, &err)) << err;
  PostofficeOptions o;
  CHECK(parseNode(FLAGS_scheduler, &o.scheduler));
  if (FLAGS_my_rank >= 0) {
    o.my_node = assembleMyNode(o.scheduler);
  } else {
    CHECK(parseNode(FLAGS_my_node, &o.my_node));
  }
  if (FLAGS_bind_to) o.my_node.set_port(FLAGS_bind_to);
  o.num_servers = FLAGS_num_servers;
  o.num_workers = FLAGS_num_workers;
  o.num_unused = FLAGS_num_unused;
  o.app_name = FLAGS_app_name;
  o.app_conf = FLAGS_app_conf;
  o.app_file = FLAGS_app_file;
  o.traffic_statistics = FLAGS_traffic_statistics;
  o.print_van = FLAGS_print_van;
  o.message_compression = FLAGS_message_compression;
  o.report_interval = FLAGS_report_interval;
  if (FLAGS_log_to_file) {
    FILE* f = fopen(("./log_" + o.my_node.id()).c_str(), "w");
    if (f) LogConfig::get().sink = f;
  }
  start(o);
}

void Postoffice::start(const PostofficeOptions& opt) {
  opt_ = opt;
  my_id_ = opt_.my_node.id();
  my_role_ = (int)opt_.my_node.role();
  setCurrent(this);
  const bool scheduler = my_role_ == (int)NodeView::SCHEDULER;
  unique_ptr<Van> van;
  if (opt_.fabric) van.reset(new InprocVan(opt_.fabric));
  else van.reset(new TcpVan());
  if (FLAGS_device == "gpu") {
    // the GPU data plane: headers over the van above, device value frames through peer memory
    int gpu = FLAGS_gpu;
    if (gpu < 0) {
      gpu = 0;
      for (char ch : my_id_) if (ch >= '0' && ch <= '9') gpu = gpu * 10 + (ch - '0');
    }
    string dir = FLAGS_peer_dir.empty() ? "/tmp/psb_peer_" + to_string(opt_.scheduler.port()) : FLAGS_peer_dir;
    van.reset(new PeerVan(std::move(van), gpu, (size_t)std::max(FLAGS_peer_arena_mb, 8) << 20, dir));
  }
  yp_.setVan(std::move(van));
  CHECK(yp_.van().init(opt_.my_node)) << "van init failed for " << opt_.my_node.ShortDebugString();
  if (opt_.print_van) yp_.van().enableLog("van_" + my_id_);
  yp_.addNode(opt_.my_node);
  if (!scheduler) yp_.addNode(opt_.scheduler);
  hb_info_->init(opt_.my_node.hostname());
  started_ = true;

  if (scheduler) {
    if (!opt_.app_file.empty())
      CHECK(readFileToString(opt_.app_file, &opt_.app_conf)) << "failed to read " << opt_.app_file;
    app_.reset(App::create(opt_.app_name, opt_.app_conf));
    CHECK(app_) << "failed to create [" << opt_.app_name << "] with conf\n" << opt_.app_conf;
  }
  recv_thread_ = std::thread(&Postoffice::recv, this);
  send_thread_ = std::thread(&Postoffice::send, this);
  if (opt_.report_interval > 0) hb_thread_ = std::thread(&Postoffice::heartbeat, this);

  if (!scheduler) {
    // announce myself; the scheduler answers with ManageApp::ADD (creates my app), then
    // ManageNode::ADD (full node table), INIT and RUN
    Task task;
    task.set_type(TaskView::MANAGE);
    task.set_request(true);
    task.set_time(0);
    task.set_customer(opt_.app_name);
    task.mutable_mng_node().set_cmd(ManageNodeView::CONNECT);
    task.mutable_mng_node().add_node().raw()->copyFrom(*myNode().raw());
    MessagePtr msg(new Message(task));
    msg->sender = my_id_;
    msg->recver = opt_.scheduler.id();
    queue(msg);

    init_app_promise_.get_future().wait();
    if (done_) return;
    CHECK(app_) << "INIT before ADD";
    app_->init();
    bool recover = false;
    {
      Lock l(mu_);
      recover = init_msg_ && init_msg_->task.msg() == "recover";
    }
    if (recover) app_->recover();
    finishBarrier(0);
    run_app_promise_.get_future().wait();
    if (done_) return;
    app_->run();
    finishBarrier(1);
    return;
  }

  // ---- scheduler ----
  {
    Task add;
    add.set_customer(app_->name());
    add.mutable_mng_node().set_cmd(ManageNodeView::ADD);
    add.mutable_mng_node().add_node().raw()->copyFrom(*myNode().raw());
    manageNode(add);
  }
  if (opt_.num_workers + opt_.num_servers == 0) {
    app_->init();
    app_->run();
    return;
  }
  nodes_are_ready_.get_future().wait();
  LI << "Scheduler has connected " << opt_.num_servers << " servers and " << opt_.num_workers
     << " workers";
  // every computation node has created its app (ManageApp::ADD was sent with time 1)
  app_->port(kCompGroup)->waitOutgoingTask(1);

  std::vector<Node> nodes = yp_.nodes();
  nodes = Postmaster::partitionServerKeyRange(nodes, Range<Key>::all());
  nodes = Postmaster::assignNodeRank(nodes);
  Task add;
  add.set_type(TaskView::MANAGE);
  add.mutable_mng_node().set_cmd(ManageNodeView::ADD);
  for (const auto& n : nodes) add.mutable_mng_node().add_node().raw()->copyFrom(*n.raw());
  {
    Task mine = add;
    mine.set_customer(app_->name());
    manageNode(mine);  // ranks and ranges for my own table too
  }
  app_->port(kCompGroup)->submitAndWait(add);

  Task init;
  init.set_type(TaskView::MANAGE);
  init.mutable_mng_app().set_cmd(ManageAppView::INIT);
  int init_wait = app_->port(kCompGroup)->submit(init);
  app_->init();
  app_->port(kCompGroup)->waitOutgoingTask(init_wait);

  Task run;
  run.set_type(TaskView::MANAGE);
  run.mutable_mng_app().set_cmd(ManageAppView::RUN);
  int run_wait = app_->port(kCompGroup)->submit(run);
  app_->run();
  app_->port(kCompGroup)->waitOutgoingTask(run_wait);
}

// the INIT / RUN requests stay unfinished until the local app returned from init() / run()
void Postoffice::finishBarrier(int which) {
  MessagePtr m;
  {
    Lock l(mu_);
    m = which == 0 ? init_msg_ : run_msg_;
  }
  if (!m) return;
  Customer* c = yp_.customer(m->task.customer());
  if (c) {
    RNodePtr r = c->exec().rnode(m->sender);
    if (r) r->finishIncomingTask(m->task.time());
  }
  reply(m->sender, m->task);
}

void Postoffice::stop() {
  const bool scheduler = my_role_ == (int)NodeView::SCHEDULER;
  if (scheduler) {
    if (opt_.num_workers + opt_.num_servers) {
      std::unique_lock<std::mutex> l(mu_);
      cv_.wait(l, [&] { return num_done_ >= opt_.num_workers + opt_.num_servers; });
    }
    Task terminate;
    terminate.set_type(TaskView::TERMINATE);
    if (opt_.num_workers + opt_.num_servers) app_->port(kLiveGroup)->submit(terminate);
    std::this_thread::sleep_for(std::chrono::milliseconds(2));
    if (opt_.traffic_statistics) yp_.van().statistic();
    LI << "System stopped";
    done_ = true;
  } else {
    Task done;
    done.set_type(TaskView::MANAGE);
    done.mutable_mng_app().set_cmd(ManageAppView::DONE);
    app_->port(app_->schedulerID())->submit(done);
    std::unique_lock<std::mutex> l(mu_);
    cv_.wait(l, [&] { return done_.load(); });
  }
}

void Postoffice::reply(const NodeID& recver, const TaskView& task, const string& reply_msg) {
  if (!task.request()) return;
  Task tk;
  tk.set_customer(task.customer());
  tk.set_request(false);
  tk.set_type(TaskView::REPLY);
  if (!reply_msg.empty()) tk.set_msg(reply_msg);
  tk.set_time(task.time());
  MessagePtr re(new Message(tk));
  re->sender = my_id_;
  re->recver = recver;
  queue(re);
}

void Postoffice::queue(const MessagePtr& msg) {
  if (msg->valid) {
    sending_queue_.push(msg);
    return;
  }
  // nothing to send (the piece intersects no key of the receiver): answer locally
  Task tk;
  tk.set_customer(msg->task.customer());
  tk.set_request(false);
  tk.set_type(TaskView::REPLY);
  tk.set_time(msg->task.time());
  MessagePtr reply(new Message(tk));
  reply->sender = msg->recver;
  reply->recver = msg->sender;
  Customer* c = yp_.customer(tk.customer());
  CHECK(c) << "unknown customer " << tk.customer();
  c->exec().accept(reply);
}

void Postoffice::send() {
  MessagePtr msg;
  while (true) {
    sending_queue_.wait_and_pop(msg);
    if (msg->terminate) break;
    size_t bytes = 0;
    Status st = yp_.van().send(msg, &bytes);
    if (!st.ok()) LL << "sending " << msg->shortDebugString() << " failed: " << st.ToString();
    hb_info_->increaseOutBytes(bytes);
  }
}

void Postoffice::recv() {
  setCurrent(this);
  while (true) {
    MessagePtr msg;
    size_t bytes = 0;
    Status st = yp_.van().recv(&msg, &bytes);
    if (!st.ok()) break;
    hb_info_->increaseInBytes(bytes);
    TaskView tk = msg->task;
    const bool request = tk.request();
    auto type = tk.type();
    if (type == TaskView::CALL_CUSTOMER || type == TaskView::REPLY) {
      Customer* c = yp_.customer(tk.customer());
      if (!c) {
        LL << my_id_ << ": customer [" << tk.customer() << "] does not exist, dropped "
           << msg->shortDebugString();
        continue;
      }
      c->exec().accept(msg);
      continue;
    }
    if (type == TaskView::HEARTBEATING) {
      HeartbeatReport rep;
      if (rep.ParseFromString(tk.msg())) dashboard_->addReport(msg->sender, rep);
      continue;
    }
    if (type == TaskView::TERMINATE) {
      if (opt_.traffic_statistics) yp_.van().statistic();
      {
        Lock l(mu_);
        done_ = true;
      }
      cv_.notify_all();
      // wake a main thread that still waits for INIT / RUN
      try { init_app_promise_.set_value(); } catch (...) {}
      try { run_app_promise_.set_value(); } catch (...) {}
      break;
    }
    if (type == TaskView::MANAGE && request) {
      if (tk.has_mng_app()) {
        auto cmd = tk.mng_app().cmd();
        if (cmd == ManageAppView::INIT || cmd == ManageAppView::RUN) {
          // answered by finishBarrier once init() / run() returned
          {
            Lock l(mu_);
            (cmd == ManageAppView::INIT ? init_msg_ : run_msg_) = msg;
          }
          try {
            (cmd == ManageAppView::INIT ? init_app_promise_ : run_app_promise_).set_value();
          } catch (const std::future_error&) {
            reply(msg->sender, tk);  // a repeated INIT / RUN: nothing left to do
          }
          continue;
        }
        manageApp(tk);
      }
      if (tk.has_mng_node()) {
        manageNode(tk);
        // a CONNECT does not come from an RNode::submit, nobody waits for its reply
        if (tk.mng_node().cmd() == ManageNodeView::CONNECT) continue;
      }
      // mark finished in the app's executor and acknowledge
      Customer* c = yp_.customerIfPresent(tk.customer());
      if (c) {
        RNodePtr r = c->exec().rnode(msg->sender);
        if (r) r->finishIncomingTask(tk.time());
      }
      reply(msg->sender, tk);
    }
  }
}

void Postoffice::manageApp(const TaskView& tk) {
  auto cmd = tk.mng_app().cmd();
  if (cmd == ManageAppView::ADD) {
    if (!app_) {
      app_.reset(App::create(tk.customer(), tk.mng_app().conf()));
      CHECK(app_) << "failed to create app " << tk.customer();
    }
  } else if (cmd == ManageAppView::DONE) {
    {
      Lock l(mu_);
      ++num_done_;
    }
    cv_.notify_all();
  } else if (cmd == ManageAppView::REMOVE) {
    app_.reset();
  }
}

void Postoffice::manageNode(const TaskView& tk) {
  ManageNodeView mng = tk.mng_node();
  switch (mng.cmd()) {
    case ManageNodeView::CONNECT: {
      CHECK(my_role_ == (int)NodeView::SCHEDULER);
      CHECK_EQ(mng.node_size(), 1);
      NodeView node = mng.node(0);
      if (node.role() == NodeView::UNUSED) {  // hot spare: remember it, do not start an app yet
        Lock l(mu_);
        unused_nodes_.push_back(Node(node));
        yp_.addNode(node);
        break;
      }
      Task add;
      add.set_customer(app_->name());
      add.mutable_mng_node().set_cmd(ManageNodeView::ADD);
      add.mutable_mng_node().add_node().raw()->copyFrom(*node.raw());
      manageNode(add);
      Task task;
      task.set_type(TaskView::MANAGE);
      task.set_time(1);
      task.mutable_mng_app().set_cmd(ManageAppView::ADD);
      task.mutable_mng_app().set_conf(opt_.app_conf);
      app_->port(node.id())->submit(task);
      bool ready = false;
      {
        Lock l(mu_);
        if (node.role() == NodeView::WORKER) ++num_connected_workers_;
        if (node.role() == NodeView::SERVER) ++num_connected_servers_;
        ready = num_connected_workers_ == opt_.num_workers &&
                num_connected_servers_ == opt_.num_servers;
      }
      if (ready) nodes_are_ready_.set_value();
      break;
    }
    case ManageNodeView::ADD:
    case ManageNodeView::UPDATE: {
      Customer* obj = yp_.customer(tk.customer());
      CHECK(obj) << "customer [" << tk.customer() << "] does not exist";
      for (int i = 0; i < mng.node_size(); ++i) {
        NodeView node = mng.node(i);
        yp_.addNode(node);
        if (node.id() == my_id_) {
          Lock l(my_node_mu_);
          opt_.my_node = node;
          my_role_ = (int)node.role();
        }
        obj->exec().add(node);
        for (const auto& c : yp_.childern(obj->name())) {
          Customer* child = yp_.customer(c);
          if (child) child->exec().add(node);
        }
      }
      break;
    }
    case ManageNodeView::REPLACE: {  // node(0) died, node(1) takes over its role / rank / range
      CHECK_EQ(mng.node_size(), 2);
      yp_.removeNode(mng.node(0).id());
      yp_.addNode(mng.node(1));
      // a spare learns its new role here, before its app exists (the factory looks at the role)
      if (mng.node(1).id() == my_id_) {
        Lock l(my_node_mu_);
        opt_.my_node = mng.node(1);
        my_role_ = (int)mng.node(1).role();
      }
      Customer* obj = yp_.customerIfPresent(tk.customer());
      if (!obj) break;
      obj->exec().replace(mng.node(0), mng.node(1));
      for (const auto& c : yp_.childern(obj->name())) {
        Customer* child = yp_.customer(c);
        if (child) child->exec().replace(mng.node(0), mng.node(1));
      }
      break;
    }
    case ManageNodeView::REMOVE: {
      Customer* obj = yp_.customer(tk.customer());
      for (int i = 0; i < mng.node_size(); ++i) {
        yp_.removeNode(mng.node(i).id());
        if (!obj) continue;
        obj->exec().remove(mng.node(i));
        for (const auto& c : yp_.childern(obj->name())) {
          Customer* child = yp_.customer(c);
          if (child) child->exec().remove(mng.node(i));
        }
      }
      break;
    }
  }
}

void Postoffice::replaceNode(const NodeView& dead, const NodeView& live) {
  CHECK(my_role_ == (int)NodeView::SCHEDULER);
  Node nw(live);
  nw.set_role(dead.role());
  nw.set_rank(dead.rank());
  nw.mutable_key().raw()->copyFrom(*dead.key().raw());
  Task t;
  t.set_type(TaskView::MANAGE);
  t.set_customer(app_->name());
  t.mutable_mng_node().set_cmd(ManageNodeView::REPLACE);
  t.mutable_mng_node().add_node().raw()->copyFrom(*dead.raw());
  t.mutable_mng_node().add_node().raw()->copyFrom(*nw.raw());
  manageNode(t);
  // the spare needs its new identity, the app, then the node table (ranks, key ranges, replica
  // wiring) ...
  app_->port(nw.id())->submitAndWait(t);
  Task addapp;
  addapp.set_type(TaskView::MANAGE);
  addapp.mutable_mng_app().set_cmd(ManageAppView::ADD);
  addapp.mutable_mng_app().set_conf(opt_.app_conf);
  app_->port(nw.id())->submitAndWait(addapp);
  Task table;
  table.set_type(TaskView::MANAGE);
  table.mutable_mng_node().set_cmd(ManageNodeView::ADD);
  for (const auto& n : yp_.nodes())
    if (n.role() != NodeView::UNUSED) table.mutable_mng_node().add_node().raw()->copyFrom(*n.raw());
  app_->port(nw.id())->submitAndWait(table);
  // ... everybody else swaps the dead node for it (unanswered requests are re-sent to it) ...
  app_->port(kLiveGroup)->submit(t);
  // ... and it initialises, reloads the shard from the replica holders and runs
  Task init;
  init.set_type(TaskView::MANAGE);
  init.set_msg("recover");
  init.mutable_mng_app().set_cmd(ManageAppView::INIT);
  app_->port(nw.id())->submitAndWait(init);
  Task run;
  run.set_type(TaskView::MANAGE);
  run.mutable_mng_app().set_cmd(ManageAppView::RUN);
  app_->port(nw.id())->submit(run);
}

string Postoffice::dashboardReport() { return dashboard_->report(); }

void Postoffice::kill() {
  {
    Lock l(mu_);
    done_ = true;
  }
  cv_.notify_all();
  try { init_app_promise_.set_value(); } catch (...) {}
  try { run_app_promise_.set_value(); } catch (...) {}
  if (started_) yp_.van_->shutdown();
}

void Postoffice::removeNode(const NodeView& dead) {
  Task t;
  t.set_type(TaskView::MANAGE);
  t.set_customer(app_->name());
  t.mutable_mng_node().set_cmd(ManageNodeView::REMOVE);
  t.mutable_mng_node().add_node().raw()->copyFrom(*dead.raw());
  manageNode(t);
  app_->port(kLiveGroup)->submit(t);
}

std::vector<string> Postoffice::deadNodes(int timeout_sec) { return dashboard_->deadNodes(timeout_sec); }

// periodic heartbeat: every node reports resource usage to the scheduler, which keeps the
// dashboard and (with -enable_fault_tolerance) swaps overdue nodes for unused ones
void Postoffice::heartbeat() {
  const bool scheduler = my_role_ == (int)NodeView::SCHEDULER;
  int tick = 0;
  while (!done_) {
    for (int i = 0; i < opt_.report_interval * 10 && !done_; ++i)
      std::this_thread::sleep_for(std::chrono::milliseconds(100));
    if (done_) break;
    HeartbeatReport rep = hb_info_->get();
    if (scheduler) {
      dashboard_->addReport(my_id_, rep);
      if (++tick % 5 == 0 && FLAGS_verbose) fprintf(stderr, "%s\n", dashboard_->report().c_str());
      if (FLAGS_enable_fault_tolerance) {
        for (const string& id : dashboard_->deadNodes(opt_.report_interval * 5)) {
          Node dead, live;
          bool have = false;
          for (const auto& n : yp_.nodes()) if (n.id() == id) dead = n;
          {
            Lock l(mu_);
            if (!unused_nodes_.empty()) { live = unused_nodes_.back(); unused_nodes_.pop_back(); have = true; }
          }
          if (!dead.id().empty() && have) {
            LL << "node " << id << " is overdue, replaced by " << live.id();
            dashboard_->forget(id);
            replaceNode(dead, live);
          }
        }
      }
    } else {
      Task tk;
      tk.set_type(TaskView::HEARTBEATING);
      tk.set_request(false);
      tk.set_msg(rep.SerializeAsString());
      MessagePtr m(new Message(tk));
      m->sender = my_id_;
      m->recver = opt_.scheduler.id();
      queue(m);
    }
  }
}

}  // namespace PS
