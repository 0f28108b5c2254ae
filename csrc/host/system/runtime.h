// The host control plane: Postoffice, YellowPages, Customer / App, Executor (per-customer
// dependency-gated task engine), RNode (remote node or group proxy), TaskTracker, Postmaster.
//
// Semantics follow the reference (ref src/system/{postoffice,yellow_pages,customer,app,
// executor,remote_node,task_tracker,postmaster}.{h,cc}):
//   * every customer exists under the same name on every node and owns one executor thread;
//   * a request carries a per-(customer, peer) timestamp and optional wait_time dependencies
//     that the RECEIVER enforces: it is not started before those earlier tasks from the same
//     sender have finished (bounded-delay consistency);
//   * a request to a group is sliced by the members' key ranges; pieces that intersect no
//     range are answered locally without touching the wire.
// Differences: Postoffice is an ordinary object (several nodes can live in one process over
// an InprocVan, which is how the unit tests run whole clusters), the van is pluggable, node
// ranks and shard order use natural node-id order (S2 before S10), REMOVE/REPLACE of nodes and
// heartbeats are implemented instead of being stubs.
#pragma once
#include <list>

#include "filter/filter.h"
#include "system/van.h"
#include "util/flags.h"

namespace PS {

DECLARE_int32(num_threads);
DECLARE_bool(verbose);

class Postoffice;
class Customer;
class Executor;
class App;

// node groups (ref src/system/executor.h:8-18)
static const NodeID kServerGroup = "all_servers";
static const NodeID kWorkerGroup = "all_workers";
static const NodeID kCompGroup = "all_comp_nodes";   // servers + workers
static const NodeID kReplicaGroup = "all_replicas";  // nodes holding a replica of my range
static const NodeID kOwnerGroup = "all_owners";      // nodes whose range I replicate
static const NodeID kLiveGroup = "all_lives";        // everything except unused nodes

typedef std::vector<Range<Key>> KeyRangeList;

// natural order of node ids: "S2" < "S10" (the reference sorts lexicographically,
// postmaster.cc + yellow_pages.h:49, which misorders shards beyond 10 servers)
struct NaturalLess {
  bool operator()(const string& a, const string& b) const;
};

// ---- TaskTracker (ref task_tracker.h:11-55) -------------------------------------------------------
class TaskTracker {
 public:
  void start(int t) { Lock l(mu_); task_[t] = false; }
  bool hasStarted(int t) { Lock l(mu_); return task_.count(t) > 0; }
  void finish(int t) {
    { Lock l(mu_); task_[t] = true; }
    cv_.notify_all();
  }
  bool hasFinished(int t) { Lock l(mu_); auto it = task_.find(t); return it != task_.end() && it->second; }
  bool tryWait(int t) { return hasFinished(t); }
  void wait(int t) {
    std::unique_lock<std::mutex> l(mu_);
    cv_.wait(l, [&] { auto it = task_.find(t); return it != task_.end() && it->second; });
  }
  bool waitFor(int t, int millis) {
    std::unique_lock<std::mutex> l(mu_);
    return cv_.wait_for(l, std::chrono::milliseconds(millis),
                        [&] { auto it = task_.find(t); return it != task_.end() && it->second; });
  }

 private:
  std::mutex mu_;
  std::condition_variable cv_;
  std::map<int, bool> task_;
};

// ---- RNode ---------------------------------------------------------------------------------------
class RNode {
 public:
  RNode(const NodeView& node, Executor& exec);
  ~RNode();
  const NodeID& id() const { return id_; }
  NodeView::Role role() const { return rnode_.role(); }
  Range<Key> keyRange() const { return rnode_.key().to<Key>(); }
  const Node& node() const { return rnode_; }
  void setNode(const NodeView& n) { rnode_ = n; }
  int size();  // members when this is a group, else 1
  int time() { Lock l(mu_); return time_; }

  // returns the timestamp used
  int submit(const MessagePtr& msg);
  int submit(MessagePtrList& msgs);
  int submit(const TaskView& task, const Message::Callback& recv_handle = Message::Callback());
  int submitAndWait(const TaskView& task, const Message::Callback& recv_handle = Message::Callback());
  int submit(const std::vector<Task>& tasks, const Message::Callback& recv_handle = Message::Callback());
  int submitAndWait(const std::vector<Task>& tasks, const Message::Callback& recv_handle = Message::Callback());

  void waitOutgoingTask(int time);
  bool tryWaitOutgoingTask(int time);
  void finishOutgoingTask(int time);
  void waitIncomingTask(int time);
  bool tryWaitIncomingTask(int time);
  void finishIncomingTask(int time);

  void encodeFilter(const MessagePtr& msg);
  void decodeFilter(const MessagePtr& msg);
  // messages sent to this node that were never answered (used when the node is replaced)
  MessagePtrList pendingMsgs();

 private:
  friend class Executor;
  Filter* findFilter(const FilterConfigView& conf);
  NodeID id_;
  Node rnode_;
  Executor& exec_;
  std::mutex mu_;
  int time_ = Message::kInvalidTime;
  TaskTracker incoming_task_, outgoing_task_;
  shared_ptr<RNode> replaces_;  // the dead node whose unanswered requests were re-sent to me
  shared_ptr<RNode> replaced_by_;  // set (with retired_, under mu_) when this node was swapped out
  bool retired_ = false;
  std::map<int, MessagePtr> pending_msgs_;
  std::map<int, Message::Callback> msg_receive_handle_, msg_finish_handle_;
  std::mutex filter_mu_;
  std::map<int, unique_ptr<Filter>> filter_;
};
typedef shared_ptr<RNode> RNodePtr;

// ---- Executor ------------------------------------------------------------------------------------
class Executor {
 public:
  explicit Executor(Customer& obj);
  ~Executor();
  void init(const std::vector<Node>& nodes);
  void stop();
  void add(const NodeView& node);
  void remove(const NodeView& node);
  void replace(const NodeView& dead, const NodeView& live);
  void updateKeyRange(const NodeView& node);

  RNodePtr rnode(const NodeID& k);
  std::vector<RNodePtr> group(const NodeID& k);
  KeyRangeList keyRanges(const NodeID& k);
  // a copy taken under the lock: rank and key range of this node change when the scheduler
  // assigns them (ManageNode ADD / UPDATE / REPLACE) while app threads are reading
  Node myNode() const { Lock l(node_mu_); return my_node_; }
  void setMyNode(const NodeView& n) { Lock l(node_mu_); my_node_ = n; }
  Customer& obj() { return obj_; }
  Postoffice& sys() { return sys_; }

  // called by the postoffice's receive thread
  void accept(const MessagePtr& msg);
  void notify() { { Lock l(recved_msg_mu_); } dag_cond_.notify_all(); }
  // payload (task.msg) of the reply being handled; valid inside recv/fin callbacks
  string lastRecvReply();
  MessagePtr activeMessage() { return active_msg_; }
  size_t queuedMessages() { Lock l(recved_msg_mu_); return recved_msgs_.size(); }

 private:
  void run();
  bool pickMessage();
  struct NodeInfo {
    RNodePtr node;
    std::vector<RNodePtr> sub_nodes;  // members sorted by key range
    KeyRangeList key_ranges;
    void addSubNode(const RNodePtr& s);
    void removeSubNode(const RNodePtr& s);
  };
  Customer& obj_;
  Postoffice& sys_;
  Node my_node_;
  mutable std::mutex node_mu_;
  std::map<NodeID, NodeInfo> nodes_;
  std::mutex recved_msg_mu_;
  std::condition_variable dag_cond_;
  std::list<MessagePtr> recved_msgs_;
  MessagePtr active_msg_;
  std::atomic<bool> done_{false};
  std::thread thread_;
};

// ---- Customer / App ------------------------------------------------------------------------------
class Customer {
 public:
  // registers under `name` with the postoffice of the calling thread (Postoffice::current())
  explicit Customer(const string& name);
  Customer(const string& name, const string& parent);
  virtual ~Customer();

  // handle a request (runs on this customer's executor thread)
  virtual void process(const MessagePtr& msg) {}
  // split a request for a group: default = one copy per member
  virtual MessagePtrList slice(const MessagePtr& msg, const KeyRangeList& krs);

  const string& name() const { return name_; }
  Executor& exec() { return *exec_; }
  Postoffice& sys() { return sys_; }
  RNodePtr port(const NodeID& k) { return exec_->rnode(k); }

  Node myNode() { return exec_->myNode(); }
  NodeID myNodeID() { return exec_->myNode().id(); }
  int myRank() { return exec_->myNode().rank(); }
  bool isWorker() { return exec_->myNode().role() == NodeView::WORKER; }
  bool isServer() { return exec_->myNode().role() == NodeView::SERVER; }
  bool isScheduler() { return exec_->myNode().role() == NodeView::SCHEDULER; }
  NodeID schedulerID();
  Range<Key> myKeyRange() { return exec_->myNode().key().to<Key>(); }
  // stops the executor thread; call from the most derived destructor when members are used
  // by process()
  void stopExecutor();

 protected:
  Postoffice& sys_;
  string name_;
  unique_ptr<Executor> exec_;
};

class App : public Customer {
 public:
  explicit App(const string& name);
  virtual ~App() {}
  virtual void init() {}
  virtual void run() {}
  // called after init() on a node that replaces a dead one: reload what the replicas hold
  virtual void recover() {}
  // default: forwards to the process-wide factory (see setFactory)
  static App* create(const string& name, const string& conf);
  typedef std::function<App*(const string& name, const string& conf)> Factory;
  static void setFactory(const Factory& f);
};

// ---- Engine (ref system.h:4-17): a named customer with no behaviour, kept as an extension point ----
class Engine : public Customer {
 public:
  explicit Engine(const string& name) : Customer(name) {}
  virtual ~Engine() {}
};

// ---- YellowPages (ref yellow_pages.{h,cc}) -----------------------------------------------------------
class YellowPages {
 public:
  void add(Customer* obj);
  void remove(const string& name);
  Customer* customer(const string& name);           // waits a little for one under construction
  Customer* customerIfPresent(const string& name);  // never waits
  void depends(const string& child, const string& parent);
  std::vector<string> childern(const string& parent);
  void addNode(const NodeView& node);
  void removeNode(const NodeID& id);
  std::vector<Node> nodes();
  bool hasNode(const NodeID& id);
  int num_workers() { Lock l(mu_); return num_workers_; }
  int num_servers() { Lock l(mu_); return num_servers_; }
  Van& van() { return *van_; }
  void setVan(unique_ptr<Van> v) { van_ = std::move(v); }

 private:
  std::mutex mu_;
  std::condition_variable cv_;
  std::map<string, Customer*> customers_;
  std::map<string, std::vector<string>> relations_;
  std::map<NodeID, Node, NaturalLess> nodes_;
  int num_workers_ = 0, num_servers_ = 0;
  unique_ptr<Van> van_;
  friend class Postoffice;
};

// ---- Postmaster (ref postmaster.{h,cc}) -------------------------------------------------------------
class Postmaster {
 public:
  // files matching `conf` divided round-robin over the workers
  static std::vector<DataConfig> partitionData(const DataConfigView& conf, int num_workers);
  // server s of S gets range.evenDivide(S, s); every other node gets the whole range
  static std::vector<Node> partitionServerKeyRange(const std::vector<Node>& nodes, Range<Key> range);
  static std::vector<Node> assignNodeRank(const std::vector<Node>& nodes);
};

// ---- Postoffice --------------------------------------------------------------------------------------
struct PostofficeOptions {
  Node my_node, scheduler;
  int num_servers = 0, num_workers = 0, num_unused = 0;
  string app_name = "app", app_conf, app_file;
  bool traffic_statistics = false, print_van = false, message_compression = true;
  int report_interval = 0;            // seconds between heartbeats, 0 = off
  shared_ptr<InprocFabric> fabric;    // non-null: in-process van instead of TCP
  App::Factory factory;               // overrides the process-wide App factory
};

class HeartbeatInfo;
class Dashboard;

class Postoffice {
 public:
  Postoffice();
  ~Postoffice();
  // process-wide default instance (what the binaries use)
  static Postoffice& instance();
  // the postoffice customers created on this thread attach to
  static Postoffice& current();
  static void setCurrent(Postoffice* p);

  // gflags front end: parses -my_node, -scheduler, -num_servers, ... then start(options)
  void start(int* argc, char*** argv);
  void start(const PostofficeOptions& opt);
  // blocks until the scheduler terminated the job (or, on the scheduler, all nodes are done)
  void stop();

  void queue(const MessagePtr& msg);
  void reply(const NodeID& recver, const TaskView& task, const string& reply_msg = "");
  template <class Msg>
  void replyProtocalMessage(const MessagePtr& msg, const Msg& proto) {
    reply(msg->sender, msg->task, proto.SerializeAsString());
    msg->replied = true;
  }
  void reply(const MessagePtr& msg, const string& reply_msg = "") {
    reply(msg->sender, msg->task, reply_msg);
    msg->replied = true;
  }

  YellowPages& yp() { return yp_; }
  Node myNode() { Lock l(my_node_mu_); return opt_.my_node; }
  const string& myID() const { return my_id_; }  // never changes after start()
  Node& scheduler() { return opt_.scheduler; }
  const PostofficeOptions& options() const { return opt_; }
  App* app() { return app_.get(); }
  bool messageCompression() const { return opt_.message_compression; }
  HeartbeatInfo& heartbeatInfo() { return *hb_info_; }
  // scheduler side: replace a dead node by an unused one / drop it (ManageNode::REPLACE/REMOVE)
  void replaceNode(const NodeView& dead, const NodeView& live);
  void removeNode(const NodeView& dead);
  std::vector<string> deadNodes(int timeout_sec);
  // the scheduler's table of the latest heartbeat reports (ref dashboard.cc)
  string dashboardReport();
  // what a crash looks like to this process: stop waiting, shut the van (fault-injection in tests)
  void kill();

 private:
  void send();
  void recv();
  void heartbeat();
  void manageApp(const TaskView& tk);
  void manageNode(const TaskView& tk);
  void addMyNode(const string& name, const NodeView& recver);
  void finishBarrier(int n);

  PostofficeOptions opt_;
  std::mutex my_node_mu_;            // guards opt_.my_node after start()
  string my_id_;                     // = opt_.my_node.id(), immutable
  std::atomic<int> my_role_{0};      // changes once at most (a spare taking over a role)
  YellowPages yp_;
  unique_ptr<App> app_;
  threadsafe_queue<MessagePtr> sending_queue_;
  std::thread recv_thread_, send_thread_, hb_thread_;
  std::atomic<bool> done_{false}, started_{false};
  std::promise<void> nodes_are_ready_, init_app_promise_, run_app_promise_;
  std::mutex mu_;
  std::condition_variable cv_;
  int num_done_ = 0, num_connected_workers_ = 0, num_connected_servers_ = 0;
  std::vector<Node> unused_nodes_;
  MessagePtr init_msg_, run_msg_;
  unique_ptr<HeartbeatInfo> hb_info_;
  unique_ptr<Dashboard> dashboard_;
};

}  // namespace PS
