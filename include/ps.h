// The "simple API" of the parameter server, source compatible with the reference's src/ps.h:
//
//   namespace PS { App* CreateServerNode(const std::string& conf); }   // user supplied
//   int WorkerNodeMain(int argc, char* argv[]);                         // user supplied
//   link with libpsb200_host + psmain (csrc/host/ps_main.cc provides main())
//
// and use KVVector / KVMap ... plus MyRank(), RankSize(), IsWorker() ... inside.
// ref src/ps.h:16-50, src/ps_main.cc:5-26, src/app/hello_world/main.cc
#pragma once
#include "parameter/kv.h"
#include "system/runtime.h"

// a worker's program: flags recognised by the runtime have been stripped from argv
int WorkerNodeMain(int argc, char* argv[]);

namespace PS {

// the app of a server node; `conf` is the string passed as -app_conf / read from -app_file
App* CreateServerNode(const std::string& conf);

inline App* MyApp() { return Postoffice::current().app(); }
inline Node MyNode() { return Postoffice::current().myNode(); }
inline std::string MyNodeID() { return MyNode().id(); }
inline int IsWorker() { return MyNode().role() == NodeView::WORKER; }
inline int IsServer() { return MyNode().role() == NodeView::SERVER; }
inline int IsScheduler() { return MyNode().role() == NodeView::SCHEDULER; }
inline Range<Key> MyKeyRange() { return MyNode().key().to<Key>(); }
inline std::string SchedulerID() { return Postoffice::current().scheduler().id(); }
inline int NextCustomerID() { static std::atomic<int> id{0}; return id++; }
// rank of this node among the nodes of its role, and the number of such nodes
inline int MyRank() { return MyNode().rank(); }
inline int RankSize() {
  YellowPages& yp = Postoffice::current().yp();
  return IsWorker() ? yp.num_workers() : (IsServer() ? yp.num_servers() : 1);
}

}  // namespace PS
