// hello world on the simple API (ref src/app/hello_world/main.cc:5-48): servers hold a
// KVVector "w" with keys 0..5 and values 0, .1, ... .5; every worker pulls four of them.
#include "../ps_main.inc"

namespace PS {
class Server : public App {
 public:
  Server() : App(FLAGS_app_name), vec_("w", FLAGS_app_name) {}
  ~Server() { stopExecutor(); }
  void init() override {
    // this server's part of keys {0..5}
    Range<Key> mine = myKeyRange();
    SArray<Key> key;
    SArray<float> val;
    for (Key k = 0; k < 6; ++k) {
      if (!mine.contains(k)) continue;
      key.pushBack(k);
      val.pushBack(0.1f * k);
    }
    vec_.key() = key;
    vec_.value() = val;
    LI << "this is server " << myRank() << " holding " << key.size() << " keys";
  }

 private:
  KVVector<Key, float> vec_;
};

App* CreateServerNode(const std::string& conf) { return new Server(); }
}  // namespace PS

int WorkerNodeMain(int argc, char* argv[]) {
  using namespace PS;
  KVVector<Key, float> vec("w", FLAGS_app_name);
  SArray<Key> key;
  if (MyRank() == 0) key = {0, 2, 4, 5};
  else key = {0, 1, 3, 4};
  vec.key() = key;
  MessagePtr msg(new Message(kServerGroup));
  msg->setKey(key);
  int t = vec.pull(msg);
  vec.waitOutMsg(kServerGroup, t);
  std::ostringstream os;
  os << MyNodeID() << ": pulled weight " << vec.value() << " at key " << vec.key();
  fprintf(stdout, "%s\n", os.str().c_str());
  fflush(stdout);
  return 0;
}
