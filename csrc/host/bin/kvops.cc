// A small program written only against ps.h that exercises the public KV contract end to end:
// timestamps returned by push / pull, wait_time dependencies between a worker's requests,
// waitOutMsg, per-key operators of KVMap (OR then PLUS), double values, and a
// KVBufferedVector whose pushes the server app collects with waitInMsg + received(t).
// Runs unchanged with -device cpu and -device gpu (shards in HBM, value frames through the
// PeerVan); tests/test_device_kv.py compares the two.
// ref: src/parameter/shared_parameter.h:17-40, kv_map.h:4-45, kv_buffered_vector.h:20-79
#include "../ps_main.inc"

namespace PS {

static const int kKeys = 1000;      // keys 0, 3, 6, ... spread over the servers' ranges
static Key keyAt(int i) { return (Key)i * (kMaxKey / kKeys); }

class OpsServer : public App {
 public:
  OpsServer() : App(FLAGS_app_name), vec_("vec", FLAGS_app_name), map_("map", FLAGS_app_name), buf_("buf", FLAGS_app_name) {}
  ~OpsServer() { stopExecutor(); }
  void init() override {
    SArray<Key> mine;
    for (int i = 0; i < kKeys; ++i)
      if (myKeyRange().contains(keyAt(i))) mine.pushBack(keyAt(i));
    vec_.key() = mine;
    vec_.value() = SArray<double>(mine.size(), 0.0);
    buf_.key() = mine;
  }
  // collect the buffered pushes of all workers at time 0 and publish their sum through vec_
  void run() override {
    buf_.waitInMsg(kWorkerGroup, 0);
    auto got = buf_.receivedOnHost(0);
    double sum = 0;
    for (double v : got.second[0]) sum += v;
    LI << "server " << myRank() << " buffered sum " << sum << " over " << got.first.size() << " keys"
       << (vec_.onDevice() ? " (HBM)" : " (host)");
    fprintf(stdout, "S%d: buffered %.1f\n", myRank(), sum);
    fflush(stdout);
  }

 private:
  KVVector<Key, double> vec_;
  KVMap<Key, int64> map_;
  KVBufferedVector<Key, double> buf_;
};

App* CreateServerNode(const std::string& conf) { return new OpsServer(); }
}  // namespace PS

int WorkerNodeMain(int argc, char* argv[]) {
  using namespace PS;
  const int me = MyRank(), workers = RankSize();
  KVVector<Key, double> vec("vec", FLAGS_app_name);
  KVMap<Key, int64> map("map", FLAGS_app_name);
  KVBufferedVector<Key, double> buf("buf", FLAGS_app_name);

  // ---- KVVector: two dependent pushes, the second must not be applied before the first ----
  SArray<Key> key;
  for (int i = me; i < kKeys; i += 2) key.pushBack(keyAt(i));  // rank 0: even, rank 1: odd, overlap for >2
  SArray<double> one(key.size(), me + 1.0), ten(key.size(), 10.0 * (me + 1));
  MessagePtr p1(new Message(kServerGroup));
  p1->setKey(key);
  p1->addValue(one);
  const int t1 = vec.push(p1);
  MessagePtr p2(new Message(kServerGroup, Message::kInvalidTime, t1));  // wait_time = t1
  p2->setKey(key);
  p2->addValue(vec.stage(ten));  // -device gpu: leaves from HBM through peer memory
  const int t2 = vec.push(p2);
  vec.waitOutMsg(kServerGroup, t2);
  vec.key() = key;
  MessagePtr q(new Message(kServerGroup, Message::kInvalidTime, t2));
  q->setKey(key);
  const int t3 = vec.pull(q);
  vec.waitOutMsg(kServerGroup, t3);
  bool vec_ok = vec.value().size() == key.size();
  // keys of my parity were pushed only by workers of my parity
  double expect = 0;
  for (int r = me % 2; r < workers; r += 2) expect += 11.0 * (r + 1);
  double mine = 11.0 * (me + 1);
  for (double v : vec.value()) vec_ok &= v >= mine - 1e-12 && v <= expect + 1e-12;

  // ---- KVMap: OR with my bit, then PLUS 16 that waits for the OR ----
  SArray<Key> mk;
  for (Key k : {Key(7), Key(1) << 40, kMaxKey - 5}) mk.pushBack(k);
  SArray<int64> bit(mk.size(), int64(1) << me), sixteen(mk.size(), 16);
  MessagePtr m1(new Message(kServerGroup));
  m1->setKey(mk);
  m1->addValue(bit);
  KVMap<Key, int64>::set(m1).set_op(OR);
  const int u1 = map.push(m1);
  MessagePtr m2(new Message(kServerGroup, Message::kInvalidTime, u1));
  m2->setKey(mk);
  m2->addValue(sixteen);
  KVMap<Key, int64>::set(m2).set_op(PLUS);
  const int u2 = map.push(m2);
  map.waitOutMsg(kServerGroup, u2);

  // ---- KVBufferedVector: one push at the agreed time 0, collected by the server app ----
  MessagePtr b(new Message(kServerGroup, 0));
  b->setKey(key);
  b->addValue(buf.stage(one));
  buf.waitOutMsg(kServerGroup, buf.push(b));

  // read the map back once every worker is through (poll: workers do not synchronise otherwise)
  const int64 want = ((int64(1) << workers) - 1) + 16 * workers;
  bool map_ok = false;
  for (int tries = 0; tries < 400 && !map_ok; ++tries) {
    MessagePtr g(new Message(kServerGroup));
    g->setKey(mk);
    map.waitOutMsg(kServerGroup, map.pull(g));
    auto snap = map.snapshot();
    map_ok = true;
    for (Key k : mk) map_ok &= snap[k] == want;
    if (!map_ok) usleep(20000);
  }
  fprintf(stdout, "%s: vec %s map %s\n", MyNodeID().c_str(), vec_ok ? "ok" : "WRONG", map_ok ? "ok" : "WRONG");
  fflush(stdout);
  return 0;
}
