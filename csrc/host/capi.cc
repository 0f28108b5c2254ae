// C ABI of the host runtime for the Python package (ctypes): utilities, parsers, filters and
// whole in-process clusters.  Everything returns 0 on success; psh_last_error() has the text.
#include <sstream>

#include "app/linear_method/async_sgd.h"
#include "app/linear_method/darlin.h"
#include "app/linear_method/linear.h"
#include "app/linear_method/model_evaluation.h"
#include "data/slot_reader.h"
#include "data/text_parser.h"
#include "filter/filter.h"
#include "parameter/kv.h"
#include "system/heartbeat.h"
#include "system/monitor.h"
#include "system/runtime.h"
#include "util/hash.h"
#include "util/localizer.h"
#include "util/sketch.h"
#include "util/snappy.h"

#define PSH_API extern "C" __attribute__((visibility("default")))

using namespace PS;

static thread_local std::string g_err;
static int fail(const std::string& m) {
  g_err = m;
  return 1;
}
static int put(const std::string& s, char* out, int64_t cap, int64_t* len) {
  if (len) *len = (int64_t)s.size();
  if ((int64_t)s.size() > cap) return fail("output buffer too small");
  memcpy(out, s.data(), s.size());
  return 0;
}

PSH_API const char* psh_last_error() { return g_err.c_str(); }
PSH_API int psh_abi_version() { return 1; }

// ---- hashing / compression ------------------------------------------------------------------------
PSH_API uint32_t psh_crc32c(const char* data, int64_t n) { return crc32c::Value(data, (size_t)n); }
PSH_API uint32_t psh_crc32c_extend(uint32_t crc, const char* data, int64_t n) { return crc32c::Extend(crc, data, (size_t)n); }
PSH_API uint32_t psh_murmur3_x86_32(const char* data, int32_t n, uint32_t seed) {
  uint32_t out;
  MurmurHash3_x86_32(data, n, seed, &out);
  return out;
}
PSH_API void psh_murmur3_x64_128(const char* data, int32_t n, uint32_t seed, uint64_t* out2) {
  MurmurHash3_x64_128(data, n, seed, out2);
}
PSH_API uint32_t psh_sketch_hash(uint64_t key) { return Sketch::hash(key); }
PSH_API int64_t psh_snappy_max_compressed(int64_t n) { return (int64_t)snappy::MaxCompressedLength((size_t)n); }
PSH_API int psh_snappy_compress(const char* src, int64_t n, char* dst, int64_t* out_len) {
  size_t m = 0;
  snappy::RawCompress(src, (size_t)n, dst, &m);
  *out_len = (int64_t)m;
  return 0;
}
PSH_API int psh_snappy_uncompressed_length(const char* src, int64_t n, int64_t* out_len) {
  size_t m = 0;
  if (!snappy::GetUncompressedLength(src, (size_t)n, &m)) return fail("bad snappy header");
  *out_len = (int64_t)m;
  return 0;
}
PSH_API int psh_snappy_uncompress(const char* src, int64_t n, char* dst) {
  return snappy::RawUncompress(src, (size_t)n, dst) ? 0 : fail("corrupt snappy stream");
}

// ---- ranges ---------------------------------------------------------------------------------------
PSH_API void psh_range_even_divide(uint64_t begin, uint64_t end, uint64_t n, uint64_t i, uint64_t* out2) {
  Range<uint64> r = Range<uint64>(begin, end).evenDivide(n, i);
  out2[0] = r.begin();
  out2[1] = r.end();
}

// ---- text format ----------------------------------------------------------------------------------
// parse and re-print (single_line=0 -> DebugString); used to check config compatibility
PSH_API int psh_textproto_normalize(const char* text, int single_line, char* out, int64_t cap, int64_t* len) {
  PbMsg m;
  std::string err;
  if (!m.parse(text, &err)) return fail(err);
  return put(m.toString(single_line != 0), out, cap, len);
}
// value of a dotted path ("linear_method.darlin.max_pass_of_data", "a.b[2].c"); empty if absent
PSH_API int psh_textproto_get(const char* text, const char* path, char* out, int64_t cap, int64_t* len) {
  PbMsg m;
  std::string err;
  if (!m.parse(text, &err)) return fail(err);
  const PbMsg* cur = &m;
  std::string p(path);
  size_t b = 0;
  std::string result;
  while (cur) {
    // components are separated by '.', but '[...]' extension names contain dots themselves
    size_t e = b;
    int depth = 0;
    while (e < p.size() && (p[e] != '.' || depth > 0)) {
      if (p[e] == '[' && e == b) depth = 1;
      if (p[e] == ']') depth = 0;
      ++e;
    }
    std::string comp = p.substr(b, e - b);
    size_t idx = 0;
    if (!comp.empty() && comp.back() == ']' && comp[0] != '[') {
      size_t lb = comp.rfind('[');
      idx = (size_t)atoi(comp.c_str() + lb + 1);
      comp = comp.substr(0, lb);
    }
    const PbMsg::Field* f = cur->find(comp, idx);
    if (!f) break;
    if (e >= p.size()) {
      result = f->is_msg ? f->msg->shortDebugString() : f->scalar;
      break;
    }
    cur = f->is_msg ? f->msg.get() : nullptr;
    b = e + 1;
  }
  return put(result, out, cap, len);
}

// ---- parsers --------------------------------------------------------------------------------------
// Parse `n_lines` newline separated lines; returns a flat description:
//   per example: n_slots, then per slot: id, n_keys, n_vals ; keys and vals appended to the arrays
PSH_API int psh_parse_text(const char* text, int32_t format, int32_t ignore_group, int64_t* meta,
                           int64_t meta_cap, uint64_t* keys, int64_t key_cap, float* vals,
                           int64_t val_cap, int64_t* counts4) {
  ExampleParser parser;
  parser.init((DataConfigView::TextFormat)format, ignore_group != 0);
  std::istringstream in(text);
  std::string line;
  int64_t nm = 0, nk = 0, nv = 0, nex = 0, bad = 0;
  Example ex;
  while (std::getline(in, line)) {
    if (line.empty()) continue;
    if (!parser.toProto(line.c_str(), &ex)) {
      ++bad;
      continue;
    }
    if (nm + 1 + 3 * (int64_t)ex.slot.size() > meta_cap) return fail("meta buffer too small");
    meta[nm++] = ex.slot_size();
    for (const Slot& s : ex.slot) {
      meta[nm++] = s.id;
      meta[nm++] = s.key_size();
      meta[nm++] = s.val_size();
      if (nk + s.key_size() > key_cap || nv + s.val_size() > val_cap) return fail("data buffer too small");
      for (uint64 k : s.key) keys[nk++] = k;
      for (float v : s.val) vals[nv++] = v;
    }
    ++nex;
  }
  counts4[0] = nex; counts4[1] = bad; counts4[2] = nk; counts4[3] = nv;
  return 0;
}

// Example binary codec round trip (RecordIO payload compatibility)
PSH_API int psh_example_roundtrip(const char* line, int32_t format, char* out, int64_t cap, int64_t* len) {
  ExampleParser parser;
  parser.init((DataConfigView::TextFormat)format, false);
  Example ex, back;
  if (!parser.toProto(line, &ex)) return fail("parse error");
  std::string bin = ex.SerializeAsString();
  if (!back.ParseFromString(bin)) return fail("binary decode failed");
  if (back.SerializeAsString() != bin) return fail("binary round trip mismatch");
  return put(bin, out, cap, len);
}

// Load text files through SlotReader (columnar cache) and return the slot arrays of one group.
// info_out receives the merged ExampleInfo as text.
PSH_API int psh_slot_reader_load(const char* data_conf_text, const char* cache_prefix, char* info_out,
                                 int64_t cap, int64_t* len) {
  DataConfig data, cache;
  std::string err;
  if (!data.raw()->parse(data_conf_text, &err)) return fail(err);
  cache.add_file(cache_prefix);
  SlotReader reader(data, cache);
  ExampleInfo info;
  reader.read(&info);
  return put(info.DebugString(), info_out, cap, len);
}
PSH_API int psh_slot_reader_arrays(const char* data_conf_text, const char* cache_prefix, int32_t slot,
                                   uint64_t* index, int64_t index_cap, int64_t* offset,
                                   int64_t offset_cap, float* value, int64_t value_cap,
                                   int64_t* counts3) {
  DataConfig data, cache;
  std::string err;
  if (!data.raw()->parse(data_conf_text, &err)) return fail(err);
  cache.add_file(cache_prefix);
  SlotReader reader(data, cache);
  reader.read(nullptr);
  SArray<uint64> idx = reader.index(slot);
  SArray<size_t> off = reader.offset(slot);
  SArray<float> val = reader.value<float>(slot);
  counts3[0] = idx.size(); counts3[1] = off.size(); counts3[2] = val.size();
  if ((int64_t)idx.size() > index_cap || (int64_t)off.size() > offset_cap || (int64_t)val.size() > value_cap)
    return fail("buffers too small");
  for (size_t i = 0; i < idx.size(); ++i) index[i] = idx[i];
  for (size_t i = 0; i < off.size(); ++i) offset[i] = (int64_t)off[i];
  for (size_t i = 0; i < val.size(); ++i) value[i] = val[i];
  return 0;
}

// ---- sketches ---------------------------------------------------------------------------------------
PSH_API int psh_countmin_filter(const uint64_t* keys, const uint8_t* counts, int64_t n, int32_t cm_n,
                                int32_t cm_k, int32_t freq, uint8_t* keep_out, uint8_t* table_out) {
  CountMin<uint64, uint8> cm;
  cm.resize(cm_n, cm_k, 254);
  for (int64_t i = 0; i < n; ++i) cm.insert(keys[i], counts[i]);
  for (int64_t i = 0; i < n; ++i) keep_out[i] = (int)cm.query(keys[i]) > freq;
  if (table_out) memcpy(table_out, cm.data().data(), cm.n());
  return 0;
}
PSH_API int psh_bloom_fpr(int64_t n_insert, int64_t n_probe, int64_t bits, int32_t k, int32_t blocked,
                          double* fpr_out) {
  BloomFilter<uint64> bf((int)bits, k);
  BlockBloomFilter<uint64> bbf((size_t)bits, k);
  for (int64_t i = 0; i < n_insert; ++i) {
    uint64 key = fmix64((uint64)i + 1);
    if (blocked) bbf.insert(key); else bf.insert(key);
  }
  for (int64_t i = 0; i < n_insert; ++i) {
    uint64 key = fmix64((uint64)i + 1);
    if (!(blocked ? bbf.query(key) : bf.query(key))) return fail("false negative");
  }
  int64_t fp = 0;
  for (int64_t i = 0; i < n_probe; ++i) {
    uint64 key = fmix64((uint64)(i + n_insert + 12345));
    fp += blocked ? bbf.query(key) : bf.query(key);
  }
  *fpr_out = (double)fp / (double)n_probe;
  return 0;
}

// ---- fixed point codec ------------------------------------------------------------------------------
PSH_API int psh_fixing_float_roundtrip(const double* in, int64_t n, int32_t nbytes, double min_v,
                                       double max_v, uint64_t seed, double* out, uint8_t* code_out) {
  if (nbytes < 1 || nbytes > 7) return fail("nbytes must be in [1,7]");
  SArray<double> orig(n);
  for (int64_t i = 0; i < n; ++i) orig[i] = in[i];
  uint64 s = seed;
  SArray<char> code = FixingFloatFilter::encodeArray<double>(orig, nbytes, min_v, max_v, &s);
  if (code_out) memcpy(code_out, code.data(), code.size());
  SArray<double> back(FixingFloatFilter::decodeArray<double>(code, nbytes, min_v, max_v));
  for (int64_t i = 0; i < n; ++i) out[i] = back[i];
  return 0;
}

// ---- evaluation --------------------------------------------------------------------------------------
PSH_API int psh_auc_accuracy(const double* label, const double* predict, int64_t n, double* out3) {
  SArray<double> y(n), p(n);
  for (int64_t i = 0; i < n; ++i) { y[i] = label[i]; p[i] = predict[i]; }
  out3[0] = Evaluation<double>::auc(y, p);
  out3[1] = Evaluation<double>::accuracy(y, p);
  AUC bucket;
  bucket.setGoodness(100000);
  AUCData d;
  bucket.compute(y, p, d);
  AUC merged;
  merged.setGoodness(100000);
  merged.merge(d);
  out3[2] = merged.evaluate();
  return 0;
}

// ---- whole clusters inside this process ----------------------------------------------------------------
// Runs 1 scheduler + S servers + W workers as threads over the in-process van (use_tcp=0) or
// over TCP on localhost ports base_port.. (use_tcp=1) with the given AppConfig text, and
// returns a report: "iter objective relative nnz_w violation nnz_active\n" per Darlin pass,
// or "loss auc" for async SGD.
namespace {
struct ClusterResult {
  std::mutex mu;
  std::string report;
};
}  // namespace

PSH_API int psh_run_cluster(const char* app_conf_text, int32_t num_servers, int32_t num_workers,
                            int32_t use_tcp, int32_t base_port, int32_t message_compression,
                            char* report_out, int64_t cap, int64_t* len) {
  std::string conf(app_conf_text);
  {
    AppConfig test;
    std::string err;
    if (!test.raw()->parse(conf, &err)) return fail("bad app config: " + err);
  }
  auto fabric = use_tcp ? shared_ptr<InprocFabric>() : std::make_shared<InprocFabric>();
  ClusterResult result;
  auto make_node = [&](const char* role, const std::string& id, int port) {
    Node n;
    NodeView::Role r;
    NodeView::Role_Parse(role, &r);
    n.set_role(r);
    n.set_id(id);
    n.set_hostname("127.0.0.1");
    n.set_port(port);
    return n;
  };
  Node sched = make_node("SCHEDULER", "H", base_port);
  std::vector<Node> nodes;
  nodes.push_back(sched);
  for (int i = 0; i < num_servers; ++i) nodes.push_back(make_node("SERVER", "S" + to_string(i), base_port + 1 + i));
  for (int i = 0; i < num_workers; ++i)
    nodes.push_back(make_node("WORKER", "W" + to_string(i), base_port + 1 + num_servers + i));

  std::vector<std::thread> threads;
  std::vector<unique_ptr<Postoffice>> offices(nodes.size());
  for (size_t i = 0; i < nodes.size(); ++i) offices[i].reset(new Postoffice());
  for (size_t i = 0; i < nodes.size(); ++i) {
    threads.emplace_back([&, i] {
      PostofficeOptions opt;
      opt.my_node = nodes[i];
      opt.scheduler = sched;
      opt.num_servers = num_servers;
      opt.num_workers = num_workers;
      opt.app_conf = conf;
      opt.fabric = fabric;
      opt.message_compression = message_compression != 0;
      opt.factory = LM::createAppFromText;
      Postoffice& po = *offices[i];
      po.start(opt);
      if (i == 0) {
        std::ostringstream os;
        if (auto* d = dynamic_cast<LM::DarlinScheduler*>(po.app())) {
          for (const auto& kv : d->progress()) {
            const BCDProgress& p = kv.second;
            os << kv.first << " " << PbMsg::fmtDouble(p.objective()) << " " << PbMsg::fmtDouble(p.relative_obj())
               << " " << p.nnz_w() << " " << PbMsg::fmtDouble(p.violation()) << " " << p.nnz_active_set() << "\n";
          }
        } else if (auto* s = dynamic_cast<LM::AsyncSGDScheduler*>(po.app())) {
          os << PbMsg::fmtDouble(s->lastLoss()) << " " << PbMsg::fmtDouble(s->lastAUC()) << "\n";
        } else if (auto* e = dynamic_cast<LM::ModelEvaluation*>(po.app())) {
          os << PbMsg::fmtDouble(e->auc()) << " " << PbMsg::fmtDouble(e->accuracy()) << " " << e->numExamples() << "\n";
        }
        Lock l(result.mu);
        result.report = os.str();
      }
      po.stop();
    });
  }
  for (auto& t : threads) t.join();
  offices.clear();
  return put(result.report, report_out, cap, len);
}

// hello-world style KV check on an in-process cluster: servers hold keys 0..n-1 with value
// 0.1*key in a KVVector, each worker pulls `pull_keys`; returns the pulled values of worker 0
namespace {
class KVServerApp : public App {
 public:
  KVServerApp(const string& name, int n) : App(name), vec_("w", name), n_(n) {}
  ~KVServerApp() { stopExecutor(); }
  void init() override {
    Range<Key> mine = myKeyRange();
    SArray<Key> key;
    SArray<float> val;
    for (Key k = 0; k < (Key)n_; ++k)
      if (mine.contains(k)) { key.pushBack(k); val.pushBack(0.1f * k); }
    vec_.key() = key;
    vec_.value() = val;
  }
  KVVector<Key, float> vec_;
  int n_;
};
class KVWorkerApp : public App {
 public:
  KVWorkerApp(const string& name, std::vector<uint64> keys, std::vector<float>* out, std::mutex* mu)
      : App(name), vec_("w", name), keys_(keys), out_(out), mu_(mu) {}
  ~KVWorkerApp() { stopExecutor(); }
  void run() override {
    SArray<Key> key;
    for (uint64 k : keys_) key.pushBack(k);
    vec_.key() = key;
    // push +1 to every key, then pull
    MessagePtr push(new Message(kServerGroup));
    push->setKey(key);
    SArray<float> ones(key.size(), 1.0f);
    push->addValue(ones);
    push->addFilter(FilterConfigView::KEY_CACHING);
    int tp = vec_.push(push);
    vec_.waitOutMsg(kServerGroup, tp);
    // workers do not synchronise with each other: pull until every worker's +1 has landed
    const float all = (float)sys_.yp().num_workers();
    for (int tries = 0; tries < 1000; ++tries) {
      vec_.value().clear();  // a pull reply is added into the local values
      MessagePtr msg(new Message(kServerGroup));
      msg->setKey(key);
      msg->addFilter(FilterConfigView::KEY_CACHING);
      int t = vec_.pull(msg);
      vec_.waitOutMsg(kServerGroup, t);
      SArray<float> v = vec_.value();
      bool complete = v.size() == key.size();
      for (size_t i = 0; complete && i < v.size(); ++i) complete = v[i] >= 0.1f * key[i] + all - 1e-3f;
      if (complete) break;
      usleep(2000);
    }
    if (myRank() == 0) {
      Lock l(*mu_);
      SArray<float> v = vec_.value();
      out_->assign(v.begin(), v.end());
    }
  }
  KVVector<Key, float> vec_;
  std::vector<uint64> keys_;
  std::vector<float>* out_;
  std::mutex* mu_;
};
}  // namespace

PSH_API int psh_run_kv_cluster(int32_t num_servers, int32_t num_workers, int32_t n_keys,
                               const uint64_t* pull_keys, int32_t n_pull, int32_t use_tcp,
                               int32_t base_port, float* out_vals) {
  auto fabric = use_tcp ? shared_ptr<InprocFabric>() : std::make_shared<InprocFabric>();
  std::vector<uint64> keys(pull_keys, pull_keys + n_pull);
  std::vector<float> got;
  std::mutex mu;
  auto factory = [&](const string& name, const string&) -> App* {
    Node me = Postoffice::current().myNode();
    if (me.role() == NodeView::SERVER) return new KVServerApp(name, n_keys);
    if (me.role() == NodeView::WORKER) return new KVWorkerApp(name, keys, &got, &mu);
    return new App(name);
  };
  std::vector<Node> nodes;
  auto add = [&](NodeView::Role r, const std::string& id, int port) {
    Node n;
    n.set_role(r);
    n.set_id(id);
    n.set_hostname("127.0.0.1");
    n.set_port(port);
    nodes.push_back(n);
  };
  add(NodeView::SCHEDULER, "H", base_port);
  for (int i = 0; i < num_servers; ++i) add(NodeView::SERVER, "S" + to_string(i), base_port + 1 + i);
  for (int i = 0; i < num_workers; ++i) add(NodeView::WORKER, "W" + to_string(i), base_port + 1 + num_servers + i);
  std::vector<unique_ptr<Postoffice>> offices(nodes.size());
  for (auto& o : offices) o.reset(new Postoffice());
  std::vector<std::thread> threads;
  for (size_t i = 0; i < nodes.size(); ++i) {
    threads.emplace_back([&, i] {
      PostofficeOptions opt;
      opt.my_node = nodes[i];
      opt.scheduler = nodes[0];
      opt.num_servers = num_servers;
      opt.num_workers = num_workers;
      opt.fabric = fabric;
      opt.factory = factory;
      offices[i]->start(opt);
      offices[i]->stop();
    });
  }
  for (auto& t : threads) t.join();
  offices.clear();
  if ((int)got.size() != n_pull) return fail("worker 0 pulled " + to_string(got.size()) + " values");
  for (int i = 0; i < n_pull; ++i) out_vals[i] = got[i];
  return 0;
}

// ---- fault-tolerance scenarios (in-process clusters) ---------------------------------------------------
namespace {
struct Gate {
  std::mutex mu;
  std::condition_variable cv;
  int stage = 0;
  void set(int s) { { Lock l(mu); stage = s; } cv.notify_all(); }
  void wait(int s) { std::unique_lock<std::mutex> l(mu); cv.wait(l, [&] { return stage >= s; }); }
};
// a server whose shard can be reloaded from the replica holders
static Key spreadKey(int i, int n) { return (Key)i * (~(Key)0 / (Key)n); }  // covers every shard
class ReplServerApp : public KVServerApp {
 public:
  using KVServerApp::KVServerApp;
  void init() override {
    Range<Key> mine = myKeyRange();
    SArray<Key> key;
    SArray<float> val;
    for (int i = 0; i < n_; ++i)
      if (mine.contains(spreadKey(i, n_))) { key.pushBack(spreadKey(i, n_)); val.pushBack(0.1f * i); }
    vec_.key() = key;
    vec_.value() = val;
  }
  void recover() override { vec_.recover(myKeyRange()); }
};
class ReplWorkerApp : public App {
 public:
  ReplWorkerApp(const string& name, int n, std::vector<float>* out, Gate* gate)
      : App(name), vec_("w", name), n_(n), out_(out), gate_(gate) {}
  ~ReplWorkerApp() { stopExecutor(); }
  void run() override {
    SArray<Key> key;
    for (int k = 0; k < n_; ++k) key.pushBack(spreadKey(k, n_));
    vec_.key() = key;
    MessagePtr push(new Message(kServerGroup));
    push->setKey(key);
    push->addValue(SArray<float>(key.size(), 1.0f));
    vec_.waitOutMsg(kServerGroup, vec_.push(push));
    gate_->set(1);   // pushed and acknowledged: the driver may now kill a server
    gate_->wait(2);  // a spare took the dead server's place
    MessagePtr pull(new Message(kServerGroup));
    pull->setKey(key);
    vec_.waitOutMsg(kServerGroup, vec_.pull(pull));
    SArray<float> v = vec_.value();
    out_->assign(v.begin(), v.end());
  }
  KVVector<Key, float> vec_;
  int n_;
  std::vector<float>* out_;
  Gate* gate_;
};
class SleepWorkerApp : public App {
 public:
  SleepWorkerApp(const string& name, int ms) : App(name), ms_(ms) {}
  void run() override { std::this_thread::sleep_for(std::chrono::milliseconds(ms_)); }
  int ms_;
};

struct MiniCluster {
  std::vector<Node> nodes;
  std::vector<unique_ptr<Postoffice>> offices;
  std::vector<std::thread> threads;
  std::vector<std::unique_ptr<std::atomic<int>>> state;  // 0 starting, 1 started, 2 stopped
  shared_ptr<InprocFabric> fabric = std::make_shared<InprocFabric>();
  void add(NodeView::Role r, const string& id) {
    Node n;
    n.set_role(r);
    n.set_id(id);
    n.set_hostname("127.0.0.1");
    n.set_port(1000 + (int)nodes.size());
    nodes.push_back(n);
  }
  int index(const string& id) const {
    for (size_t i = 0; i < nodes.size(); ++i) if (nodes[i].id() == id) return (int)i;
    return -1;
  }
  void launch(int num_servers, int num_workers, int num_unused, int report_interval, App::Factory factory) {
    offices.resize(nodes.size());
    for (auto& o : offices) o.reset(new Postoffice());
    for (size_t i = 0; i < nodes.size(); ++i) state.emplace_back(new std::atomic<int>(0));
    for (size_t i = 0; i < nodes.size(); ++i) {
      threads.emplace_back([=] {
        PostofficeOptions opt;
        opt.my_node = nodes[i];
        opt.scheduler = nodes[0];
        opt.num_servers = num_servers;
        opt.num_workers = num_workers;
        opt.num_unused = num_unused;
        opt.report_interval = report_interval;
        opt.fabric = fabric;
        opt.factory = factory;
        opt.print_van = getenv("PSH_PRINT_VAN") != nullptr;
        offices[i]->start(opt);
        state[i]->store(1);
        offices[i]->stop();
        state[i]->store(2);
      });
    }
  }
  // empty string: every node stopped; otherwise the nodes that did not within `seconds` (their
  // threads and postoffices are leaked on purpose so the caller can report instead of hanging)
  string join(int seconds = 60) {
    string stuck;
    for (int tick = 0; tick < seconds * 100; ++tick) {
      bool all = true;
      for (auto& s : state) all = all && s->load() == 2;
      if (all) break;
      std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    for (size_t i = 0; i < nodes.size(); ++i)
      if (state[i]->load() != 2) stuck += nodes[i].id() + (state[i]->load() == 0 ? "(in start) " : "(in stop) ");
    if (stuck.empty()) {
      for (auto& t : threads) t.join();
      offices.clear();
    } else {
      for (auto& t : threads) t.detach();
      for (auto& o : offices) o.release();
    }
    return stuck;
  }
};
}  // namespace

// Servers S0,S1 hold keys 0..n-1 (value 0.1*key) with one replica each, a spare U0 idles.  The
// worker adds 1 to every key; then S1 is killed, the scheduler swaps U0 in (ManageNode::REPLACE),
// U0 reloads S1's shard from the replica holder, and the worker pulls all keys.
// out_vals[k]: 0.1*k + 1 on the surviving shard, 1 (= what the replica saw pushed) on the recovered one.
PSH_API int psh_run_replace_scenario(int32_t n_keys, float* out_vals, uint64_t* s1_range2) {
  const int saved = FLAGS_num_replicas;
  FLAGS_num_replicas = 1;
  std::vector<float> got;
  Gate gate;
  MiniCluster c;
  c.add(NodeView::SCHEDULER, "H");
  c.add(NodeView::SERVER, "S0");
  c.add(NodeView::SERVER, "S1");
  c.add(NodeView::WORKER, "W0");
  c.add(NodeView::UNUSED, "U0");
  auto factory = [&](const string& name, const string&) -> App* {
    Node me = Postoffice::current().myNode();
    if (me.role() == NodeView::SERVER) return new ReplServerApp(name, n_keys);
    if (me.role() == NodeView::WORKER) return new ReplWorkerApp(name, n_keys, &got, &gate);
    return new App(name);
  };
  c.launch(2, 1, 1, 0, factory);
  gate.wait(1);
  Postoffice& sched = *c.offices[0];
  Node dead, live;
  for (int tries = 0; tries < 500 && live.id().empty(); ++tries) {  // the spare may connect late
    for (const auto& n : sched.yp().nodes()) {
      if (n.id() == "S1") dead = n;
      if (n.id() == "U0") live = n;
    }
    if (live.id().empty()) std::this_thread::sleep_for(std::chrono::milliseconds(10));
  }
  if (live.id().empty() || dead.id().empty()) return fail("spare or victim never registered");
  // which of the n spread keys live on S1
  if (s1_range2) {
    Range<Key> r(dead.key().begin(), dead.key().end());
    int lo = n_keys, hi = 0;
    for (int i = 0; i < n_keys; ++i)
      if (r.contains(spreadKey(i, n_keys))) { lo = std::min(lo, i); hi = std::max(hi, i + 1); }
    s1_range2[0] = lo; s1_range2[1] = hi;
  }
  // replication is asynchronous (a push is acknowledged before its copy reaches the replica
  // holder): give the forwarded pushes time to land before the crash
  std::this_thread::sleep_for(std::chrono::milliseconds(300));
  c.offices[c.index("S1")]->kill();
  sched.replaceNode(dead, live);
  gate.set(2);
  string stuck = c.join(30);
  FLAGS_num_replicas = saved;
  if (!stuck.empty()) return fail("nodes did not stop: " + stuck);
  if ((int)got.size() != n_keys) return fail("worker pulled " + to_string(got.size()) + " values");
  for (int i = 0; i < n_keys; ++i) out_vals[i] = got[i];
  return 0;
}

// Every node reports to the scheduler each `interval_s` seconds; returns the dashboard table
// after `run_ms` and, in *n_dead, how many nodes look dead with a zero-second timeout margin
// `dead_timeout_s` (a node that never reported is not counted).
PSH_API int psh_run_heartbeat_scenario(int32_t interval_s, int32_t run_ms, int32_t dead_timeout_s,
                                       char* out, int64_t cap, int64_t* len, int32_t* n_dead) {
  MiniCluster c;
  c.add(NodeView::SCHEDULER, "H");
  c.add(NodeView::SERVER, "S0");
  c.add(NodeView::WORKER, "W0");
  c.add(NodeView::WORKER, "W1");
  auto factory = [&](const string& name, const string&) -> App* {
    Node me = Postoffice::current().myNode();
    if (me.role() == NodeView::WORKER) return new SleepWorkerApp(name, run_ms);
    return new App(name);
  };
  c.launch(1, 2, 0, interval_s, factory);
  std::this_thread::sleep_for(std::chrono::milliseconds(run_ms * 9 / 10));
  string rep = c.offices[0]->dashboardReport();
  if (n_dead) *n_dead = (int32_t)c.offices[0]->deadNodes(dead_timeout_s).size();
  string stuck = c.join(30);
  if (!stuck.empty()) return fail("nodes did not stop: " + stuck);
  return put(rep, out, cap, len);
}

// ---- KVMap: per-key operators applied in timestamp order ------------------------------------------------
namespace {
class MapServerApp : public App {
 public:
  explicit MapServerApp(const string& name) : App(name), map_("m", name) {}
  ~MapServerApp() { stopExecutor(); }
  KVMap<Key, double> map_;
};
class MapWorkerApp : public App {
 public:
  MapWorkerApp(const string& name, int n, std::vector<double>* out) : App(name), map_("m", name), n_(n), out_(out) {}
  ~MapWorkerApp() { stopExecutor(); }
  void run() override {
    SArray<Key> key;
    for (int k = 0; k < n_; ++k) key.pushBack(spreadKey(k, n_));
    const std::pair<Operator, double> steps[] = {{PLUS, 5}, {TIMES, 3}, {MINUS, 1}, {DIVIDE, 2}};
    int t = -1;
    for (const auto& st : steps) {   // not waited for one by one: the executor keeps the order
      MessagePtr push(new Message(kServerGroup));
      push->setKey(key);
      push->addValue(SArray<double>(key.size(), st.second));
      KVMap<Key, double>::set(push).set_op(st.first);
      t = map_.push(push);
    }
    map_.waitOutMsg(kServerGroup, t);
    MessagePtr pull(new Message(kServerGroup));
    pull->setKey(key);
    map_.waitOutMsg(kServerGroup, map_.pull(pull));
    MessagePtr mine(new Message(kServerGroup));   // read my cached copy back through getValue
    mine->setKey(key);
    map_.getValue(mine);
    SArray<double> v(mine->value[0]);
    out_->assign(v.begin(), v.end());
  }
  KVMap<Key, double> map_;
  int n_;
  std::vector<double>* out_;
};
}  // namespace

// 2 servers hold an (initially empty) KVMap; one worker pushes +5, *3, -1, /2 on n spread keys and
// pulls: every value must be ((0 + 5) * 3 - 1) / 2 = 7
PSH_API int psh_run_kvmap_scenario(int32_t n_keys, double* out_vals) {
  std::vector<double> got;
  MiniCluster c;
  c.add(NodeView::SCHEDULER, "H");
  c.add(NodeView::SERVER, "S0");
  c.add(NodeView::SERVER, "S1");
  c.add(NodeView::WORKER, "W0");
  auto factory = [&](const string& name, const string&) -> App* {
    Node me = Postoffice::current().myNode();
    if (me.role() == NodeView::SERVER) return new MapServerApp(name);
    if (me.role() == NodeView::WORKER) return new MapWorkerApp(name, n_keys, &got);
    return new App(name);
  };
  c.launch(2, 1, 0, 0, factory);
  string stuck = c.join(30);
  if (!stuck.empty()) return fail("nodes did not stop: " + stuck);
  if ((int)got.size() != n_keys) return fail("worker read " + to_string(got.size()) + " values");
  for (int i = 0; i < n_keys; ++i) out_vals[i] = got[i];
  return 0;
}

// ---- van ring / key-range reassignment (ref src/test/van_test.cc, reassign_server_key_range.cc) -------
// n vans in one process; a message with a key array and a value array of `payload` elements goes
// around the ring n times (every van forwards what it receives to its successor).  Returns 0 when
// the payload arrives intact at node 0; *bytes = what node 0 sent.
PSH_API int psh_van_ring(int32_t n, int32_t use_tcp, int32_t base_port, int32_t payload, int64_t* bytes) {
  if (n < 2) return fail("ring needs at least 2 nodes");
  auto fabric = std::make_shared<InprocFabric>();
  std::vector<Node> nodes(n);
  std::vector<unique_ptr<Van>> vans(n);
  for (int i = 0; i < n; ++i) {
    nodes[i].set_role(NodeView::WORKER);
    nodes[i].set_id("N" + to_string(i));
    nodes[i].set_hostname("127.0.0.1");
    nodes[i].set_port(base_port + i);
    if (use_tcp) vans[i].reset(new TcpVan());
    else vans[i].reset(new InprocVan(fabric));
    if (!vans[i]->init(nodes[i])) return fail("van init failed for " + nodes[i].id());
  }
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j)
      if (i != j && !vans[i]->connect(nodes[j])) return fail("connect failed");
  SArray<Key> key(payload);
  SArray<float> val(payload);
  for (int i = 0; i < payload; ++i) { key[i] = (Key)i * 7919u; val[i] = 0.5f * i; }
  const int laps = n;
  std::atomic<int> errors(0);
  std::vector<std::thread> th;
  for (int i = 1; i < n; ++i) {
    th.emplace_back([&, i] {
      for (int lap = 0; lap < laps; ++lap) {
        MessagePtr m;
        size_t b = 0;
        if (!vans[i]->recv(&m, &b).ok()) { ++errors; return; }
        m->sender = nodes[i].id();
        m->recver = nodes[(i + 1) % n].id();
        if (!vans[i]->send(m, &b).ok()) { ++errors; return; }
      }
    });
  }
  MessagePtr cur(new Message());
  cur->task.set_type(TaskView::CALL_CUSTOMER);
  cur->task.set_customer("ring");
  cur->setKey(key);
  cur->addValue(val);
  size_t sent = 0;
  for (int lap = 0; lap < laps; ++lap) {
    cur->sender = nodes[0].id();
    cur->recver = nodes[1].id();
    size_t b = 0;
    if (!vans[0]->send(cur, &b).ok()) return fail("send failed");
    sent += b;
    MessagePtr back;
    if (!vans[0]->recv(&back, &b).ok()) return fail("recv failed");
    cur = back;
  }
  for (auto& t : th) t.join();
  for (auto& v : vans) v->shutdown();
  if (bytes) *bytes = (int64_t)sent;
  if (errors) return fail("a ring member failed");
  SArray<Key> k2(cur->key);
  if (cur->value.size() != 1) return fail("value array lost");
  SArray<float> v2(cur->value[0]);
  if (k2.size() != (size_t)payload || v2.size() != (size_t)payload) return fail("payload size changed");
  for (int i = 0; i < payload; ++i)
    if (k2[i] != key[i] || v2[i] != val[i]) return fail("payload corrupted at " + to_string(i));
  if (cur->task.customer() != "ring") return fail("task header lost");
  return 0;
}

namespace {
// scheduler: repartition the servers' key ranges inside [lo, hi) and broadcast ManageNode::UPDATE
class ReassignRoot : public App {
 public:
  ReassignRoot(const string& name, Key lo, Key hi) : App(name), lo_(lo), hi_(hi) {}
  void init() override {
    auto nodes = Postmaster::partitionServerKeyRange(sys_.yp().nodes(), Range<Key>(lo_, hi_));
    Task task;
    task.set_type(TaskView::MANAGE);
    task.mutable_mng_node().set_cmd(ManageNodeView::UPDATE);
    for (const auto& n : nodes) task.mutable_mng_node().add_node().raw()->copyFrom(*n.raw());
    port(kLiveGroup)->submitAndWait(task);
  }
  Key lo_, hi_;
};
class ReassignSlave : public App {
 public:
  ReassignSlave(const string& name, std::map<string, std::pair<Key, Key>>* out, std::mutex* mu)
      : App(name), out_(out), mu_(mu) {}
  void run() override {
    Range<Key> r = myKeyRange();
    Lock l(*mu_);
    (*out_)[myNodeID()] = std::make_pair(r.begin(), r.end());
  }
  std::map<string, std::pair<Key, Key>>* out_;
  std::mutex* mu_;
};
}  // namespace

// every node's key range as it sees it in run(), after the scheduler reassigned [lo, hi) in init():
// lines "<id> <begin> <end>"
PSH_API int psh_run_reassign_scenario(int32_t num_servers, int32_t num_workers, uint64_t lo, uint64_t hi,
                                      char* out, int64_t cap, int64_t* len) {
  std::map<string, std::pair<Key, Key>> got;
  std::mutex mu;
  MiniCluster c;
  c.add(NodeView::SCHEDULER, "H");
  for (int i = 0; i < num_servers; ++i) c.add(NodeView::SERVER, "S" + to_string(i));
  for (int i = 0; i < num_workers; ++i) c.add(NodeView::WORKER, "W" + to_string(i));
  auto factory = [&](const string& name, const string&) -> App* {
    Node me = Postoffice::current().myNode();
    if (me.role() == NodeView::SCHEDULER) return new ReassignRoot(name, lo, hi);
    return new ReassignSlave(name, &got, &mu);
  };
  c.launch(num_servers, num_workers, 0, 0, factory);
  string stuck = c.join(30);
  if (!stuck.empty()) return fail("nodes did not stop: " + stuck);
  std::ostringstream os;
  for (const auto& kv : got) os << kv.first << " " << kv.second.first << " " << kv.second.second << "\n";
  return put(os.str(), out, cap, len);
}

// ---- native unit checks (ref src/test/{bitmap,shared_array,sparse_matrix,localizer,common}_test.cc) --------
// Runs the C++-side checks of one area ("bitmap", "sarray", "matrix", "localizer", "match", "concurrent")
// or "all"; returns 0 and an empty message when every check holds, else the first failure.
namespace {
struct Checks {
  std::ostringstream err;
  int n = 0;
  void expect(bool ok, const char* what, int line) {
    ++n;
    if (!ok && err.str().empty()) err << what << " (capi.cc:" << line << ")";
  }
};
#define EXPECT(c) ck.expect((c), #c, __LINE__)

void checkBitmap(Checks& ck) {
  Bitmap b(100);
  EXPECT(b.size() == 100 && b.nnz() == 0);
  for (uint32 i = 0; i < 100; i += 3) b.set(i);
  EXPECT(b.nnz() == 34 && b.test(99) && !b.test(98) && b[0]);
  b.clear(0);
  EXPECT(!b[0] && b.nnz() == 33);
  b.fill(true);
  EXPECT(b.nnz() == 100);       // bits beyond size() stay clear
  b.resize(33, true);
  EXPECT(b.nnz() == 33 && b.memSize() == 8);
}

void checkSArray(Checks& ck) {
  SArray<int> a = {1, 3, 5, 7, 9, 11};
  SArray<int> b = {3, 4, 5, 6, 11, 20};
  EXPECT((a.setIntersection(b) == SArray<int>({3, 5, 11})));
  EXPECT((a.setUnion(b) == SArray<int>({1, 3, 4, 5, 6, 7, 9, 11, 20})));
  SArray<int> seg = a.segment(SizeR(1, 4));
  EXPECT(seg.size() == 3 && seg[0] == 3 && seg[2] == 7);
  seg[0] = 33;                                   // zero-copy view
  EXPECT(a[1] == 33);
  a[1] = 3;
  SizeR r = a.findRange(Range<int>(4, 10));
  EXPECT(r.begin() == 2 && r.end() == 5);
  SArray<double> d(1000);
  for (size_t i = 0; i < d.size(); ++i) d[i] = (i % 7) * 0.5;
  SArray<char> z = d.compressTo();
  SArray<double> e;
  e.uncompressFrom(z.data(), z.size());
  EXPECT(e == d && z.size() < d.size() * sizeof(double));
  SArray<uint64> bits(d);                        // constructing from another element type
  EXPECT(bits.size() == d.size() && bits[0] == 0);  // reinterprets the same bytes (ref _inl.h:73-80)
  SArray<char> raw(d);                           // shares memory
  EXPECT(raw.size() == d.size() * sizeof(double));
  EXPECT(d.nnz() == 1000 - 143);
  SArray<int> g;
  g.pushBack(4); g.pushBack(5); g.append(SArray<int>({6, 7}));
  EXPECT(g.size() == 4 && g.back() == 7 && g.front() == 4);
}

void checkMatrix(Checks& ck) {
  // 3 x 4 CSR:  [1 0 2 0; 0 3 0 0; 4 0 5 6]
  MatrixInfo info;
  info.set_type(MatrixInfoView::SPARSE);
  info.set_row_major(true);
  info.mutable_row().set_begin(0); info.mutable_row().set_end(3);
  info.mutable_col().set_begin(0); info.mutable_col().set_end(4);
  info.set_nnz(6); info.set_sizeof_index(4); info.set_sizeof_value(8);
  SArray<size_t> off = {0, 2, 3, 6};
  SArray<uint32> idx = {0, 2, 1, 0, 2, 3};
  SArray<double> val = {1, 2, 3, 4, 5, 6};
  MatrixPtr<double> X(new SparseMatrix<uint32, double>(info, off, idx, val));
  SArray<double> x = {1, 2, 3, 4};
  EXPECT((X->times(x) == SArray<double>({7, 6, 43})));
  SArray<double> y = {1, 1, 1};
  EXPECT((X->transTimes(y) == SArray<double>({5, 3, 7, 6})));
  MatrixPtr<double> C = X->toColMajor();
  EXPECT(C && C->colMajor() && (C->times(x) == SArray<double>({7, 6, 43})));
  EXPECT((C->transTimes(y) == SArray<double>({5, 3, 7, 6})));
  MatrixPtr<double> blk = C->colBlock(SizeR(2, 4));
  EXPECT(blk->cols() == 2 && (blk->times(SArray<double>({3, 4})) == SArray<double>({6, 0, 39})));
  MatrixPtr<double> rows = X->rowBlock(SizeR(1, 3));
  EXPECT(rows->rows() == 2 && (rows->times(x) == SArray<double>({6, 43})));
  MatrixPtr<double> sq = X->dotTimes(X);
  EXPECT((sq->times(SArray<double>({1, 1, 1, 1})) == SArray<double>({5, 9, 77})));
  DenseMatrix<double> D(2, 3);
  for (int i = 0; i < 6; ++i) D.value()[i] = i + 1;         // [1 2 3; 4 5 6]
  const Matrix<double>& Dm = D;  // the SArray overloads live in the base class
  EXPECT((Dm.times(SArray<double>({1, 0, 1})) == SArray<double>({4, 10})));
  EXPECT((Dm.transTimes(SArray<double>({1, 1})) == SArray<double>({5, 7, 9})));
}

void checkLocalizer(Checks& ck) {
  SArray<uint64> idx = {50, 7, 7, 900, 50, 7, 3};
  SArray<uint64> uniq;
  SArray<uint8> cnt;
  Localizer<uint64, double> lc;
  lc.countUniqIndex(idx, &uniq, &cnt);
  EXPECT((uniq == SArray<uint64>({3, 7, 50, 900})) && (cnt == SArray<uint8>({1, 3, 2, 1})));
  // 300 copies of one key saturate an 8-bit count instead of wrapping to 44
  SArray<uint64> many(300);
  many.setValue(5);
  SArray<uint64> u2;
  SArray<uint8> c2;
  Localizer<uint64, double> l2;
  l2.countUniqIndex(many, &u2, &c2);
  EXPECT(u2.size() == 1 && c2[0] == 255);
  // remap rows {50 7 | 7 900 50 | 7 3} onto the dictionary {7, 50}: columns 0/1, others dropped
  MatrixInfo info;
  info.set_type(MatrixInfoView::SPARSE);
  info.set_row_major(true);
  info.mutable_row().set_begin(0); info.mutable_row().set_end(3);
  info.mutable_col().set_begin(0); info.mutable_col().set_end(1000);
  info.set_nnz(7);
  SArray<size_t> off = {0, 2, 5, 7};
  SArray<double> val = {1, 2, 3, 4, 5, 6, 7};
  MatrixPtr<double> X = lc.remapIndex(info, off, idx, val, SArray<uint64>({7, 50}));
  EXPECT(X && X->rows() == 3 && X->cols() == 2 && X->nnz() == 5);
  EXPECT((X->times(SArray<double>({1, 10})) == SArray<double>({12, 53, 6})));
}

void checkMatch(Checks& ck) {
  SArray<uint64> sk = {2, 4, 6, 8};
  SArray<double> sv = {20, 40, 60, 80};
  SArray<uint64> dk = {1, 2, 3, 6, 7, 8, 9};
  SArray<double> dv;
  size_t n = parallelOrderedMatch<uint64, double>(sk, sv, dk, &dv);
  EXPECT(n == 3 && (dv == SArray<double>({0, 20, 0, 60, 0, 80, 0})));
  n = parallelOrderedMatch<uint64, double, OpPlus<double>>(sk, sv, dk, &dv);
  EXPECT(n == 3 && dv[1] == 40 && dv[5] == 160);
  SArray<uint64> ko;
  SArray<double> vo;
  parallelUnion<uint64, double>(sk, sv, SArray<uint64>({4, 5}), SArray<double>({1, 2}), &ko, &vo);
  EXPECT((ko == SArray<uint64>({2, 4, 5, 6, 8})) && (vo == SArray<double>({20, 41, 2, 60, 80})));
  SArray<int> big(200000);
  uint32 s = 12345;
  for (size_t i = 0; i < big.size(); ++i) { s = s * 1664525u + 1013904223u; big[i] = (int)(s >> 8); }
  parallelSort(&big, 4, [](int a, int b) { return a < b; });
  EXPECT(std::is_sorted(big.begin(), big.end()));
}

void checkConcurrent(Checks& ck) {
  std::atomic<int> sum(0);
  {
    ThreadPool pool(4);
    for (int i = 1; i <= 100; ++i) pool.add([&sum, i] { sum += i; });
    pool.startWorkers();
  }
  EXPECT(sum == 5050);
  threadsafe_queue<int> q;
  std::thread prod([&] { for (int i = 0; i < 1000; ++i) q.push(i); });
  long got = 0;
  for (int i = 0; i < 1000; ++i) { int v; q.wait_and_pop(v); got += v; }
  prod.join();
  EXPECT(got == 499500);
  Barrier bar(3);
  std::atomic<int> leaders(0);
  std::vector<std::thread> th;
  for (int i = 0; i < 3; ++i) th.emplace_back([&] { if (bar.block()) ++leaders; });
  for (auto& t : th) t.join();
  EXPECT(leaders == 1);
  Range<uint64> r(10, 110);
  EXPECT(r.evenDivide(4, 3).end() == 110 && r.evenDivide(4, 0).begin() == 10 && r.size() == 100);
  EXPECT(r.setIntersection(Range<uint64>(100, 500)) == Range<uint64>(100, 110));
  Timer t;
  t.start();
  std::this_thread::sleep_for(std::chrono::milliseconds(20));
  EXPECT(t.stop() >= 0.015);
}
#undef EXPECT
}  // namespace

PSH_API int psh_selftest(const char* area, int32_t* n_checks) {
  string a = area ? area : "all";
  Checks ck;
  if (a == "all" || a == "bitmap") checkBitmap(ck);
  if (a == "all" || a == "sarray") checkSArray(ck);
  if (a == "all" || a == "matrix") checkMatrix(ck);
  if (a == "all" || a == "localizer") checkLocalizer(ck);
  if (a == "all" || a == "match") checkMatch(ck);
  if (a == "all" || a == "concurrent") checkConcurrent(ck);
  if (n_checks) *n_checks = ck.n;
  if (ck.n == 0) return fail("unknown self-test area '" + a + "'");
  if (!ck.err.str().empty()) return fail("check failed: " + ck.err.str());
  return 0;
}

// ---- diagnostics ---------------------------------------------------------------------------------------
#include <execinfo.h>
#include <signal.h>
static void psh_crash_handler(int sig) {
  void* bt[64];
  int n = backtrace(bt, 64);
  fprintf(stderr, "psb200 host runtime: fatal signal %d\n", sig);
  backtrace_symbols_fd(bt, n, 2);
  _exit(128 + sig);
}
// print a native backtrace on SIGSEGV / SIGABRT (tests and tools call this first)
PSH_API void psh_install_crash_handler() {
  signal(SIGSEGV, psh_crash_handler);
  signal(SIGABRT, psh_crash_handler);
  signal(SIGFPE, psh_crash_handler);
}
