// build/selftest: the native unit checks and the in-process cluster scenarios in one binary, so
// that they can run under ThreadSanitizer / AddressSanitizer (tools/sanitize.py) - the reference has
// no sanitizer story at all.
//   build/selftest [all|units|kv|kvmap|ring|replace|heartbeat|reassign|apps]
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <unistd.h>
#include <string>
#include <vector>

extern "C" {
const char* psh_last_error();
int psh_selftest(const char* area, int32_t* n_checks);
int psh_run_kv_cluster(int32_t num_servers, int32_t num_workers, int32_t n_keys, const uint64_t* pull_keys,
                       int32_t n_pull, int32_t use_tcp, int32_t base_port, float* out_vals);
int psh_van_ring(int32_t n, int32_t use_tcp, int32_t base_port, int32_t payload, int64_t* bytes);
int psh_run_replace_scenario(int32_t n_keys, float* out_vals, uint64_t* s1_range2);
int psh_run_heartbeat_scenario(int32_t interval_s, int32_t run_ms, int32_t dead_timeout_s, char* out,
                               int64_t cap, int64_t* len, int32_t* n_dead);
int psh_run_cluster(const char* app_conf_text, int32_t num_servers, int32_t num_workers, int32_t use_tcp,
                    int32_t base_port, int32_t message_compression, char* report_out, int64_t cap, int64_t* len);
int psh_run_kvmap_scenario(int32_t n_keys, double* out_vals);
int psh_run_reassign_scenario(int32_t num_servers, int32_t num_workers, uint64_t lo, uint64_t hi, char* out,
                              int64_t cap, int64_t* len);
}

// a few hundred separable libsvm examples in `parts` files under dir/train
static void writeData(const std::string& dir, int parts, int rows, int feats) {
  std::string cmd = "mkdir -p " + dir + "/train " + dir + "/model " + dir + "/cache";
  if (system(cmd.c_str()) != 0) return;
  uint32_t s = 2024;
  auto rnd = [&] { s = s * 1664525u + 1013904223u; return s >> 8; };
  std::vector<float> w(feats);
  for (auto& x : w) x = (rnd() % 7 == 0) ? ((int)(rnd() % 200) - 100) / 25.0f : 0.f;
  for (int p = 0; p < parts; ++p) {
    FILE* f = fopen((dir + "/train/part-" + std::to_string(p)).c_str(), "w");
    if (!f) return;
    for (int r = 0; r < rows; ++r) {
      std::vector<int> idx;
      for (int j = 0; j < feats; ++j) if (rnd() % 8 == 0) idx.push_back(j);
      if (idx.empty()) idx.push_back((int)(rnd() % feats));
      float m = 0;
      std::vector<float> v(idx.size());
      for (size_t k = 0; k < idx.size(); ++k) { v[k] = (rnd() % 100 + 1) / 100.f; m += v[k] * w[idx[k]]; }
      fprintf(f, "%s", m + ((int)(rnd() % 100) - 50) / 200.f > 0 ? "+1" : "-1");
      for (size_t k = 0; k < idx.size(); ++k) fprintf(f, " %d:%g", idx[k] + 1, v[k]);
      fprintf(f, "\n");
    }
    fclose(f);
  }
}

static int failures = 0;
static void report(const char* what, int rc) {
  if (rc) { ++failures; fprintf(stderr, "FAIL %s: %s\n", what, psh_last_error()); }
  else fprintf(stderr, "ok   %s\n", what);
}

int main(int argc, char** argv) {
  std::string which = argc > 1 ? argv[1] : "all";
  auto want = [&](const char* s) { return which == "all" || which == s; };
  if (want("units")) {
    int32_t n = 0;
    report("units", psh_selftest("all", &n));
  }
  if (want("kv")) {
    uint64_t keys[4] = {0, 2, 4, 5};
    float out[4];
    int rc = psh_run_kv_cluster(3, 2, 6, keys, 4, 0, 0, out);
    if (!rc && !(out[1] > 2.19f && out[1] < 2.21f)) { rc = 1; fprintf(stderr, "pulled %g for key 2, expected 0.2 + one push per worker\n", out[1]); }
    report("kv cluster (in-process van)", rc);
    rc = psh_run_kv_cluster(2, 2, 6, keys, 4, 1, 24500 + (int)(getpid() % 2000), out);
    report("kv cluster (tcp van)", rc);
  }
  if (want("ring")) {
    int64_t bytes = 0;
    report("van ring (in-process)", psh_van_ring(4, 0, 0, 2000, &bytes));
    report("van ring (tcp)", psh_van_ring(3, 1, 26500 + (int)(getpid() % 2000), 2000, &bytes));
  }
  if (want("kvmap")) {
    std::vector<double> vals(16);
    int rc = psh_run_kvmap_scenario(16, vals.data());
    for (double v : vals) if (!rc && v != 7.0) rc = 1;
    report("kvmap operators in order", rc);
  }
  if (want("reassign")) {
    std::vector<char> buf(1 << 16);
    int64_t len = 0;
    report("key-range reassignment", psh_run_reassign_scenario(3, 2, 0, 100, buf.data(), buf.size(), &len));
  }
  if (want("replace")) {
    std::vector<float> vals(40);
    uint64_t rng[2];
    report("node replacement with replica recovery", psh_run_replace_scenario(40, vals.data(), rng));
  }
  if (want("heartbeat")) {
    std::vector<char> buf(1 << 16);
    int64_t len = 0;
    int32_t dead = 0;
    report("heartbeat / dashboard", psh_run_heartbeat_scenario(1, 1500, 60, buf.data(), buf.size(), &len, &dead));
  }
  if (want("apps")) {
    const std::string dir = "/tmp/psb200_selftest_" + std::to_string(getpid());
    writeData(dir, 4, 300, 120);
    const std::string head = "linear_method { training_data { format: TEXT text: LIBSVM file: \"" + dir +
        "/train/part.*\" } model_output { format: TEXT file: \"" + dir + "/model/m\" } loss { type: LOGIT } ";
    const std::string darlin = head + "penalty { type: L1 lambda: .5 } learning_rate { type: CONSTANT alpha: 1 } "
        "darlin { max_pass_of_data: 4 epsilon: 1e-7 feature_block_ratio: 1 max_block_delay: 1 "
        "local_cache { format: BIN file: \"" + dir + "/cache/c\" } } }";
    const std::string sgd = head + "penalty { type: L2 lambda: .0001 } learning_rate { type: DECAY alpha: .5 beta: 1 } "
        "async_sgd { algo: FTRL minibatch: 100 num_data_pass: 2 max_delay: 2 report_interval: 0 "
        "fixing_float_by_nbytes: 1 tail_feature_freq: 1 } }";
    std::vector<char> buf(1 << 20);
    int64_t len = 0;
    report("darlin cluster (2 servers, 2 workers, in-process van)",
           psh_run_cluster(darlin.c_str(), 2, 2, 0, 0, 1, buf.data(), buf.size(), &len));
    report("darlin cluster (tcp van)",
           psh_run_cluster(darlin.c_str(), 2, 2, 1, 28500 + (int)(getpid() % 2000), 1, buf.data(), buf.size(), &len));
    report("async sgd cluster (1 server, 2 workers)",
           psh_run_cluster(sgd.c_str(), 1, 2, 0, 0, 1, buf.data(), buf.size(), &len));
    if (system(("rm -rf " + dir).c_str()) != 0) ++failures;
  }
  fprintf(stderr, "%d failure(s)\n", failures);
  return failures ? 1 : 0;
}
