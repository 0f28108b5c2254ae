// Per-node resource sampling (HeartbeatInfo) and the scheduler-side table (Dashboard).
// ref src/system/heartbeat_info.{h,cc}, src/system/dashboard.{h,cc}.  In the reference both
// exist but every call site is commented out (postoffice.cc:220-249, 355-398); here
// -report_interval N wires them: nodes send HEARTBEATING every N seconds, the scheduler keeps
// the table, prints it with -verbose and uses it for failure detection.
#pragma once
#include <sys/time.h>

#include "proto/messages.h"
#include "util/file.h"

namespace PS {

class HeartbeatInfo {
 public:
  void init(const string& hostname) {
    hostname_ = hostname;
    start_ = tic();
    sampleCpu(&last_proc_, &last_host_total_, &last_host_idle_);
  }
  void increaseInBytes(size_t d) { in_bytes_ += d; }
  void increaseOutBytes(size_t d) { out_bytes_ += d; }
  void startBusy() { Lock l(mu_); busy_.start(); }
  void stopBusy() { Lock l(mu_); busy_.stop(); }

  HeartbeatReport get() {
    HeartbeatReport r;
    Lock l(mu_);
    r.set_hostname(hostname_);
    struct timeval tv;
    gettimeofday(&tv, nullptr);
    r.set_seconds_since_epoch((uint64)tv.tv_sec);
    double total = toc(start_);
    r.set_total_time_milli((uint64)(total * 1e3));
    r.set_busy_time_milli((uint64)(busy_.get() * 1e3));
    r.set_net_in_mb(in_bytes_.exchange(0) >> 20);
    r.set_net_out_mb(out_bytes_.exchange(0) >> 20);
    uint64 proc, host_total, host_idle;
    sampleCpu(&proc, &host_total, &host_idle);
    uint64 dt = host_total - last_host_total_;
    if (dt) {
      long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
      r.set_process_cpu_usage((uint64)(100.0 * ncpu * (proc - last_proc_) / dt));
      r.set_host_cpu_usage((uint64)(100.0 * (dt - (host_idle - last_host_idle_)) / dt));
    }
    last_proc_ = proc; last_host_total_ = host_total; last_host_idle_ = host_idle;
    uint64 rss_kb = 0, virt_kb = 0;
    procStatus(&rss_kb, &virt_kb);
    r.set_process_rss_mb(rss_kb >> 10);
    r.set_process_virt_mb(virt_kb >> 10);
    uint64 mem_total = 0, mem_avail = 0;
    memInfo(&mem_total, &mem_avail);
    if (mem_total) {
      r.set_host_in_use_gb((mem_total - mem_avail) >> 20);
      r.set_host_in_use_percentage(100 * (mem_total - mem_avail) / mem_total);
    }
    return r;
  }

 private:
  static void sampleCpu(uint64* proc, uint64* host_total, uint64* host_idle) {
    *proc = *host_total = *host_idle = 0;
    string s;
    if (readFileToString("/proc/self/stat", &s)) {
      size_t p = s.rfind(')');
      auto t = split(p == string::npos ? s : s.substr(p + 2), ' ', true);
      if (t.size() > 12) *proc = strtoull(t[11].c_str(), nullptr, 10) + strtoull(t[12].c_str(), nullptr, 10);
    }
    if (readFileToString("/proc/stat", &s)) {
      auto t = split(s.substr(0, s.find('\n')), ' ', true);
      for (size_t i = 1; i < t.size(); ++i) {
        uint64 v = strtoull(t[i].c_str(), nullptr, 10);
        *host_total += v;
        if (i == 4 || i == 5) *host_idle += v;
      }
    }
  }
  static void procStatus(uint64* rss_kb, uint64* virt_kb) {
    string s;
    if (!readFileToString("/proc/self/status", &s)) return;
    for (const auto& line : split(s, '\n', true)) {
      if (line.compare(0, 6, "VmRSS:") == 0) *rss_kb = strtoull(line.c_str() + 6, nullptr, 10);
      if (line.compare(0, 7, "VmSize:") == 0) *virt_kb = strtoull(line.c_str() + 7, nullptr, 10);
    }
  }
  static void memInfo(uint64* total_kb, uint64* avail_kb) {
    string s;
    if (!readFileToString("/proc/meminfo", &s)) return;
    for (const auto& line : split(s, '\n', true)) {
      if (line.compare(0, 9, "MemTotal:") == 0) *total_kb = strtoull(line.c_str() + 9, nullptr, 10);
      if (line.compare(0, 13, "MemAvailable:") == 0) *avail_kb = strtoull(line.c_str() + 13, nullptr, 10);
    }
  }
  std::mutex mu_;
  string hostname_;
  TimePoint start_ = tic();
  Timer busy_;
  std::atomic<size_t> in_bytes_{0}, out_bytes_{0};
  uint64 last_proc_ = 0, last_host_total_ = 0, last_host_idle_ = 0;
};

class Dashboard {
 public:
  void addReport(const string& node_id, const HeartbeatReportView& report) {
    Lock l(mu_);
    auto& e = data_[node_id];
    e.report = report;
    e.last_seen = tic();
  }
  void addTask(const string& node_id, int task_id) {
    Lock l(mu_);
    data_[node_id].report.set_task_id(task_id);
  }
  void forget(const string& node_id) {
    Lock l(mu_);
    data_.erase(node_id);
  }
  // nodes that reported before but not within the last `timeout_sec` seconds
  std::vector<string> deadNodes(int timeout_sec) {
    Lock l(mu_);
    std::vector<string> r;
    for (auto& kv : data_)
      if (toc(kv.second.last_seen) > timeout_sec) r.push_back(kv.first);
    return r;
  }
  string title() const {
    return "  Node  Task  MyCPU(%) MyRSS(M) MyVir(M) BusyTime(%) InMB OutMB HostCPU(%) HostUseGB HostUse(%)";
  }
  string report() {
    Lock l(mu_);
    std::ostringstream os;
    os << title() << "\n";
    for (auto& kv : data_) {
      const auto& r = kv.second.report;
      char buf[512];
      double busy = r.total_time_milli() ? 100.0 * r.busy_time_milli() / r.total_time_milli() : 0;
      snprintf(buf, sizeof(buf), "%6s %5d %9llu %8llu %8llu %11.1f %4llu %5llu %10llu %9llu %10llu",
               kv.first.c_str(), r.task_id(), (unsigned long long)r.process_cpu_usage(),
               (unsigned long long)r.process_rss_mb(), (unsigned long long)r.process_virt_mb(), busy,
               (unsigned long long)r.net_in_mb(), (unsigned long long)r.net_out_mb(),
               (unsigned long long)r.host_cpu_usage(), (unsigned long long)r.host_in_use_gb(),
               (unsigned long long)r.host_in_use_percentage());
      os << buf << "\n";
    }
    return os.str();
  }
  size_t size() {
    Lock l(mu_);
    return data_.size();
  }

 private:
  struct NaturalIdLess {
    bool operator()(const string& a, const string& b) const {
      // role letter first, then numeric suffix: S2 before S10
      if (a.empty() || b.empty() || a[0] != b[0]) return a < b;
      return atoi(a.c_str() + 1) < atoi(b.c_str() + 1) || (atoi(a.c_str() + 1) == atoi(b.c_str() + 1) && a < b);
    }
  };
  struct Entry {
    HeartbeatReport report;
    TimePoint last_seen = tic();
  };
  std::mutex mu_;
  std::map<string, Entry, NaturalIdLess> data_;
};

}  // namespace PS
