// Base definitions of the host runtime: integer types, Key, logging / CHECK macros, flags.
//
// The host runtime keeps the reference's public vocabulary (namespace PS, Key = uint64,
// LI / LL logging, CHECK_*), but depends on nothing outside libstdc++, pthreads and zlib:
// glog, gflags and protobuf (ref src/util/common.h:34-44) are replaced by the small
// facilities in this directory.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <future>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <tuple>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace PS {

typedef int8_t int8;
typedef int16_t int16;
typedef int32_t int32;
typedef int64_t int64;
typedef uint8_t uint8;
typedef uint16_t uint16;
typedef uint32_t uint32;
typedef uint64_t uint64;

// uint64 keys everywhere (ref src/util/common.h:54)
typedef uint64 Key;
static const Key kMaxKey = ~0ULL;
static const uint64 kuint64max = ~0ULL;
static const uint32 kuint32max = ~0U;

typedef std::lock_guard<std::mutex> Lock;
using std::string;
using std::shared_ptr;
using std::unique_ptr;
using std::pair;
using std::make_pair;
using std::to_string;
typedef std::vector<string> StringList;

// ---------------------------------------------------------------------------------------------
// logging: LI (info) / LL (error) streams and fatal CHECKs, thread-safe line output
// ---------------------------------------------------------------------------------------------
enum LogSeverity { LOG_INFO = 0, LOG_ERROR = 1, LOG_FATAL = 2 };

struct LogConfig {
  static LogConfig& get() {
    static LogConfig c;
    return c;
  }
  std::mutex mu;
  FILE* sink = stderr;    // -log_to_file redirects (ref postoffice.cc:48-54)
  int min_severity = 0;   // info is shown by default like glog's logtostderr in local.sh
  string prefix;          // node id once known
};

class LogMessage {
 public:
  LogMessage(const char* file, int line, LogSeverity sev) : sev_(sev) {
    const char* base = strrchr(file, '/');
    os_ << "[" << (sev == LOG_INFO ? "I" : sev == LOG_ERROR ? "E" : "F") << " "
        << (base ? base + 1 : file) << ":" << line << "] ";
  }
  ~LogMessage() noexcept(false) {
    LogConfig& c = LogConfig::get();
    if ((int)sev_ >= c.min_severity) {
      Lock l(c.mu);
      fprintf(c.sink, "%s%s\n", c.prefix.c_str(), os_.str().c_str());
      fflush(c.sink);
    }
    if (sev_ == LOG_FATAL) {
      if (c.sink != stderr) fprintf(stderr, "%s%s\n", c.prefix.c_str(), os_.str().c_str());
      abort();
    }
  }
  std::ostream& stream() { return os_; }

 private:
  LogSeverity sev_;
  std::ostringstream os_;
};

struct LogVoidify {
  void operator&(std::ostream&) {}
};

#define LI ::PS::LogMessage(__FILE__, __LINE__, ::PS::LOG_INFO).stream()
#define LL ::PS::LogMessage(__FILE__, __LINE__, ::PS::LOG_ERROR).stream()
#define LOG_FATAL_STREAM ::PS::LogMessage(__FILE__, __LINE__, ::PS::LOG_FATAL).stream()

#define CHECK(cond) \
  (cond) ? (void)0 : ::PS::LogVoidify() & LOG_FATAL_STREAM << "Check failed: " #cond " "
#define CHECK_OP(a, b, op)                                                             \
  ((a)op(b)) ? (void)0                                                                 \
             : ::PS::LogVoidify() & LOG_FATAL_STREAM << "Check failed: " #a " " #op " " #b \
                                                       << " (" << (a) << " vs " << (b) << ") "
#define CHECK_EQ(a, b) CHECK_OP(a, b, ==)
#define CHECK_NE(a, b) CHECK_OP(a, b, !=)
#define CHECK_LT(a, b) CHECK_OP(a, b, <)
#define CHECK_LE(a, b) CHECK_OP(a, b, <=)
#define CHECK_GT(a, b) CHECK_OP(a, b, >)
#define CHECK_GE(a, b) CHECK_OP(a, b, >=)
#define CHECK_NOTNULL(p) ::PS::CheckNotNull(__FILE__, __LINE__, #p, (p))

template <typename T>
T CheckNotNull(const char* file, int line, const char* name, T&& t) {
  if (t == nullptr) LogMessage(file, line, LOG_FATAL).stream() << name << " must not be NULL";
  return std::forward<T>(t);
}

#define DISALLOW_COPY_AND_ASSIGN(T) \
  T(const T&) = delete;             \
  void operator=(const T&) = delete

// population count of a 32-bit word (ref common.h:86-90)
inline int32 NumberOfSetBits(uint32 i) { return __builtin_popcount(i); }

// "[n]: a b c ... x y z" (ref common.h:92-105)
template <typename V>
inline string dbstr(const V* data, int n, int m = 5) {
  std::stringstream ss;
  ss << "[" << n << "]: ";
  if (n < 2 * m) {
    for (int i = 0; i < n; ++i) ss << data[i] << " ";
  } else {
    for (int i = 0; i < m; ++i) ss << data[i] << " ";
    ss << "... ";
    for (int i = n - m; i < n; ++i) ss << data[i] << " ";
  }
  return ss.str();
}

// ---------------------------------------------------------------------------------------------
// timers (ref common.h tic/toc, resource_usage.h Timer / MilliTimer)
// ---------------------------------------------------------------------------------------------
typedef std::chrono::steady_clock::time_point TimePoint;
inline TimePoint tic() { return std::chrono::steady_clock::now(); }
inline double toc(TimePoint start) {
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - start).count();
}

class Timer {
 public:
  void start() { tp_ = tic(); }
  void restart() { time_ = 0; start(); }
  void reset() { time_ = 0; }
  double stop() { time_ += toc(tp_); return time_; }
  double get() const { return time_; }
  double getAndRestart() { double t = stop(); restart(); return t; }

 private:
  TimePoint tp_ = tic();
  double time_ = 0;
};

class ScopedTimer {
 public:
  explicit ScopedTimer(double* acc) : acc_(acc), tp_(tic()) {}
  ~ScopedTimer() { *acc_ += toc(tp_); }

 private:
  double* acc_;
  TimePoint tp_;
};

}  // namespace PS
