// Thread pool, barrier, thread-safe queues, producer/consumer prefetcher.
// ref src/util/threadpool.{h,cc}, src/util/barrier.h, src/util/threadsafe_queue.h,
//     src/util/threadsafe_limited_queue.h, src/util/producer_consumer.h
#pragma once
#include <deque>
#include <queue>

#include "util/common.h"

namespace PS {

// add() tasks, then startWorkers(); the destructor (or wait()) joins.  Tasks added after the
// start are picked up as well, which the reference's pool does not allow.
class ThreadPool {
 public:
  explicit ThreadPool(int num_workers) : num_workers_(std::max(1, num_workers)) {}
  ~ThreadPool() { wait(); }
  typedef std::function<void()> Task;
  void add(const Task& task) {
    {
      Lock l(mu_);
      tasks_.push_back(task);
    }
    cv_.notify_one();
  }
  void startWorkers() {
    Lock l(mu_);
    if (started_) return;
    started_ = true;
    for (int i = 0; i < num_workers_; ++i) threads_.emplace_back([this] { run(); });
  }
  void wait() {
    startWorkers();
    {
      Lock l(mu_);
      done_ = true;
    }
    cv_.notify_all();
    for (auto& t : threads_)
      if (t.joinable()) t.join();
    threads_.clear();
  }

 private:
  void run() {
    while (true) {
      Task task;
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [this] { return done_ || !tasks_.empty(); });
        if (tasks_.empty()) return;
        task = std::move(tasks_.front());
        tasks_.pop_front();
      }
      task();
    }
  }
  int num_workers_;
  bool started_ = false, done_ = false;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<Task> tasks_;
  std::vector<std::thread> threads_;
};

// run fn(i) for i in [0,n) on up to `threads` threads and wait
inline void parallelFor(int n, int threads, const std::function<void(int)>& fn) {
  if (n <= 0) return;
  if (threads <= 1 || n == 1) {
    for (int i = 0; i < n; ++i) fn(i);
    return;
  }
  std::atomic<int> next(0);
  std::vector<std::thread> ts;
  int nt = std::min(threads, n);
  for (int t = 0; t < nt; ++t)
    ts.emplace_back([&] {
      for (int i = next++; i < n; i = next++) fn(i);
    });
  for (auto& t : ts) t.join();
}

// reusable barrier for a fixed number of threads
class Barrier {
 public:
  explicit Barrier(int n) : n_(n) {}
  bool block() {  // returns true for exactly one of the released threads
    std::unique_lock<std::mutex> l(mu_);
    int gen = gen_;
    if (++count_ == n_) {
      count_ = 0;
      ++gen_;
      cv_.notify_all();
      return true;
    }
    cv_.wait(l, [&] { return gen != gen_; });
    return false;
  }

 private:
  int n_, count_ = 0, gen_ = 0;
  std::mutex mu_;
  std::condition_variable cv_;
};

template <typename T>
class threadsafe_queue {
 public:
  void push(T v) {
    {
      Lock l(mu_);
      q_.push(std::move(v));
    }
    cv_.notify_one();
  }
  void wait_and_pop(T& v) {
    std::unique_lock<std::mutex> l(mu_);
    cv_.wait(l, [this] { return !q_.empty(); });
    v = std::move(q_.front());
    q_.pop();
  }
  bool try_pop(T& v) {
    Lock l(mu_);
    if (q_.empty()) return false;
    v = std::move(q_.front());
    q_.pop();
    return true;
  }
  // false on timeout
  bool wait_and_pop_for(T& v, int millis) {
    std::unique_lock<std::mutex> l(mu_);
    if (!cv_.wait_for(l, std::chrono::milliseconds(millis), [this] { return !q_.empty(); }))
      return false;
    v = std::move(q_.front());
    q_.pop();
    return true;
  }
  bool empty() const {
    Lock l(mu_);
    return q_.empty();
  }
  size_t size() const {
    Lock l(mu_);
    return q_.size();
  }

 private:
  mutable std::mutex mu_;
  std::queue<T> q_;
  std::condition_variable cv_;
};

// bounded by the total "capacity" (e.g. bytes) of queued items, not by their number
template <typename T>
class threadsafe_limited_queue {
 public:
  threadsafe_limited_queue() {}
  explicit threadsafe_limited_queue(size_t capacity) { setMaxCapacity(capacity); }
  void setMaxCapacity(size_t capacity) { max_capacity_ = capacity; }

  void push(T v, size_t capacity, bool finished = false) {
    std::unique_lock<std::mutex> l(mu_);
    full_cv_.wait(l, [&] { return cur_capacity_ + capacity <= max_capacity_ || q_.empty(); });
    q_.push(std::make_pair(std::move(v), capacity));
    cur_capacity_ += capacity;
    done_ = finished;
    l.unlock();
    empty_cv_.notify_all();
  }
  void setFinished() {
    {
      Lock l(mu_);
      done_ = true;
    }
    empty_cv_.notify_all();
  }
  // false once the producer finished and the queue drained
  bool pop(T& v) {
    std::unique_lock<std::mutex> l(mu_);
    empty_cv_.wait(l, [this] { return done_ || !q_.empty(); });
    if (q_.empty()) return false;
    v = std::move(q_.front().first);
    cur_capacity_ -= q_.front().second;
    q_.pop();
    l.unlock();
    full_cv_.notify_all();
    return true;
  }
  size_t size() const {
    Lock l(mu_);
    return q_.size();
  }
  bool empty() const { return size() == 0; }

 private:
  mutable std::mutex mu_;
  bool done_ = false;
  size_t max_capacity_ = 1 << 30, cur_capacity_ = 0;
  std::queue<std::pair<T, size_t>> q_;
  std::condition_variable empty_cv_, full_cv_;
};

// a producer thread fills a byte-bounded queue; pop() blocks the consumer
template <class V>
class ProducerConsumer {
 public:
  ProducerConsumer() { setCapacity(1000); }
  explicit ProducerConsumer(int capacity_in_mb) { setCapacity(capacity_in_mb); }
  ~ProducerConsumer() { if (thr_.joinable()) thr_.join(); }
  void setCapacity(int mb) { queue_.setMaxCapacity((size_t)mb * 1000000); }
  // func(&item, &size_in_bytes) returns false when exhausted
  void startProducer(const std::function<bool(V*, size_t*)>& func) {
    thr_ = std::thread([this, func] {
      V entry;
      bool done = false;
      while (!done) {
        size_t size = 0;
        done = !func(&entry, &size);
        if (!done) queue_.push(entry, size, false);
      }
      queue_.setFinished();
    });
  }
  bool pop(V* data) { return queue_.pop(*data); }

 private:
  threadsafe_limited_queue<V> queue_;
  std::thread thr_;
};

}  // namespace PS
