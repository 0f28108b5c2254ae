// Progress monitoring: slaves report a progress message to a master customer, which merges
// per sender and prints at a fixed interval.
// ref src/system/monitor.h:5-77 (MonitorMaster / MonitorSlaver, used by async SGD) and
//     src/system/dist_monitor.h:6-65 (DistMonitor: own reporter / printer threads).
#pragma once
#include "system/runtime.h"

namespace PS {

template <typename Progress>
class MonitorMaster : public Customer {
 public:
  MonitorMaster(const string& my_name, const string& parent_name) : Customer(my_name, parent_name) {}
  explicit MonitorMaster(Customer* parent)
      : MonitorMaster(parent->name() + "monitor", parent->name()) {}
  virtual ~MonitorMaster() { stopExecutor(); }

  typedef std::unordered_map<NodeID, Progress> AllProgress;
  typedef std::function<void(double time, AllProgress*)> Printer;
  typedef std::function<void(const Progress& src, Progress* dst)> Merger;
  void setPrinter(double time_interval, Printer printer) {
    Lock l(mu_);
    timer_.restart();
    printer_ = printer;
    interval_ = time_interval;
  }
  void setMerger(Merger merger) { Lock l(mu_); merger_ = merger; }

  void process(const MessagePtr& msg) override {
    Progress prog;
    if (!prog.ParseFromString(msg->task.msg())) return;
    Lock l(mu_);
    if (merger_) merger_(prog, &progress_[msg->sender]);
    else progress_[msg->sender] = prog;
    double time = timer_.stop();
    if (time > interval_ && printer_) {
      total_time_ += time;
      printer_(total_time_, &progress_);
      timer_.restart();
    } else {
      timer_.start();
    }
  }
  // print whatever has been merged so far (end of training)
  void flush() {
    Lock l(mu_);
    if (printer_ && !progress_.empty()) {
      total_time_ += timer_.stop();
      printer_(total_time_, &progress_);
      timer_.restart();
    }
  }

 private:
  std::mutex mu_;
  AllProgress progress_;
  double interval_ = 1;
  Timer timer_;
  double total_time_ = 0;
  Merger merger_;
  Printer printer_;
};

template <typename Progress>
class MonitorSlaver : public Customer {
 public:
  MonitorSlaver(const NodeID& master, const string& my_name, const string& parent_name)
      : Customer(my_name, parent_name), master_(master) {}
  MonitorSlaver(const NodeID& master, Customer* parent)
      : MonitorSlaver(master, parent->name() + "monitor", parent->name()) {}
  virtual ~MonitorSlaver() { stopExecutor(); }

  void report(const Progress& prog) {
    Task report;
    report.set_type(TaskView::CALL_CUSTOMER);
    report.set_msg(prog.SerializeAsString());
    RNodePtr master = port(master_);
    if (master) master->submit(report);
  }

 protected:
  NodeID master_;
};

// Same idea with its own threads: every node calls set() whenever it likes, a reporter thread
// ships the latest value every `interval` seconds, the master prints on its own thread.
template <typename Progress>
class DistMonitor : public Customer {
 public:
  DistMonitor(const NodeID& master, const string& my_name, const string& parent_name,
              double interval = 1)
      : Customer(my_name, parent_name), master_(master), interval_(interval) {}
  virtual ~DistMonitor() {
    done_ = true;
    if (thr_.joinable()) thr_.join();
    stopExecutor();
  }
  typedef std::function<void(double, std::map<NodeID, Progress>*)> Printer;
  void startMaster(Printer printer) {
    printer_ = printer;
    thr_ = std::thread([this] {
      Timer t;
      t.start();
      while (!done_) {
        std::this_thread::sleep_for(std::chrono::milliseconds((int)(interval_ * 1000)));
        Lock l(mu_);
        if (printer_ && !all_.empty()) printer_(t.stop(), &all_);
        t.start();
      }
    });
  }
  void startReporter() {
    thr_ = std::thread([this] {
      while (!done_) {
        std::this_thread::sleep_for(std::chrono::milliseconds((int)(interval_ * 1000)));
        Progress p;
        {
          Lock l(mu_);
          if (!dirty_) continue;
          p = mine_;
          dirty_ = false;
        }
        Task report;
        report.set_type(TaskView::CALL_CUSTOMER);
        report.set_msg(p.SerializeAsString());
        RNodePtr m = port(master_);
        if (m) m->submit(report);
      }
    });
  }
  void set(const Progress& p) { Lock l(mu_); mine_ = p; dirty_ = true; }
  void process(const MessagePtr& msg) override {
    Progress p;
    if (!p.ParseFromString(msg->task.msg())) return;
    Lock l(mu_);
    all_[msg->sender] = p;
  }

 private:
  NodeID master_;
  double interval_;
  std::atomic<bool> done_{false};
  std::mutex mu_;
  Progress mine_;
  bool dirty_ = false;
  std::map<NodeID, Progress> all_;
  Printer printer_;
  std::thread thr_;
};

}  // namespace PS
