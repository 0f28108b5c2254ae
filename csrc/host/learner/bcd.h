// Block coordinate descent over the parameter server: the parts that do not depend on the
// objective.  ref src/learner/bcd.h:10-518, src/learner/proto/bcd.proto
//
// Building blocks (each is a small object instead of code spread over three roles):
//   BcdClock          scheduler time t  ->  model-channel times (push 10t, updated 10t+1, pull 10t+2)
//   BlockPlanner      feature groups -> key-range blocks, block order of a pass, prior blocks
//   ProgressBoard     merges the nodes' BCDProgress per pass and prints the table column by column
//   PrepSchedule      the timestamps of the two preprocessing handshakes, shared by server and
//                     worker so that the protocol is written down once
//   BCDNode           command table: a role registers one handler per BCDCall command
//   BCDScheduler / BCDServer<V> / BCDWorker<V>   the three roles
//
// The GPU engine (parameter_server_b200/models/darlin.py + csrc/cuda/darlin.cu) implements the
// same roles on the device; this host version runs over the van, with its model shards in HBM
// when started with -device gpu (KVBufferedVector's device column store).
#pragma once
#include <cstdarg>

#include "data/slot_reader.h"
#include "parameter/kv.h"
#include "util/localizer.h"

namespace PS {

struct BcdClock {
  static const int kStride = 10;
  static int push(int scheduler_time) { return scheduler_time * kStride; }
  static int updated(int push_time) { return push_time + 1; }  // raised by the server when the block is applied
  static int pull(int push_time) { return push_time + 2; }     // waits for updated()
};

// Preprocessing runs two handshakes per feature group, all groups of round 1 before round 2:
//   round 1  workers push (keys, counts) at count(i); servers release the filtered-key pull
//   round 2  workers push the kept keys at keys(i); servers build their shard, release the
//            initial-weight pull
struct PrepSchedule {
  int base, groups;
  int count(int i) const { return base + BcdClock::kStride * i; }
  int keys(int i) const { return base + BcdClock::kStride * (groups + i); }
};

struct FeatureBlock {
  int group;
  Range<Key> keys;
};

// ref BCDScheduler::divideFeatureBlocks (bcd.h:78-122)
class BlockPlanner {
 public:
  // a group whose rows carry more than one non-zero is cut into ceil(nnz/row * ratio) equal key
  // ranges; groups with (about) one non-zero per row have independent coordinates: one block
  void build(const ExampleInfoView& info, const BCDConfigView& conf, std::mt19937* rng) {
    blocks_.clear();
    prior_.clear();
    for (int s = 0; s < info.slot_size(); ++s) {
      const SlotInfoView slot = info.slot(s);
      if (slot.id() == 0) continue;  // the label slot
      CHECK(slot.has_nnz_ele() && slot.has_nnz_ex()) << "slot " << slot.id() << " has no statistics";
      const double per_row = (double)slot.nnz_ele() / (double)slot.nnz_ex();
      const int pieces = per_row > 1 + 1e-6
                             ? std::max(1, (int)std::ceil(per_row * conf.feature_block_ratio()))
                             : 1;
      const Range<Key> whole(slot.min_key(), slot.max_key());
      for (int p = 0; p < pieces; ++p) {
        const Range<Key> piece = whole.evenDivide(pieces, p);
        if (!piece.empty()) blocks_.push_back(FeatureBlock{(int)slot.id(), piece});
      }
    }
    // prior groups are swept several times before the first regular pass
    for (int g = 0; g < conf.prior_fea_group_size(); ++g) {
      std::vector<int> mine;
      for (size_t b = 0; b < blocks_.size(); ++b)
        if (blocks_[b].group == conf.prior_fea_group(g)) mine.push_back((int)b);
      if (mine.empty()) continue;
      prior_groups_.push_back(to_string(conf.prior_fea_group(g)));
      for (int sweep = 0; sweep < conf.num_iter_for_prior_fea_group(); ++sweep) {
        if (conf.random_feature_block_order()) std::shuffle(mine.begin(), mine.end(), *rng);
        prior_.insert(prior_.end(), mine.begin(), mine.end());
      }
    }
  }
  // block ids of pass `iter`; the first numPrior(iter) of them are prior blocks
  std::vector<int> pass(int iter, bool shuffle, std::mt19937* rng) const {
    std::vector<int> order(blocks_.size());
    for (size_t b = 0; b < order.size(); ++b) order[b] = (int)b;
    if (shuffle) std::shuffle(order.begin(), order.end(), *rng);
    if (iter == 0) order.insert(order.begin(), prior_.begin(), prior_.end());
    return order;
  }
  size_t numPrior(int iter) const { return iter == 0 ? prior_.size() : 0; }
  size_t size() const { return blocks_.size(); }
  const FeatureBlock& operator[](int b) const { return blocks_[b]; }
  const std::vector<string>& priorGroups() const { return prior_groups_; }

 private:
  std::vector<FeatureBlock> blocks_;
  std::vector<int> prior_;
  std::vector<string> prior_groups_;
};

// ref BCDScheduler::mergeProgress / showTime / showObjective (bcd.h:131-178)
class ProgressBoard {
 public:
  // a column of the per-pass table: three header lines and a cell
  struct Column {
    string head[3];
    std::function<string(int iter, const BCDProgress&)> cell;
  };
  ProgressBoard() {
    columns_.push_back(Column{{"     |        training        |  sparsity ",
                               "iter |  objective    relative |     |w|_0 ",
                               " ----+------------------------+-----------"},
                              [](int iter, const BCDProgress& p) {
                                return fmt("%4d | %.5e  %.3e |%10lu ", iter, p.objective(), p.relative_obj(),
                                           (size_t)p.nnz_w());
                              }});
    columns_.push_back(Column{{"|    time (sec.)", "|(app:min max) total", "+-----------------"},
                              [this](int iter, const BCDProgress& p) {
                                double lo = 0, hi = 0;
                                for (int i = 0; i < p.busy_time_size(); ++i) {
                                  lo = i ? std::min(lo, p.busy_time(i)) : p.busy_time(i);
                                  hi = std::max(hi, p.busy_time(i));
                                }
                                const double before = iter > 0 ? rows_[iter - 1].total_time() : 0;
                                return fmt("|%6.1f%6.1f%6.1f", lo, hi, p.total_time() - before);
                              }});
  }
  // columns are printed in insertion order; `at` = position of the new column
  void insertColumn(size_t at, const Column& c) { columns_.insert(columns_.begin() + at, c); }
  void startClock() { clock_.restart(); }

  // one node's report of pass `iter`
  void merge(int iter, const BCDProgress& got) {
    Lock l(mu_);
    BCDProgress& row = rows_[iter];
    row.set_objective(row.objective() + got.objective());
    row.set_nnz_w(row.nnz_w() + got.nnz_w());
    row.set_nnz_active_set(row.nnz_active_set() + got.nnz_active_set());
    row.set_violation(std::max(row.violation(), got.violation()));
    if (got.busy_time_size() > 0) row.add_busy_time(got.busy_time(0));
    row.set_total_time(clock_.stop());
    clock_.start();
    row.set_relative_obj(iter == 0 ? 1 : rows_[iter - 1].objective() / row.objective() - 1);
  }
  void print(int iter) {
    Lock l(mu_);
    if (iter == 0)
      for (int h = 0; h < 3; ++h) {
        for (const Column& c : columns_) fputs(c.head[h].c_str(), stderr);
        fputc('\n', stderr);
      }
    for (const Column& c : columns_) fputs(c.cell(iter, rows_[iter]).c_str(), stderr);
    fputc('\n', stderr);
  }
  BCDProgress row(int iter) { Lock l(mu_); return rows_[iter]; }
  const std::map<int, BCDProgress>& rows() const { return rows_; }

  static string fmt(const char* f, ...) {
    char buf[256];
    va_list ap;
    va_start(ap, f);
    vsnprintf(buf, sizeof(buf), f, ap);
    va_end(ap);
    return buf;
  }

 private:
  std::mutex mu_;
  std::map<int, BCDProgress> rows_;
  std::vector<Column> columns_;
  Timer clock_;
};

// what scheduler, server and worker have in common: the config and a command table
class BCDNode : public App {
 public:
  BCDNode(const string& name, const BCDConfigView& conf) : App(name), bcd_conf_(conf) {
    if (!bcd_conf_.has_local_cache()) bcd_conf_.mutable_local_cache().add_file("/tmp/bcd_");
  }
  // a BCDCall carries a command; `time` is the model-channel time of the scheduler's task
  void process(const MessagePtr& msg) override {
    CHECK(msg->task.has_bcd());
    const BCDCallView call = msg->task.bcd();
    auto h = handlers_.find((int)call.cmd());
    if (h != handlers_.end()) h->second(msg, call, BcdClock::push(msg->task.time()));
  }

 protected:
  typedef std::function<void(const MessagePtr&, const BCDCallView&, int time)> Handler;
  void on(BCDCallView::Command cmd, Handler h) { handlers_[(int)cmd] = std::move(h); }
  void useGroups(const BCDCallView& call) {
    fea_grp_.clear();
    for (int i = 0; i < call.fea_grp_size(); ++i) fea_grp_.push_back(call.fea_grp(i));
  }
  template <class Progress>
  void answer(const MessagePtr& msg, const Progress& p) { sys_.replyProtocalMessage(msg, p); }

  BCDConfig bcd_conf_;
  std::vector<int> fea_grp_;

 private:
  std::map<int, Handler> handlers_;
};

// ---------------------------------------------------------------------------------------------------
class BCDScheduler : public BCDNode {
 public:
  BCDScheduler(const string& name, const BCDConfigView& conf) : BCDNode(name, conf) {}
  virtual ~BCDScheduler() {}
  const std::map<int, BCDProgress>& progress() const { return board_.rows(); }

 protected:
  // the files go round robin to the workers (LOAD_DATA), then every computation node
  // localises them (PREPROCESS_DATA); ref bcd.h:32-76
  void loadTrainingData(const DataConfigView& train) {
    const auto t_load = tic();
    const std::vector<DataConfig> parts = Postmaster::partitionData(train, sys_.yp().num_workers());
    std::vector<Task> load(parts.size());
    for (size_t w = 0; w < parts.size(); ++w) {
      BCDCallView call = load[w].mutable_bcd();
      call.set_cmd(BCDCallView::LOAD_DATA);
      call.mutable_data().raw()->copyFrom(*parts[w].raw());
    }
    int cached = 0;
    port(kWorkerGroup)->submitAndWait(load, [&]() {
      LoadDataReturn got;
      CHECK(got.ParseFromString(exec().lastRecvReply()));
      Lock l(info_mu_);
      train_info_ = mergeExampleInfo(train_info_, got.example_info());
      cached += got.hit_cache();
    });
    CHECK(cached == 0 || cached == (int)load.size()) << "some workers hit their cache, some did not: clear the local caches";
    LI << "Loaded " << train_info_.num_ex() << " examples in " << toc(t_load) << " sec"
       << (cached ? " (from the local caches)" : "");
    for (int s = 0; s < train_info_.slot_size(); ++s)
      if (train_info_.slot(s).id() != 0) fea_grp_.push_back(train_info_.slot(s).id());

    const auto t_prep = tic();
    Task prep;
    BCDCallView call = prep.mutable_bcd();
    call.set_cmd(BCDCallView::PREPROCESS_DATA);
    for (int g : fea_grp_) call.add_fea_grp(g);
    call.set_hit_cache(cached > 0);
    port(kCompGroup)->submitAndWait(prep);
    LI << "Preprocessing is finished in " << toc(t_prep) << " sec";
    if (bcd_conf_.tail_feature_freq())
      LI << "Features with frequency <= " << bcd_conf_.tail_feature_freq() << " are filtered";
  }

  void divideFeatureBlocks() {
    planner_.build(train_info_, bcd_conf_, &rng_);
    LI << "Features are partitioned into " << planner_.size() << " blocks";
    if (!planner_.priorGroups().empty()) LI << "Prior feature groups: " << join(planner_.priorGroups(), ", ");
    board_.startClock();
  }

  void saveModel(const DataConfigView& where) {
    Task save;
    save.mutable_bcd().set_cmd(BCDCallView::SAVE_MODEL);
    save.mutable_bcd().mutable_data().raw()->copyFrom(*where.raw());
    port(kServerGroup)->submitAndWait(save);
  }

  // reply handler of EVALUATE_PROGRESS
  void mergeProgress(int iter) {
    BCDProgress got;
    CHECK(got.ParseFromString(exec().lastRecvReply()));
    board_.merge(iter, got);
  }

  BlockPlanner planner_;
  ProgressBoard board_;
  ExampleInfo train_info_;
  std::mutex info_mu_;
  std::mt19937 rng_{20140624};
};

// ---------------------------------------------------------------------------------------------------
template <typename V>
class BCDServer : public BCDNode {
 public:
  BCDServer(const string& name, const BCDConfigView& conf) : BCDNode(name, conf), model_(name + "_model", name) {
    on(BCDCallView::PREPROCESS_DATA, [this](const MessagePtr&, const BCDCallView& c, int t) { preprocessData(t, c); });
    on(BCDCallView::UPDATE_MODEL, [this](const MessagePtr&, const BCDCallView& c, int t) { updateModel(t, c); });
    // a save needs no progress reply (the reference falls through into EVALUATE_PROGRESS, bcd.h:211-218)
    on(BCDCallView::SAVE_MODEL, [this](const MessagePtr&, const BCDCallView& c, int) {
      CHECK(c.has_data());
      saveModel(c.data());
    });
    on(BCDCallView::EVALUATE_PROGRESS, [this](const MessagePtr& m, const BCDCallView&, int) {
      BCDProgress p;
      evaluate(&p);
      answer(m, p);
    });
  }
  virtual ~BCDServer() { stopExecutor(); }

 protected:
  virtual void updateModel(int time, const BCDCallView& call) = 0;
  virtual void evaluate(BCDProgress* prog) = 0;

  // the server's half of PrepSchedule: wait for every worker, then let their pulls through
  virtual void preprocessData(int time, const BCDCallView& call) {
    useGroups(call);
    const PrepSchedule at{time, (int)fea_grp_.size()};
    if (!call.hit_cache())
      for (int i = 0; i < at.groups; ++i) {  // counts are in the sketch (SharedParameter::tailFilterStep)
        model_.waitInMsg(kWorkerGroup, at.count(i));
        model_.finish(kWorkerGroup, BcdClock::updated(at.count(i)));
      }
    for (int i = 0; i < at.groups; ++i) {  // the kept keys were joined into my key list
      model_.waitInMsg(kWorkerGroup, at.keys(i));
      const int chl = fea_grp_[i];
      model_.clearTailFilter(chl, i + 1 == at.groups);
      SArray<V>& w = model_.value(chl);
      w.resize(model_.key(chl).size());
      w.setValue(bcd_conf_.init_w());
      model_.finish(kWorkerGroup, BcdClock::updated(at.keys(i)));
    }
  }

  // ref bcd.h:251-272: "<key>\t<value>" per weight that is neither zero nor NaN, one file per server
  void saveModel(const DataConfigView& where) {
    if (where.format() != DataConfigView::TEXT) return;
    CHECK(where.file_size()) << "model_output has no file";
    const string file = where.file(0) + "_" + myNodeID();
    if (!dirExists(getPath(file))) createDir(getPath(file));
    std::ofstream out(file);
    CHECK(out.good()) << "cannot write " << file;
    for (int g : fea_grp_) {
      const SArray<Key> key = model_.key(g);
      const SArray<V> w = model_.value(g);
      CHECK_EQ(key.size(), w.size());
      for (size_t i = 0; i < key.size(); ++i)
        if (w[i] != 0 && w[i] == w[i]) out << key[i] << "\t" << (double)w[i] << "\n";
    }
    LI << myNodeID() << " written the model to " << file;
  }

  KVBufferedVector<Key, V> model_;
};

// ---------------------------------------------------------------------------------------------------
template <typename V>
class BCDWorker : public BCDNode {
 public:
  BCDWorker(const string& name, const BCDConfigView& conf) : BCDNode(name, conf), model_(name + "_model", name) {
    on(BCDCallView::LOAD_DATA, [this](const MessagePtr& m, const BCDCallView& c, int) {
      CHECK(c.has_data());
      LoadDataReturn ret;
      ExampleInfo info;
      loadData(c.data(), &info);
      ret.mutable_example_info().raw()->copyFrom(*info.raw());
      ret.set_hit_cache(0);  // a worker-level matrix cache is a stub in the reference too (bcd.h:467-505);
                             // the per-file slot cache of SlotReader does the heavy lifting
      answer(m, ret);
    });
    on(BCDCallView::PREPROCESS_DATA, [this](const MessagePtr&, const BCDCallView& c, int t) { preprocessData(t, c); });
    on(BCDCallView::UPDATE_MODEL, [this](const MessagePtr& m, const BCDCallView& c, int t) {
      computeGradient(t, c, m);
      m->finished = false;  // the block is done on this worker when its pull came back
    });
    on(BCDCallView::EVALUATE_PROGRESS, [this](const MessagePtr& m, const BCDCallView&, int) {
      BCDProgress p;
      evaluate(&p);
      answer(m, p);
    });
  }
  virtual ~BCDWorker() { stopExecutor(); }

 protected:
  virtual void computeGradient(int time, const BCDCallView& call, MessagePtr msg) = 0;
  virtual void evaluate(BCDProgress* prog) = 0;

  void loadData(const DataConfigView& data, ExampleInfo* info) {
    slot_reader_.init(data, bcd_conf_.local_cache());
    slot_reader_.read(info);
  }

  // the worker's half of PrepSchedule (ref bcd.h:324-465)
  virtual void preprocessData(int time, const BCDCallView& call) {
    useGroups(call);
    const PrepSchedule at{time, (int)fea_grp_.size()};
    const int window = std::max(1, bcd_conf_.max_num_parallel_groups_in_preprocessing());
    std::vector<int> filtered(at.groups, -1);  // time of the filtered-key pull of each group
    if (!call.hit_cache())
      for (int i = 0; i < at.groups; ++i) {
        filtered[i] = countAndFilter(fea_grp_[i], at.count(i));
        if (i >= window) model_.waitOutMsg(kServerGroup, filtered[i - window]);  // bounds the memory in flight
      }
    std::vector<std::promise<void>> ready(at.groups);
    for (int i = 0; i < at.groups; ++i) {
      if (filtered[i] >= 0 && i >= at.groups - window) model_.waitOutMsg(kServerGroup, filtered[i]);
      adoptKeysAndFetchWeights(fea_grp_[i], at.keys(i), &ready[i]);
    }
    if (!call.hit_cache()) y_ = slot_reader_.value<V>(0);
    for (auto& r : ready) r.get_future().wait();
  }

  // round 1 of a group: unique keys with counts go to the servers' sketch; the keys that are
  // frequent enough come back and the group is localised against them.  Returns the pull's time.
  int countAndFilter(int grp, int t) {
    auto localizer = std::make_shared<Localizer<Key, V>>();
    const SArray<Key> index = slot_reader_.index(grp);
    SArray<Key> uniq;
    SArray<uint8> count;
    localizer->countUniqIndex(index, &uniq, &count);

    MessagePtr tell(new Message(kServerGroup, t));
    tell->setKey(uniq);
    tell->addValue(count);
    tell->task.set_key_channel(grp);
    tell->addFilter(FilterConfigView::KEY_CACHING);
    TailKeyFilterView sketch = model_.set(tell).mutable_tail_filter();
    sketch.set_insert_count(true);
    sketch.set_countmin_k(bcd_conf_.countmin_k());
    sketch.set_countmin_n((int)(uniq.size() * bcd_conf_.countmin_n_ratio()));
    CHECK_EQ(t, model_.push(tell));

    MessagePtr ask(new Message(kServerGroup, BcdClock::pull(t), BcdClock::updated(t)));
    ask->setKey(uniq);
    ask->task.set_key_channel(grp);
    ask->addFilter(FilterConfigView::KEY_CACHING);
    model_.set(ask).mutable_tail_filter().set_query_key(bcd_conf_.tail_feature_freq());
    ask->fin_handle = [this, grp, localizer, index]() {
      // global keys -> positions in the kept key list (now model_.key(grp)), then by column
      MatrixPtr<V> X = localizer->remapIndex(slot_reader_.info<V>(grp), slot_reader_.offset(grp), index,
                                             slot_reader_.value<V>(grp), model_.key(grp));
      slot_reader_.clear(grp);
      if (!X) return;
      if (bcd_conf_.feature_block_ratio()) X = X->toColMajor();
      Lock l(mu_);
      X_[grp] = X;
    };
    CHECK_EQ(BcdClock::pull(t), model_.pull(ask));
    return BcdClock::pull(t);
  }

  // round 2 of a group: the servers learn my keys, then hand out the initial weights; dual_
  // starts as X w summed over the groups (the application turns it into what it needs)
  void adoptKeysAndFetchWeights(int grp, int t, std::promise<void>* done) {
    MessagePtr tell(new Message(kServerGroup, t));
    tell->setKey(model_.key(grp));
    tell->task.set_key_channel(grp);
    tell->addFilter(FilterConfigView::KEY_CACHING);
    CHECK_EQ(t, model_.push(tell));

    const int tp = BcdClock::pull(t);
    MessagePtr ask(new Message(kServerGroup, tp, BcdClock::updated(t)));
    ask->setKey(model_.key(grp));
    ask->task.set_key_channel(grp);
    ask->addFilter(FilterConfigView::KEY_CACHING).set_clear_cache_if_done(true);
    ask->fin_handle = [this, grp, tp, done]() {
      if (const size_t n = model_.key(grp).size()) {
        auto w0 = model_.receivedOnHost(tp);
        CHECK_EQ(w0.second.size(), (size_t)1);
        CHECK_EQ(w0.first.size(), n);
        model_.value(grp) = w0.second[0];
        MatrixPtr<V> X;
        { Lock l(mu_); X = X_[grp]; }
        CHECK(X) << "group " << grp << " has no localized matrix";
        if (dual_.empty()) dual_ = SArray<V>(X->rows(), 0);
        CHECK_EQ(dual_.size(), X->rows());
        if (bcd_conf_.init_w().type() != ParameterInitConfigView::ZERO) {
          const SArray<V> xw = X->times(model_.value(grp));
          for (size_t r = 0; r < xw.size(); ++r) dual_[r] += xw[r];
        }
      }
      done->set_value();
    };
    CHECK_EQ(tp, model_.pull(ask));
  }

  SlotReader slot_reader_;
  std::map<int, MatrixPtr<V>> X_;  // slot id -> column-major localized matrix
  SArray<V> y_;
  SArray<V> dual_;
  std::mutex mu_;
  KVBufferedVector<Key, V> model_;
};

}  // namespace PS
