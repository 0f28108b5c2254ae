// Block coordinate descent: scheduler / server / worker skeletons (CPU plumbing path).
// ref src/learner/bcd.h:10-518, src/learner/proto/bcd.proto
// The GPU path implements the same roles in parameter_server_b200/models/darlin.py on top of
// csrc/cuda/darlin.cu; this file is the host version that runs over the van (and doubles as
// the end-to-end oracle of the protocol: timestamps 10t / 10t+1 / 10t+2 per scheduler time t).
#pragma once
#include "data/slot_reader.h"
#include "parameter/kv.h"
#include "util/localizer.h"

namespace PS {

class BCDCommon {
 public:
  explicit BCDCommon(const BCDConfigView& conf) : bcd_conf_(conf) {
    if (!bcd_conf_.has_local_cache()) bcd_conf_.mutable_local_cache().add_file("/tmp/bcd_");
  }
  virtual ~BCDCommon() {}

 protected:
  std::vector<int> fea_grp_;
  const int time_ratio_ = 10;
  BCDConfig bcd_conf_;
};

// ---------------------------------------------------------------------------------------------------
class BCDScheduler : public App, public BCDCommon {
 public:
  BCDScheduler(const string& name, const BCDConfigView& conf) : App(name), BCDCommon(conf) {}
  virtual ~BCDScheduler() {}

  // LOAD_DATA to the workers (files divided round robin), then PREPROCESS_DATA to all
  void loadTrainingData(const DataConfigView& train) {
    auto load_time = tic();
    auto confs = Postmaster::partitionData(train, sys_.yp().num_workers());
    std::vector<Task> loads(confs.size());
    for (size_t i = 0; i < confs.size(); ++i) {
      BCDCallView bcd = loads[i].mutable_bcd();
      bcd.set_cmd(BCDCallView::LOAD_DATA);
      bcd.mutable_data().raw()->copyFrom(*confs[i].raw());
    }
    int hit_cache = 0;
    port(kWorkerGroup)->submitAndWait(loads, [this, &hit_cache]() {
      LoadDataReturn info;
      CHECK(info.ParseFromString(exec().lastRecvReply()));
      Lock l(progress_mu_);
      g_train_info_ = mergeExampleInfo(g_train_info_, info.example_info());
      hit_cache += info.hit_cache();
    });
    if (hit_cache > 0) {
      CHECK_EQ(hit_cache, (int)loads.size()) << "clear the local caches";
      LI << "Hit local caches for the training data";
    }
    LI << "Loaded " << g_train_info_.num_ex() << " examples in " << toc(load_time) << " sec";

    for (int i = 0; i < g_train_info_.slot_size(); ++i) {
      SlotInfoView info = g_train_info_.slot(i);
      if (info.id() == 0) continue;  // the label
      fea_grp_.push_back(info.id());
    }
    auto prep_time = tic();
    Task preprocess;
    BCDCallView prep = preprocess.mutable_bcd();
    prep.set_cmd(BCDCallView::PREPROCESS_DATA);
    for (int grp : fea_grp_) prep.add_fea_grp(grp);
    prep.set_hit_cache(hit_cache > 0);
    port(kCompGroup)->submitAndWait(preprocess);
    LI << "Preprocessing is finished in " << toc(prep_time) << " sec";
    if (bcd_conf_.tail_feature_freq())
      LI << "Features with frequency <= " << bcd_conf_.tail_feature_freq() << " are filtered";
  }

  // every feature group is cut into ceil(nnz_per_row * feature_block_ratio) equal key ranges
  void divideFeatureBlocks() {
    for (int i = 0; i < g_train_info_.slot_size(); ++i) {
      SlotInfoView info = g_train_info_.slot(i);
      if (info.id() == 0) continue;
      CHECK(info.has_nnz_ele());
      CHECK(info.has_nnz_ex());
      double nnz_per_row = (double)info.nnz_ele() / (double)info.nnz_ex();
      int n = 1;
      if (nnz_per_row > 1 + 1e-6)  // only groups whose features are correlated are split
        n = std::max((int)std::ceil(nnz_per_row * bcd_conf_.feature_block_ratio()), 1);
      for (int k = 0; k < n; ++k) {
        auto block = Range<Key>(info.min_key(), info.max_key()).evenDivide(n, k);
        if (block.empty()) continue;
        fea_blk_.push_back(std::make_pair(info.id(), block));
      }
    }
    LI << "Features are partitioned into " << fea_blk_.size() << " blocks";
    for (size_t i = 0; i < fea_blk_.size(); ++i) blk_order_.push_back((int)i);
    std::vector<string> hit_blk;
    for (int i = 0; i < bcd_conf_.prior_fea_group_size(); ++i) {
      int grp_id = bcd_conf_.prior_fea_group(i);
      std::vector<int> tmp;
      for (size_t k = 0; k < fea_blk_.size(); ++k)
        if (fea_blk_[k].first == grp_id) tmp.push_back((int)k);
      if (tmp.empty()) continue;
      hit_blk.push_back(to_string(grp_id));
      for (int j = 0; j < bcd_conf_.num_iter_for_prior_fea_group(); ++j) {
        if (bcd_conf_.random_feature_block_order()) std::shuffle(tmp.begin(), tmp.end(), rng_);
        prior_blk_order_.insert(prior_blk_order_.end(), tmp.begin(), tmp.end());
      }
    }
    if (!hit_blk.empty()) LI << "Prior feature groups: " << join(hit_blk, ", ");
    total_timer_.restart();
  }

  void saveModel(const DataConfigView& data) {
    Task task;
    task.mutable_bcd().set_cmd(BCDCallView::SAVE_MODEL);
    task.mutable_bcd().mutable_data().raw()->copyFrom(*data.raw());
    port(kServerGroup)->submitAndWait(task);
  }
  const std::map<int, BCDProgress>& progress() const { return g_progress_; }

 protected:
  void mergeProgress(int iter) {
    BCDProgress recv;
    CHECK(recv.ParseFromString(exec().lastRecvReply()));
    Lock l(progress_mu_);
    BCDProgress& p = g_progress_[iter];
    p.set_objective(p.objective() + recv.objective());
    p.set_nnz_w(p.nnz_w() + recv.nnz_w());
    if (recv.busy_time_size() > 0) p.add_busy_time(recv.busy_time(0));
    p.set_total_time(total_timer_.stop());
    total_timer_.start();
    p.set_relative_obj(iter == 0 ? 1 : g_progress_[iter - 1].objective() / p.objective() - 1);
    p.set_violation(std::max(p.violation(), recv.violation()));
    p.set_nnz_active_set(p.nnz_active_set() + recv.nnz_active_set());
  }
  void showTime(int iter) {
    if (iter == -3) fprintf(stderr, "|    time (sec.)\n");
    else if (iter == -2) fprintf(stderr, "|(app:min max) total\n");
    else if (iter == -1) fprintf(stderr, "+-----------------\n");
    else {
      BCDProgress& prog = g_progress_[iter];
      double ttl_t = prog.total_time() - (iter > 0 ? g_progress_[iter - 1].total_time() : 0);
      double mn = 1e300, mx = 0;
      for (int i = 0; i < prog.busy_time_size(); ++i) {
        mn = std::min(mn, prog.busy_time(i));
        mx = std::max(mx, prog.busy_time(i));
      }
      if (prog.busy_time_size() == 0) mn = 0;
      fprintf(stderr, "|%6.1f%6.1f%6.1f\n", mn, mx, ttl_t);
    }
  }
  void showObjective(int iter) {
    if (iter == -3) fprintf(stderr, "     |        training        |  sparsity ");
    else if (iter == -2) fprintf(stderr, "iter |  objective    relative |     |w|_0 ");
    else if (iter == -1) fprintf(stderr, " ----+------------------------+-----------");
    else {
      BCDProgress& prog = g_progress_[iter];
      fprintf(stderr, "%4d | %.5e  %.3e |%10lu ", iter, prog.objective(), prog.relative_obj(),
              (size_t)prog.nnz_w());
    }
  }

  std::mutex progress_mu_;
  std::map<int, BCDProgress> g_progress_;
  Timer total_timer_;
  typedef std::vector<std::pair<int, Range<Key>>> FeatureBlocks;
  FeatureBlocks fea_blk_;
  std::vector<int> blk_order_, prior_blk_order_;
  ExampleInfo g_train_info_;
  std::mt19937 rng_{20140624};
};

// ---------------------------------------------------------------------------------------------------
template <typename V>
class BCDServer : public App, public BCDCommon {
 public:
  BCDServer(const string& name, const BCDConfigView& conf)
      : App(name), BCDCommon(conf), model_(name + "_model", name) {}
  virtual ~BCDServer() { stopExecutor(); }

  void process(const MessagePtr& msg) override {
    CHECK(msg->task.has_bcd());
    BCDCallView bcd = msg->task.bcd();
    int time = msg->task.time() * time_ratio_;
    switch (bcd.cmd()) {
      case BCDCallView::PREPROCESS_DATA: preprocessData(time, bcd); break;
      case BCDCallView::UPDATE_MODEL: updateModel(time, bcd); break;
      case BCDCallView::SAVE_MODEL:
        // the reference falls through into EVALUATE_PROGRESS here (bcd.h:211-218); a save does
        // not need a progress reply, so it ends here
        CHECK(bcd.has_data());
        saveModel(bcd.data());
        break;
      case BCDCallView::EVALUATE_PROGRESS: {
        BCDProgress prog;
        evaluate(&prog);
        sys_.replyProtocalMessage(msg, prog);
        break;
      }
      default: break;
    }
  }

 protected:
  virtual void updateModel(int time, const BCDCallView& call) = 0;
  virtual void evaluate(BCDProgress* prog) = 0;

  virtual void preprocessData(int time, const BCDCallView& call) {
    int grp_size = call.fea_grp_size();
    fea_grp_.clear();
    for (int i = 0; i < grp_size; ++i) fea_grp_.push_back(call.fea_grp(i));
    bool hit_cache = call.hit_cache();
    // round 1: key counts arrive (tail filter), then the filtered-key pulls are released
    for (int i = 0; i < grp_size; ++i, time += time_ratio_) {
      if (hit_cache) continue;
      model_.waitInMsg(kWorkerGroup, time);
      model_.finish(kWorkerGroup, time + 1);
    }
    // round 2: the kept keys arrive, weights are initialised, the initial pulls are released
    for (int i = 0; i < grp_size; ++i, time += time_ratio_) {
      model_.waitInMsg(kWorkerGroup, time);
      int chl = fea_grp_[i];
      model_.clearTailFilter(chl);
      model_.value(chl).resize(model_.key(chl).size());
      model_.value(chl).setValue(bcd_conf_.init_w());
      model_.finish(kWorkerGroup, time + 1);
    }
  }

  // "key\tvalue" per non-zero, non-NaN weight into <file>_<node id>
  void saveModel(const DataConfigView& output) {
    if (output.format() != DataConfigView::TEXT) return;
    CHECK(output.file_size());
    string file = output.file(0) + "_" + myNodeID();
    if (!dirExists(getPath(file))) createDir(getPath(file));
    std::ofstream out(file);
    CHECK(out.good()) << "cannot write " << file;
    for (int grp : fea_grp_) {
      SArray<Key> key = model_.key(grp);
      SArray<V> value = model_.value(grp);
      CHECK_EQ(key.size(), value.size());
      for (size_t i = 0; i < key.size(); ++i) {
        double v = value[i];
        if (v != 0 && !(v != v)) out << key[i] << "\t" << v << "\n";
      }
    }
    LI << myNodeID() << " written the model to " << file;
  }

  KVBufferedVector<Key, V> model_;
};

// ---------------------------------------------------------------------------------------------------
template <typename V>
class BCDWorker : public App, public BCDCommon {
 public:
  BCDWorker(const string& name, const BCDConfigView& conf)
      : App(name), BCDCommon(conf), model_(name + "_model", name) {}
  virtual ~BCDWorker() { stopExecutor(); }

  void process(const MessagePtr& msg) override {
    CHECK(msg->task.has_bcd());
    BCDCallView bcd = msg->task.bcd();
    int time = msg->task.time() * time_ratio_;
    switch (bcd.cmd()) {
      case BCDCallView::LOAD_DATA: {
        LoadDataReturn ret;
        int hit_cache = 0;
        CHECK(bcd.has_data());
        ExampleInfo info;
        loadData(bcd.data(), &info, &hit_cache);
        ret.mutable_example_info().raw()->copyFrom(*info.raw());
        ret.set_hit_cache(hit_cache);
        sys_.replyProtocalMessage(msg, ret);
        break;
      }
      case BCDCallView::PREPROCESS_DATA: preprocessData(time, bcd); break;
      case BCDCallView::UPDATE_MODEL:
        computeGradient(time, bcd, msg);
        msg->finished = false;  // finished by the pull callback
        break;
      case BCDCallView::EVALUATE_PROGRESS: {
        BCDProgress prog;
        evaluate(&prog);
        sys_.replyProtocalMessage(msg, prog);
        break;
      }
      default: break;
    }
  }

 protected:
  virtual void computeGradient(int time, const BCDCallView& bcd, MessagePtr msg) = 0;
  virtual void evaluate(BCDProgress* prog) = 0;

  void loadData(const DataConfigView& data, ExampleInfo* info, int* hit_cache) {
    *hit_cache = 0;  // a worker-level matrix cache is a stub in the reference too (bcd.h:467-505);
                     // the per-file slot cache of SlotReader does the heavy lifting
    slot_reader_.init(data, bcd_conf_.local_cache());
    slot_reader_.read(info);
  }

  virtual void preprocessData(int time, const BCDCallView& call) {
    int grp_size = call.fea_grp_size();
    fea_grp_.clear();
    for (int i = 0; i < grp_size; ++i) fea_grp_.push_back(call.fea_grp(i));
    bool hit_cache = call.hit_cache();
    int max_parallel = std::max(1, bcd_conf_.max_num_parallel_groups_in_preprocessing());
    std::vector<int> pull_time(grp_size);

    // ---- round 1: count unique keys, push counts, pull the keys that pass the tail filter ----
    for (int i = 0; i < grp_size; ++i, time += time_ratio_) {
      if (hit_cache) continue;
      int grp = fea_grp_[i];
      SArray<Key> uniq_key;
      SArray<uint8> key_cnt;
      auto localizer = std::make_shared<Localizer<Key, V>>();
      SArray<Key> index = slot_reader_.index(grp);
      localizer->countUniqIndex(index, &uniq_key, &key_cnt);

      MessagePtr count(new Message(kServerGroup, time));
      count->setKey(uniq_key);
      count->addValue(key_cnt);
      count->task.set_key_channel(grp);
      count->addFilter(FilterConfigView::KEY_CACHING);
      TailKeyFilterView tail = model_.set(count).mutable_tail_filter();
      tail.set_insert_count(true);
      tail.set_countmin_k(bcd_conf_.countmin_k());
      tail.set_countmin_n((int)(uniq_key.size() * bcd_conf_.countmin_n_ratio()));
      CHECK_EQ(time, model_.push(count));

      MessagePtr filter(new Message(kServerGroup, time + 2, time + 1));
      filter->setKey(uniq_key);
      filter->task.set_key_channel(grp);
      filter->addFilter(FilterConfigView::KEY_CACHING);
      model_.set(filter).mutable_tail_filter().set_query_key(bcd_conf_.tail_feature_freq());
      filter->fin_handle = [this, grp, localizer, index]() {
        // localize: global keys -> positions in the kept key list, then column-major
        SArray<size_t> offset = slot_reader_.offset(grp);
        SArray<V> value = slot_reader_.value<V>(grp);
        MatrixInfo info = slot_reader_.info<V>(grp);
        MatrixPtr<V> X = localizer->remapIndex(info, offset, index, value, model_.key(grp));
        slot_reader_.clear(grp);
        if (!X) return;
        if (bcd_conf_.feature_block_ratio()) X = X->toColMajor();
        Lock l(mu_);
        X_[grp] = X;
      };
      CHECK_EQ(time + 2, model_.pull(filter));
      pull_time[i] = time + 2;
      if (i >= max_parallel) model_.waitOutMsg(kServerGroup, pull_time[i - max_parallel]);
    }

    // ---- round 2: push the kept keys (servers build their key lists), pull the initial weights
    std::vector<std::promise<void>> wait_dual(grp_size);
    for (int i = 0; i < grp_size; ++i, time += time_ratio_) {
      if (!hit_cache && i >= grp_size - max_parallel) model_.waitOutMsg(kServerGroup, pull_time[i]);
      int grp = fea_grp_[i];
      MessagePtr push_key(new Message(kServerGroup, time));
      push_key->setKey(model_.key(grp));
      push_key->task.set_key_channel(grp);
      push_key->addFilter(FilterConfigView::KEY_CACHING);
      CHECK_EQ(time, model_.push(push_key));

      MessagePtr pull_val(new Message(kServerGroup, time + 2, time + 1));
      pull_val->setKey(model_.key(grp));
      pull_val->task.set_key_channel(grp);
      pull_val->addFilter(FilterConfigView::KEY_CACHING).set_clear_cache_if_done(true);
      pull_val->fin_handle = [this, grp, time, &wait_dual, i]() {
        size_t n = model_.key(grp).size();
        if (n) {
          auto init_w = model_.received(time + 2);
          CHECK_EQ(init_w.second.size(), (size_t)1);
          CHECK_EQ(n, init_w.first.size());
          model_.value(grp) = init_w.second[0];
          MatrixPtr<V> X;
          {
            Lock l(mu_);
            X = X_[grp];
          }
          CHECK(X) << "group " << grp << " has no localized matrix";
          if (dual_.empty()) {
            dual_.resize(X->rows());
            dual_.setZero();
          } else {
            CHECK_EQ(dual_.size(), X->rows());
          }
          if (bcd_conf_.init_w().type() != ParameterInitConfigView::ZERO) {
            SArray<V> xw = X->times(model_.value(grp));  // dual_ accumulates X*w over the groups
            for (size_t r = 0; r < xw.size(); ++r) dual_[r] += xw[r];
          }
        }
        wait_dual[i].set_value();
      };
      CHECK_EQ(time + 2, model_.pull(pull_val));
      pull_time[i] = time + 2;
    }
    if (!hit_cache) {
      y_ = slot_reader_.value<V>(0);
    }
    for (int i = 0; i < grp_size; ++i) wait_dual[i].get_future().wait();
  }

  SlotReader slot_reader_;
  std::map<int, MatrixPtr<V>> X_;  // slot id -> column-major localized matrix
  SArray<V> y_;
  SArray<V> dual_;                 // exp(y .* (X w)) once the app initialised it
  std::mutex mu_;
  KVBufferedVector<Key, V> model_;
};

}  // namespace PS
