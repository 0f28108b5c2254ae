// Darlin: delayed block proximal gradient for L1 logistic regression — host (CPU) roles.
// ref src/app/linear_method/darlin.h:12-514
// The same algorithm on the device: csrc/cuda/darlin.cu (kernels) driven by
// parameter_server_b200/models/darlin.py; tests compare both against one oracle.
#pragma once
#include "app/linear_method/loss.h"
#include "learner/bcd.h"
#include "util/sketch.h"

namespace PS {
namespace LM {
typedef double Real;

class LinearMethod {
 public:
  explicit LinearMethod(const ConfigView& conf) : conf_(conf) {}
  virtual ~LinearMethod() {}

 protected:
  Config conf_;
};

class DarlinCommon {
 protected:
  static Real newDelta(Real delta_max, Real delta_w) {
    return std::min(delta_max, 2 * fabs(delta_w) + .1);
  }
  std::unordered_map<int, Bitmap> active_set_;
  std::unordered_map<int, SArray<Real>> delta_;
  SparseFilter kkt_filter_;
  Real kkt_filter_threshold_ = 1e20;
};

// ---------------------------------------------------------------------------------------------------
class DarlinScheduler : public BCDScheduler, DarlinCommon, LinearMethod {
 public:
  DarlinScheduler(const string& name, const ConfigView& conf)
      : BCDScheduler(name, conf.darlin()), LinearMethod(conf) {}
  virtual ~DarlinScheduler() { stopExecutor(); }

  void run() override {
    CHECK_EQ((int)conf_.loss().type(), (int)LossConfigView::LOGIT);
    CHECK_EQ((int)conf_.penalty().type(), (int)PenaltyConfigView::L1);
    LI << "Train l_1 logistic regression by block coordinate descent";
    loadTrainingData(conf_.training_data());
    divideFeatureBlocks();

    BCDConfigView darlin = conf_.darlin();
    const int tau = darlin.max_block_delay();
    LI << "Maximal allowed delay: " << tau;
    const bool random_blk_order = darlin.random_feature_block_order();
    if (!random_blk_order) LI << "Warning: Randomized block order often acclerates the convergence.";
    kkt_filter_threshold_ = 1e20;
    bool reset_kkt_filter = false;

    const int max_iter = darlin.max_pass_of_data();
    RNodePtr pool = port(kCompGroup);
    // a timestamp large enough not to collide with the preprocessing stage
    int time = std::max(10000, pool->time() + (int)fea_grp_.size() * time_ratio_);
    const int first_time = time + 1;
    for (int iter = 0; iter < max_iter; ++iter) {
      std::vector<int> order = blk_order_;
      if (random_blk_order) std::shuffle(order.begin(), order.end(), rng_);
      if (iter == 0) order.insert(order.begin(), prior_blk_order_.begin(), prior_blk_order_.end());

      for (size_t i = 0; i < order.size(); ++i) {
        Task update;
        BCDCallView cmd = update.mutable_bcd();
        cmd.set_cmd(BCDCallView::UPDATE_MODEL);
        if (i == 0) {  // the KKT filter threshold travels with the first block of a pass
          cmd.set_kkt_filter_threshold(kkt_filter_threshold_);
          if (reset_kkt_filter) cmd.set_reset_kkt_filter(true);
        }
        const auto& blk = fea_blk_[order[i]];
        cmd.mutable_key().from(blk.second);
        cmd.add_fea_grp(blk.first);
        update.set_time(time + 1);
        // bounded delay: block t+1 may start once block t-tau has finished on that node
        int wait_time = time - tau;
        if (iter == 0 && i < prior_blk_order_.size()) wait_time = time;  // no delay for priors
        if (wait_time < first_time) wait_time = Message::kInvalidTime;
        update.add_wait_time(wait_time);
        time = pool->submit(update);
      }

      Task eval;
      eval.mutable_bcd().set_cmd(BCDCallView::EVALUATE_PROGRESS);
      if (time - tau >= first_time) eval.add_wait_time(time - tau);
      time = pool->submitAndWait(eval, [this, iter]() { mergeProgress(iter); });
      showProgress(iter);

      Real vio = g_progress_[iter].violation();
      Real ratio = bcd_conf_.kkt_filter_threshold_ratio();
      kkt_filter_threshold_ = vio / (Real)g_train_info_.num_ex() * ratio;

      Real rel = g_progress_[iter].relative_obj();
      if (rel > 0 && rel <= darlin.epsilon()) {
        if (reset_kkt_filter) {
          LI << "Stopped: relative objective <= " << darlin.epsilon();
          break;
        }
        reset_kkt_filter = true;  // one more pass over all coordinates before stopping
      } else {
        reset_kkt_filter = false;
      }
      if (iter == max_iter - 1) LI << "Reached maximal " << max_iter << " data passes";
    }
    if (conf_.has_model_output()) saveModel(conf_.model_output());
  }

 protected:
  void showKKTFilter(int iter) {
    if (iter == -3) fprintf(stderr, "|      KKT filter     ");
    else if (iter == -2) fprintf(stderr, "| threshold  #activet ");
    else if (iter == -1) fprintf(stderr, "+---------------------");
    else fprintf(stderr, "| %.1e %11llu ", kkt_filter_threshold_,
                 (unsigned long long)g_progress_[iter].nnz_active_set());
  }
  void showProgress(int iter) {
    int s = iter == 0 ? -3 : iter;
    for (int i = s; i <= iter; ++i) {
      showObjective(i);
      showKKTFilter(i);
      showTime(i);
    }
  }
};

// ---------------------------------------------------------------------------------------------------
class DarlinServer : public BCDServer<Real>, DarlinCommon, LinearMethod {
 public:
  DarlinServer(const string& name, const ConfigView& conf)
      : BCDServer<Real>(name, conf.darlin()), LinearMethod(conf) {}
  virtual ~DarlinServer() { stopExecutor(); }

 protected:
  void preprocessData(int time, const BCDCallView& call) override {
    BCDServer<Real>::preprocessData(time, call);
    for (int grp : fea_grp_) {
      size_t n = model_.key(grp).size();
      active_set_[grp].resize(n, true);
      delta_[grp].resize(n);
      delta_[grp].setValue(bcd_conf_.delta_init_value());
    }
  }

  void updateModel(int time, const BCDCallView& call) override {
    if (call.has_kkt_filter_threshold()) {
      kkt_filter_threshold_ = call.kkt_filter_threshold();
      violation_ = 0;
    }
    if (call.reset_kkt_filter())
      for (int grp : fea_grp_) active_set_[grp].fill(true);
    CHECK_EQ(call.fea_grp_size(), 1);
    int grp = call.fea_grp(0);
    Range<Key> g_key_range = call.key().to<Key>();
    SizeR col_range = model_.find(grp, g_key_range);
    // none of my business
    if (myKeyRange().setIntersection(g_key_range).empty()) return;
    // every worker's gradient for this block has been merged into the buffer
    model_.waitInMsg(kWorkerGroup, time);
    if (!col_range.empty()) {
      auto data = model_.received(time);
      CHECK_EQ(col_range, data.first);
      CHECK_EQ(data.second.size(), (size_t)2);
      updateWeight(grp, col_range, data.second[0], data.second[1]);
    }
    model_.finish(kWorkerGroup, time + 1);  // releases the workers' pulls
  }

  // proximal Newton step per coordinate with a trust region and the KKT filter
  void updateWeight(int grp, SizeR range, SArray<Real> G, SArray<Real> U) {
    CHECK_EQ(G.size(), range.size());
    CHECK_EQ(U.size(), range.size());
    const Real eta = conf_.learning_rate().alpha();
    const Real lambda = conf_.penalty().lambda(0);
    const Real delta_max = bcd_conf_.delta_max_value();
    SArray<Real>& value = model_.value(grp);
    Bitmap& active = active_set_[grp];
    SArray<Real>& delta = delta_[grp];
    for (size_t i = 0; i < range.size(); ++i) {
      const size_t k = i + range.begin();
      if (!active.test(k)) continue;
      const Real g = G[i], u = U[i] / eta + 1e-10;
      const Real g_pos = g + lambda, g_neg = g - lambda;
      Real& w = value[k];
      Real d = -w, vio = 0;
      if (w == 0) {
        if (g_pos < 0) {
          vio = -g_pos;
        } else if (g_neg > 0) {
          vio = g_neg;
        } else if (g_pos > kkt_filter_threshold_ && g_neg < -kkt_filter_threshold_) {
          active.clear(k);
          kkt_filter_.mark(&w);
          continue;
        }
      }
      violation_ = std::max(violation_, vio);
      if (g_pos <= u * w) d = -g_pos / u;
      else if (g_neg >= u * w) d = -g_neg / u;
      d = std::min(delta[k], std::max(-delta[k], d));
      delta[k] = newDelta(delta_max, d);
      w += d;
    }
  }

  void evaluate(BCDProgress* prog) override {
    size_t nnz_w = 0, nnz_as = 0;
    Real objv = 0;
    for (int grp : fea_grp_) {
      const SArray<Real>& value = model_.value(grp);
      for (size_t i = 0; i < value.size(); ++i) {
        Real w = value[i];
        if (kkt_filter_.marked(w) || w == 0) continue;
        ++nnz_w;
        objv += fabs(w);
      }
      nnz_as += active_set_[grp].nnz();
    }
    prog->set_objective(objv * conf_.penalty().lambda(0));
    prog->set_nnz_w(nnz_w);
    prog->set_violation(violation_);
    prog->set_nnz_active_set(nnz_as);
  }

  Real violation_ = 0;
};

// ---------------------------------------------------------------------------------------------------
class DarlinWorker : public BCDWorker<Real>, DarlinCommon, LinearMethod {
 public:
  DarlinWorker(const string& name, const ConfigView& conf)
      : BCDWorker<Real>(name, conf.darlin()), LinearMethod(conf) {}
  virtual ~DarlinWorker() { stopExecutor(); }

 protected:
  void preprocessData(int time, const BCDCallView& call) override {
    BCDWorker<Real>::preprocessData(time, call);
    // dual_ = exp(y .* (X w)); with zero initial weights that is all ones
    if (bcd_conf_.init_w().type() == ParameterInitConfigView::ZERO) {
      dual_.setValue(1);
    } else {
      for (size_t i = 0; i < dual_.size(); ++i) dual_[i] = exp(y_[i] * dual_[i]);
    }
    for (int grp : fea_grp_) {
      size_t n = model_.key(grp).size();
      active_set_[grp].resize(n, true);
      delta_[grp].resize(n);
      delta_[grp].setValue(bcd_conf_.delta_init_value());
    }
  }

  void computeGradient(int time, const BCDCallView& call, MessagePtr msg) override {
    if (call.reset_kkt_filter())
      for (int grp : fea_grp_) active_set_[grp].fill(true);
    CHECK_EQ(call.fea_grp_size(), 1);
    int grp = call.fea_grp(0);
    Range<Key> g_key_range = call.key().to<Key>();
    SizeR col_range = model_.find(grp, g_key_range);
    computeAndPushGradient(time, g_key_range, grp, col_range);
    pullAndUpdateDual(time + 2, g_key_range, grp, col_range, msg);
  }

  void computeAndPushGradient(int time, Range<Key> g_key_range, int grp, SizeR col_range) {
    SArray<Real> G(col_range.size(), 0), U(col_range.size(), 0);
    {
      Lock l(mu_);  // dual_ is shared with the pull callback (model executor thread)
      busy_timer_.start();
      if (!col_range.empty()) {
        int nt = col_range.size() < 64 ? 1 : std::max(1, (int)FLAGS_num_threads);
        parallelFor(nt, nt, [&](int t) {
          SizeR thr = col_range.evenDivide(nt, t);
          if (!thr.empty()) gradientOfColumns(grp, thr, G.data() - col_range.begin(), U.data() - col_range.begin());
        });
      }
      busy_timer_.stop();
    }
    MessagePtr push(new Message(kServerGroup, time));
    push->setKey(model_.key(grp).segment(col_range));
    push->addValue({G, U});
    push->task.mutable_key_range().from(g_key_range);
    push->task.set_key_channel(grp);
    push->addFilter(FilterConfigView::KEY_CACHING);
    CHECK_EQ(time, model_.push(push));
  }

  // g_j = -sum_i y_i tau_i x_ij,  u_j = sum_i min(tau_i (1-tau_i) e^{|x_ij| delta_j}, 1/4) x_ij^2
  void gradientOfColumns(int grp, SizeR cols, Real* G, Real* U) {
    auto X = std::static_pointer_cast<SparseMatrix<uint32, Real>>(X_[grp]);
    CHECK(X->colMajor());
    const Bitmap& active = active_set_[grp];
    const SArray<Real>& delta = delta_[grp];
    const size_t* offset = X->offset().data();
    const uint32* index = X->index().data();
    const Real* value = X->value().data();
    const bool binary = X->binary();
    const Real* y = y_.data();
    for (size_t j = cols.begin(); j < cols.end(); ++j) {
      if (!active.test(j)) {
        kkt_filter_.mark(G + j);
        kkt_filter_.mark(U + j);
        continue;
      }
      Real g = 0, u = 0;
      const Real d = binary ? exp(delta[j]) : delta[j];
      for (size_t o = offset[j]; o < offset[j + 1]; ++o) {
        const uint32 i = index[o];
        const Real tau = 1 / (1 + dual_[i]);
        if (binary) {
          g -= y[i] * tau;
          u += std::min(tau * (1 - tau) * d, .25);
        } else {
          const Real v = value[o];
          g -= y[i] * tau * v;
          u += std::min(tau * (1 - tau) * exp(fabs(v) * d), .25) * v * v;
        }
      }
      G[j] = g;
      U[j] = u;
    }
  }

  void pullAndUpdateDual(int time, Range<Key> g_key_range, int grp, SizeR col_range,
                         const MessagePtr& msg) {
    MessagePtr pull(new Message(kServerGroup, time, time - 1));
    pull->setKey(model_.key(grp).segment(col_range));
    pull->task.mutable_key_range().from(g_key_range);
    pull->task.set_key_channel(grp);
    pull->addFilter(FilterConfigView::KEY_CACHING);
    pull->fin_handle = [this, time, grp, col_range, msg]() {
      if (!col_range.empty()) {
        auto data = model_.received(time);
        CHECK_EQ(col_range, data.first);
        CHECK_EQ(data.second.size(), (size_t)1);
        updateDual(grp, col_range, data.second[0]);
      }
      // only now is the scheduler's UPDATE_MODEL task finished on this worker
      port(msg->sender)->finishIncomingTask(msg->task.time());
      sys_.reply(msg->sender, msg->task);
    };
    CHECK_EQ(time, model_.pull(pull));
  }

  void updateDual(int grp, SizeR col_range, SArray<Real> new_w) {
    SArray<Real>& cur_w = model_.value(grp);
    Bitmap& active = active_set_[grp];
    SArray<Real>& delta = delta_[grp];
    const Real delta_max = bcd_conf_.delta_max_value();
    SArray<Real> delta_w(new_w.size());
    for (size_t i = 0; i < new_w.size(); ++i) {
      const size_t j = col_range.begin() + i;
      if (kkt_filter_.marked(new_w[i])) {
        active.clear(j);
        cur_w[j] = 0;
        delta_w[i] = 0;
        continue;
      }
      delta_w[i] = new_w[i] - cur_w[j];
      delta[j] = newDelta(delta_max, delta_w[i]);
      cur_w[j] = new_w[i];
    }
    auto X = std::static_pointer_cast<SparseMatrix<uint32, Real>>(X_[grp]);
    CHECK(X && X->colMajor());
    const size_t* offset = X->offset().data();
    const uint32* index = X->index().data();
    const Real* value = X->value().data();
    const bool binary = X->binary();
    const Real* y = y_.data();
    Lock l(mu_);
    busy_timer_.start();
    int nt = std::max(1, (int)FLAGS_num_threads);
    const size_t rows = X->rows();
    parallelFor(nt, nt, [&](int t) {  // every thread owns a slice of the rows: no write conflicts
      SizeR rr = SizeR(0, rows).evenDivide(nt, t);
      for (size_t jj = 0; jj < col_range.size(); ++jj) {
        const size_t j = col_range.begin() + jj;
        const Real wd = delta_w[jj];
        if (wd == 0 || !active.test(j)) continue;
        for (size_t o = offset[j]; o < offset[j + 1]; ++o) {
          const uint32 i = index[o];
          if (!rr.contains(i)) continue;
          dual_[i] *= binary ? exp(y[i] * wd) : exp(y[i] * wd * value[o]);
        }
      }
    });
    busy_timer_.stop();
  }

  void evaluate(BCDProgress* prog) override {
    Lock l(mu_);
    busy_timer_.start();
    Real s = 0;
    for (size_t i = 0; i < dual_.size(); ++i) s += log(1 + 1 / dual_[i]);
    prog->set_objective(s);
    prog->add_busy_time(busy_timer_.stop());
    busy_timer_.restart();
  }

  Timer busy_timer_;
};

}  // namespace LM
}  // namespace PS
