// Darlin: delayed block proximal gradient for L1 logistic regression, host roles.
// ref src/app/linear_method/darlin.h:12-514 (the arithmetic), src/learner/bcd.h (the roles)
//
// The arithmetic lives in DarlinMath as three free-standing kernels over plain arrays -
// column gradients, the proximal Newton step of a coordinate range, the dual rescaling - so
// that the roles below only move data: the scheduler issues the passes (PassIssuer, StopRule),
// the worker turns a block into (G, U) and a pulled weight block into a dual update, the server
// turns summed (G, U) into new weights.  The device engine runs the same three kernels as
// csrc/cuda/darlin.cu; tests compare both against one PyTorch oracle.
//
// With -device gpu the model shards are KVBufferedVector columns in HBM and the (G, U) / weight
// frames travel through the PeerVan; the math of this file stays on the host.
#pragma once
#include "app/linear_method/loss.h"
#include "learner/bcd.h"
#include "util/sketch.h"

namespace PS {
namespace LM {
typedef double Real;

// ---- the numerical core -----------------------------------------------------------------------------
struct DarlinMath {
  // trust region of a coordinate after it moved by d (ref darlin.h:23-25)
  static Real trustRegion(Real delta_max, Real d) { return std::min(delta_max, 2 * fabs(d) + .1); }

  // per active column j of [lo, hi):  g_j = -sum_i y_i tau_i x_ij
  //                                  u_j = sum_i min(tau_i (1 - tau_i) e^{|x_ij| delta_j}, 1/4) x_ij^2
  // with tau_i = 1 / (1 + dual_i).  Inactive columns get the NaN mark.  G, U are indexed from lo.
  // ref darlin.h:381-427
  template <bool kBinary>
  static void columnGradients(const SparseMatrix<uint32, Real>& X, const Real* y, const Real* dual,
                              const Bitmap& active, const Real* delta, size_t lo, size_t hi, Real* G, Real* U) {
    const size_t* colp = X.offset().data();
    const uint32* row = X.index().data();
    const Real* val = X.value().data();
    for (size_t j = lo; j < hi; ++j) {
      Real& g = G[j - lo];
      Real& u = U[j - lo];
      if (!active.test(j)) {
        SparseFilter::mark(&g);
        SparseFilter::mark(&u);
        continue;
      }
      g = u = 0;
      const Real stretch = kBinary ? exp(delta[j]) : delta[j];
      for (size_t e = colp[j]; e < colp[j + 1]; ++e) {
        const uint32 i = row[e];
        const Real tau = 1 / (1 + dual[i]);
        const Real curv = tau * (1 - tau);
        if (kBinary) {
          g -= y[i] * tau;
          u += std::min(curv * stretch, .25);
        } else {
          const Real x = val[e];
          g -= y[i] * tau * x;
          u += std::min(curv * exp(fabs(x) * stretch), .25) * x * x;
        }
      }
    }
  }

  struct StepParams {
    Real lambda, eta, delta_max, kkt_threshold;
  };
  // coordinates [first, first + n) of a shard: soft-thresholded Newton step inside the trust
  // region; a zero weight whose gradient sits well inside [-lambda, lambda] leaves the active set
  // (KKT filter) and is marked NaN.  Returns the largest KKT violation seen.  ref darlin.h:206-246
  static Real proximalStep(const StepParams& p, size_t first, size_t n, const Real* G, const Real* U,
                           Real* w, Real* delta, Bitmap* active) {
    Real worst = 0;
    for (size_t i = 0; i < n; ++i) {
      const size_t k = first + i;
      if (!active->test(k)) continue;
      const Real u = U[i] / p.eta + 1e-10;
      const Real up = G[i] + p.lambda, down = G[i] - p.lambda;  // one-sided gradients at w
      Real& wk = w[k];
      if (wk == 0) {
        if (up < 0) worst = std::max(worst, -up);
        else if (down > 0) worst = std::max(worst, down);
        else if (up > p.kkt_threshold && down < -p.kkt_threshold) {
          active->clear(k);
          SparseFilter::mark(&wk);
          continue;
        }
      }
      Real d = -wk;  // lands on zero unless one side's Newton step stays on its side
      if (up <= u * wk) d = -up / u;
      else if (down >= u * wk) d = -down / u;
      d = std::max(-delta[k], std::min(delta[k], d));
      delta[k] = trustRegion(p.delta_max, d);
      wk += d;
    }
    return worst;
  }

  // rows [r_lo, r_hi) of dual *= exp(y_i dw_j x_ij) over the columns of the block (dw indexed
  // from lo).  Threads own disjoint row slices, so there is nothing to lock.  ref darlin.h:472-502
  template <bool kBinary>
  static void rescaleDual(const SparseMatrix<uint32, Real>& X, const Real* y, const Bitmap& active,
                          const Real* dw, size_t lo, size_t hi, size_t r_lo, size_t r_hi, Real* dual) {
    const size_t* colp = X.offset().data();
    const uint32* row = X.index().data();
    const Real* val = X.value().data();
    for (size_t j = lo; j < hi; ++j) {
      const Real d = dw[j - lo];
      if (d == 0 || !active.test(j)) continue;
      for (size_t e = colp[j]; e < colp[j + 1]; ++e) {
        const uint32 i = row[e];
        if (i < r_lo || i >= r_hi) continue;
        dual[i] *= exp(y[i] * d * (kBinary ? 1 : val[e]));
      }
    }
  }
};

// the coordinate state a server shard and a worker both keep per feature group
struct CoordinateState {
  Bitmap active;
  SArray<Real> delta;
  void reset(size_t n, Real delta_init) {
    active.resize(n, true);
    delta = SArray<Real>(n, delta_init);
  }
};

// ---- scheduler ------------------------------------------------------------------------------------------
// "stop when the relative objective change stayed below epsilon for two passes in a row, the
// second of which ran over all coordinates again" (ref darlin.h:103-118)
class StopRule {
 public:
  explicit StopRule(Real epsilon) : eps_(epsilon) {}
  // returns true when training should stop; resetFilter() tells how to run the next pass
  bool after(Real relative_objective) {
    const bool small = relative_objective > 0 && relative_objective <= eps_;
    const bool stop = small && reset_;
    reset_ = small;
    return stop;
  }
  bool resetFilter() const { return reset_; }

 private:
  Real eps_;
  bool reset_ = false;
};

class DarlinScheduler : public BCDScheduler {
 public:
  DarlinScheduler(const string& name, const ConfigView& conf) : BCDScheduler(name, conf.darlin()), conf_(conf) {
    board_.insertColumn(1, ProgressBoard::Column{
        {"|      KKT filter     ", "| threshold  #activet ", "+---------------------"},
        [this](int, const BCDProgress& p) {
          return ProgressBoard::fmt("| %.1e %11llu ", kkt_threshold_, (unsigned long long)p.nnz_active_set());
        }});
  }
  virtual ~DarlinScheduler() { stopExecutor(); }

  void run() override {
    CHECK_EQ((int)conf_.loss().type(), (int)LossConfigView::LOGIT) << "Darlin is logistic loss only";
    CHECK_EQ((int)conf_.penalty().type(), (int)PenaltyConfigView::L1) << "Darlin is L1 only";
    LI << "Train l_1 logistic regression by block coordinate descent";
    loadTrainingData(conf_.training_data());
    divideFeatureBlocks();
    const BCDConfigView darlin = conf_.darlin();
    tau_ = darlin.max_block_delay();
    LI << "Maximal allowed delay: " << tau_;
    if (!darlin.random_feature_block_order()) LI << "Warning: Randomized block order often acclerates the convergence.";

    nodes_ = port(kCompGroup);
    // block tasks start well above the timestamps the preprocessing used
    time_ = std::max(10000, nodes_->time() + (int)fea_grp_.size() * BcdClock::kStride);
    first_ = time_ + 1;
    StopRule rule(darlin.epsilon());
    const int passes = darlin.max_pass_of_data();
    for (int iter = 0; iter < passes; ++iter) {
      issuePass(iter, rule.resetFilter());
      const BCDProgress row = board_.row(iter);
      board_.print(iter);
      kkt_threshold_ = row.violation() / (Real)train_info_.num_ex() * bcd_conf_.kkt_filter_threshold_ratio();
      if (rule.after(row.relative_obj())) {
        LI << "Stopped: relative objective <= " << darlin.epsilon();
        break;
      }
      if (iter + 1 == passes) LI << "Reached maximal " << passes << " data passes";
    }
    if (conf_.has_model_output()) saveModel(conf_.model_output());
  }

 private:
  // one UPDATE_MODEL per block, each allowed to start once the block tau steps earlier has
  // finished on the receiving node (prior blocks: no delay), then EVALUATE_PROGRESS behind the
  // same bound.  ref darlin.h:58-103
  void issuePass(int iter, bool reset_filter) {
    const std::vector<int> order = planner_.pass(iter, conf_.darlin().random_feature_block_order(), &rng_);
    const size_t n_prior = planner_.numPrior(iter);
    for (size_t i = 0; i < order.size(); ++i) {
      const FeatureBlock& blk = planner_[order[i]];
      Task step;
      BCDCallView call = step.mutable_bcd();
      call.set_cmd(BCDCallView::UPDATE_MODEL);
      call.mutable_key().from(blk.keys);
      call.add_fea_grp(blk.group);
      if (i == 0) {  // per-pass settings ride on the first block
        call.set_kkt_filter_threshold(kkt_threshold_);
        if (reset_filter) call.set_reset_kkt_filter(true);
      }
      step.set_time(time_ + 1);
      step.add_wait_time(dependency(i < n_prior ? 0 : tau_));
      time_ = nodes_->submit(step);
    }
    Task eval;
    eval.mutable_bcd().set_cmd(BCDCallView::EVALUATE_PROGRESS);
    const int dep = dependency(tau_);
    if (dep != Message::kInvalidTime) eval.add_wait_time(dep);
    time_ = nodes_->submitAndWait(eval, [this, iter]() { mergeProgress(iter); });
  }
  // the task `delay` steps before the next one, if this run issued it
  int dependency(int delay) const { return time_ - delay >= first_ ? time_ - delay : Message::kInvalidTime; }

  Config conf_;
  RNodePtr nodes_;
  int time_ = 0, first_ = 0, tau_ = 0;
  Real kkt_threshold_ = 1e20;
};

// ---- server --------------------------------------------------------------------------------------------
class DarlinServer : public BCDServer<Real> {
 public:
  DarlinServer(const string& name, const ConfigView& conf) : BCDServer<Real>(name, conf.darlin()), conf_(conf) {}
  virtual ~DarlinServer() { stopExecutor(); }

 protected:
  void preprocessData(int time, const BCDCallView& call) override {
    BCDServer<Real>::preprocessData(time, call);
    for (int g : fea_grp_) state_[g].reset(model_.key(g).size(), bcd_conf_.delta_init_value());
  }

  // ref darlin.h:175-204
  void updateModel(int time, const BCDCallView& call) override {
    if (call.has_kkt_filter_threshold()) {  // first block of a pass
      kkt_threshold_ = call.kkt_filter_threshold();
      violation_ = 0;
    }
    if (call.reset_kkt_filter())
      for (int g : fea_grp_) state_[g].active.fill(true);
    CHECK_EQ(call.fea_grp_size(), 1);
    const int grp = call.fea_grp(0);
    const Range<Key> block = call.key().to<Key>();
    if (myKeyRange().setIntersection(block).empty()) return;  // the block lies outside my shard
    const SizeR mine = model_.find(grp, block);
    model_.waitInMsg(kWorkerGroup, time);  // every worker's (G, U) has been summed into the buffer
    if (!mine.empty()) {
      auto sums = model_.receivedOnHost(time);
      CHECK_EQ(sums.first, mine);
      CHECK_EQ(sums.second.size(), (size_t)2);
      CHECK_EQ(sums.second[0].size(), mine.size());
      CoordinateState& st = state_[grp];
      const DarlinMath::StepParams p{conf_.penalty().lambda(0), conf_.learning_rate().alpha(),
                                     bcd_conf_.delta_max_value(), kkt_threshold_};
      const Real vio = DarlinMath::proximalStep(p, mine.begin(), mine.size(), sums.second[0].data(),
                                                sums.second[1].data(), model_.value(grp).data(),
                                                st.delta.data(), &st.active);
      violation_ = std::max(violation_, vio);
    }
    model_.finish(kWorkerGroup, BcdClock::updated(time));  // lets the workers' pulls through
  }

  // ref darlin.h:248-265
  void evaluate(BCDProgress* prog) override {
    size_t nnz = 0, n_active = 0;
    Real l1 = 0;
    for (int g : fea_grp_) {
      for (Real w : model_.value(g))
        if (w != 0 && !SparseFilter::marked(w)) { ++nnz; l1 += fabs(w); }
      n_active += state_[g].active.nnz();
    }
    prog->set_objective(l1 * conf_.penalty().lambda(0));
    prog->set_nnz_w(nnz);
    prog->set_violation(violation_);
    prog->set_nnz_active_set(n_active);
  }

 private:
  Config conf_;
  std::unordered_map<int, CoordinateState> state_;
  Real kkt_threshold_ = 1e20, violation_ = 0;
};

// ---- worker --------------------------------------------------------------------------------------------
class DarlinWorker : public BCDWorker<Real> {
 public:
  DarlinWorker(const string& name, const ConfigView& conf) : BCDWorker<Real>(name, conf.darlin()), conf_(conf) {}
  virtual ~DarlinWorker() { stopExecutor(); }

 protected:
  void preprocessData(int time, const BCDCallView& call) override {
    BCDWorker<Real>::preprocessData(time, call);
    // BCDWorker left X w in dual_; Darlin keeps exp(y .* X w)
    for (size_t i = 0; i < dual_.size(); ++i) dual_[i] = exp(y_[i] * dual_[i]);
    for (int g : fea_grp_) state_[g].reset(model_.key(g).size(), bcd_conf_.delta_init_value());
  }

  // a block on the worker: push (G, U) at `time`, then pull the new weights behind the server's
  // "updated" mark; the scheduler's task is finished by the pull's callback.  ref darlin.h:297-342
  void computeGradient(int time, const BCDCallView& call, MessagePtr task) override {
    if (call.reset_kkt_filter()) {
      Lock l(mu_);  // pulls of the previous pass' last blocks may still be landing (tau > 0)
      for (int g : fea_grp_) state_[g].active.fill(true);
    }
    CHECK_EQ(call.fea_grp_size(), 1);
    const int grp = call.fea_grp(0);
    const Range<Key> block = call.key().to<Key>();
    const SizeR cols = model_.find(grp, block);

    SArray<Real> G(cols.size(), 0), U(cols.size(), 0);
    if (!cols.empty()) {
      Lock l(mu_);  // dual_ is also written by the pull callback (model executor thread)
      busy_.start();
      const auto X = matrix(grp);
      const CoordinateState& st = state_[grp];
      const int threads = cols.size() < 64 ? 1 : std::max(1, (int)FLAGS_num_threads);
      parallelFor(threads, threads, [&](int t) {
        const SizeR part = cols.evenDivide(threads, t);
        if (part.empty()) return;
        Real* g = G.data() + (part.begin() - cols.begin());
        Real* u = U.data() + (part.begin() - cols.begin());
        if (X->binary())
          DarlinMath::columnGradients<true>(*X, y_.data(), dual_.data(), st.active, st.delta.data(), part.begin(), part.end(), g, u);
        else
          DarlinMath::columnGradients<false>(*X, y_.data(), dual_.data(), st.active, st.delta.data(), part.begin(), part.end(), g, u);
      });
      busy_.stop();
    }
    auto addressed = [&](int t, int wait) {
      MessagePtr m(new Message(kServerGroup, t, wait));
      m->setKey(model_.key(grp).segment(cols));
      m->task.mutable_key_range().from(block);
      m->task.set_key_channel(grp);
      m->addFilter(FilterConfigView::KEY_CACHING);
      return m;
    };
    MessagePtr push = addressed(time, Message::kInvalidTime);
    push->addValue({model_.stage(G), model_.stage(U)});  // -device gpu: through peer memory
    CHECK_EQ(time, model_.push(push));

    const int tp = BcdClock::pull(time);
    MessagePtr pull = addressed(tp, BcdClock::updated(time));
    pull->fin_handle = [this, tp, grp, cols, task]() {
      if (!cols.empty()) {
        auto got = model_.receivedOnHost(tp);
        CHECK_EQ(got.first, cols);
        CHECK_EQ(got.second.size(), (size_t)1);
        adoptWeights(grp, cols, got.second[0]);
      }
      port(task->sender)->finishIncomingTask(task->task.time());
      sys_.reply(task->sender, task->task);
    };
    CHECK_EQ(tp, model_.pull(pull));
  }

  // ref darlin.h:429-502: a NaN weight means "the server filtered this coordinate"
  void adoptWeights(int grp, SizeR cols, const SArray<Real>& fresh) {
    const auto X = matrix(grp);
    // the active bitmap shares words with the columns of other blocks, which a gradient task of
    // the app executor may be reading (tau > 0): the whole update runs under the lock
    Lock l(mu_);
    CoordinateState& st = state_[grp];
    SArray<Real>& w = model_.value(grp);
    const Real delta_max = bcd_conf_.delta_max_value();
    SArray<Real> dw(fresh.size(), 0);
    for (size_t i = 0; i < fresh.size(); ++i) {
      const size_t j = cols.begin() + i;
      if (SparseFilter::marked(fresh[i])) {
        st.active.clear(j);
        w[j] = 0;
        continue;
      }
      dw[i] = fresh[i] - w[j];
      st.delta[j] = DarlinMath::trustRegion(delta_max, dw[i]);
      w[j] = fresh[i];
    }
    busy_.start();
    const int threads = std::max(1, (int)FLAGS_num_threads);
    const SizeR rows(0, X->rows());
    parallelFor(threads, threads, [&](int t) {
      const SizeR r = rows.evenDivide(threads, t);
      if (X->binary())
        DarlinMath::rescaleDual<true>(*X, y_.data(), st.active, dw.data(), cols.begin(), cols.end(), r.begin(), r.end(), dual_.data());
      else
        DarlinMath::rescaleDual<false>(*X, y_.data(), st.active, dw.data(), cols.begin(), cols.end(), r.begin(), r.end(), dual_.data());
    });
    busy_.stop();
  }

  // ref darlin.h:504-511
  void evaluate(BCDProgress* prog) override {
    Lock l(mu_);
    busy_.start();
    Real loss = 0;
    for (Real d : dual_) loss += log(1 + 1 / d);
    prog->set_objective(loss);
    prog->add_busy_time(busy_.stop());
    busy_.restart();
  }

 private:
  std::shared_ptr<SparseMatrix<uint32, Real>> matrix(int grp) {
    auto X = std::static_pointer_cast<SparseMatrix<uint32, Real>>(X_[grp]);
    CHECK(X && X->colMajor()) << "group " << grp << " is not localized by column";
    return X;
  }
  Config conf_;
  std::unordered_map<int, CoordinateState> state_;
  Timer busy_;
};

}  // namespace LM
}  // namespace PS
