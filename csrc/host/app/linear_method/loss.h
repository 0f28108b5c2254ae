// Scalar losses over (y, Xw), penalties and learning rates of the linear method.
// ref src/app/linear_method/loss.h:10-129, penalty.h:9-78, learning_rate.h:7-29
// Device counterparts: csrc/cuda/sgd.cu (logit / square-hinge gradient kernels, FTRL / AdaGrad /
// SGD updates with the same proximal operator).
#pragma once
#include "proto/messages.h"
#include "util/matrix.h"

namespace PS {
namespace LM {

template <typename T>
class Loss {
 public:
  virtual ~Loss() {}
  // objective value given labels y and margins Xw
  virtual T evaluate(const SArray<T>& y, const SArray<T>& Xw) = 0;
  // gradient (and optional diagonal Hessian) with respect to w
  virtual void compute(const SArray<T>& y, const MatrixPtr<T>& X, const SArray<T>& Xw,
                       SArray<T> gradient, SArray<T> diag_hessian = SArray<T>()) = 0;
};
template <typename T> using LossPtr = shared_ptr<Loss<T>>;

// sum_i log(1 + exp(-y_i <x_i,w>))
template <typename T>
class LogitLoss : public Loss<T> {
 public:
  T evaluate(const SArray<T>& y, const SArray<T>& Xw) override {
    CHECK_EQ(y.size(), Xw.size());
    T s = 0;
    for (size_t i = 0; i < y.size(); ++i) {
      T m = -y[i] * Xw[i];
      s += m > 0 ? m + log1p(exp(-m)) : log1p(exp(m));  // overflow-safe log(1+e^m)
    }
    return s;
  }
  void compute(const SArray<T>& y, const MatrixPtr<T>& X, const SArray<T>& Xw, SArray<T> gradient,
               SArray<T> diag_hessian) override {
    CHECK_EQ(y.size(), Xw.size());
    CHECK_EQ(y.size(), X->rows());
    SArray<T> tau(y.size());
    for (size_t i = 0; i < y.size(); ++i) tau[i] = 1 / (1 + exp(y[i] * Xw[i]));
    if (gradient.size()) {
      CHECK_EQ(gradient.size(), X->cols());
      SArray<T> r(y.size());
      for (size_t i = 0; i < y.size(); ++i) r[i] = -y[i] * tau[i];
      X->trans()->times(r.data(), gradient.data());
    }
    if (diag_hessian.size()) {
      CHECK_EQ(diag_hessian.size(), X->cols());
      SArray<T> h(y.size());
      for (size_t i = 0; i < y.size(); ++i) h[i] = tau[i] * (1 - tau[i]);
      X->dotTimes(X)->trans()->times(h.data(), diag_hessian.data());
    }
  }
};

// sum_i max(0, 1 - y_i <x_i,w>)^2
template <typename T>
class SquareHingeLoss : public Loss<T> {
 public:
  T evaluate(const SArray<T>& y, const SArray<T>& Xw) override {
    T s = 0;
    for (size_t i = 0; i < y.size(); ++i) {
      T m = std::max((T)0, 1 - y[i] * Xw[i]);
      s += m * m;
    }
    return s;
  }
  void compute(const SArray<T>& y, const MatrixPtr<T>& X, const SArray<T>& Xw, SArray<T> gradient,
               SArray<T> diag_hessian) override {
    if (!gradient.size()) return;
    SArray<T> r(y.size());
    for (size_t i = 0; i < y.size(); ++i) r[i] = -2 * y[i] * std::max((T)0, 1 - y[i] * Xw[i]);
    X->trans()->times(r.data(), gradient.data());
    if (diag_hessian.size()) {
      SArray<T> h(y.size());
      for (size_t i = 0; i < y.size(); ++i) h[i] = (1 - y[i] * Xw[i]) > 0 ? 2 : 0;
      X->dotTimes(X)->trans()->times(h.data(), diag_hessian.data());
    }
  }
};

// 0.5 * sum_i (y_i - <x_i,w>)^2  (SQUARE is declared but unimplemented in the reference, loss.h:119-127)
template <typename T>
class SquareLoss : public Loss<T> {
 public:
  T evaluate(const SArray<T>& y, const SArray<T>& Xw) override {
    T s = 0;
    for (size_t i = 0; i < y.size(); ++i) s += (y[i] - Xw[i]) * (y[i] - Xw[i]);
    return s / 2;
  }
  void compute(const SArray<T>& y, const MatrixPtr<T>& X, const SArray<T>& Xw, SArray<T> gradient,
               SArray<T> diag_hessian) override {
    if (gradient.size()) {
      SArray<T> r(y.size());
      for (size_t i = 0; i < y.size(); ++i) r[i] = Xw[i] - y[i];
      X->trans()->times(r.data(), gradient.data());
    }
    if (diag_hessian.size()) {
      SArray<T> h(y.size(), 1);
      X->dotTimes(X)->trans()->times(h.data(), diag_hessian.data());
    }
  }
};

template <typename T>
LossPtr<T> createLoss(const LossConfigView& config) {
  switch (config.type()) {
    case LossConfigView::LOGIT: return LossPtr<T>(new LogitLoss<T>());
    case LossConfigView::SQUARE_HINGE: return LossPtr<T>(new SquareHingeLoss<T>());
    case LossConfigView::SQUARE: return LossPtr<T>(new SquareLoss<T>());
    default: CHECK(false) << "unsupported loss " << config.DebugString();
  }
  return LossPtr<T>();
}

// ---- penalties ------------------------------------------------------------------------------------
template <typename T>
class Penalty {
 public:
  virtual ~Penalty() {}
  virtual T eval(const SArray<T>& w) = 0;
  // argmin_x 0.5/eta (x - z)^2 + h(x)
  virtual T proximal(T z, T eta) = 0;
};

// lambda1 |x| + lambda2 x^2
template <typename T>
class ElasticNet : public Penalty<T> {
 public:
  ElasticNet(T lambda1, T lambda2) : lambda1_(lambda1), lambda2_(lambda2) {
    CHECK_GE(lambda1, 0);
    CHECK_GE(lambda2, 0);
  }
  T eval(const SArray<T>& w) override {  // the reference returns 0 here (penalty.h:36, TODO)
    T s = 0;
    for (size_t i = 0; i < w.size(); ++i)
      if (w[i] == w[i]) s += lambda1_ * fabs(w[i]) + lambda2_ * w[i] * w[i];
    return s;
  }
  T proximal(T z, T eta) override {
    T leta = lambda1_ * eta;
    if (z <= leta && z >= -leta) return 0;
    return z > 0 ? (z - leta) / (1 + lambda2_ * eta) : (z + leta) / (1 + lambda2_ * eta);
  }
  T lambda1() const { return lambda1_; }
  T lambda2() const { return lambda2_; }

 private:
  T lambda1_, lambda2_;
};
template <typename T> using PenaltyPtr = shared_ptr<Penalty<T>>;

template <typename T>
PenaltyPtr<T> createPenalty(const PenaltyConfigView& conf) {
  CHECK_GE(conf.lambda_size(), 1);
  switch (conf.type()) {
    case PenaltyConfigView::L1: {
      T l1 = conf.lambda(0);
      T l2 = conf.lambda_size() > 1 ? conf.lambda(1) : 0;
      return PenaltyPtr<T>(new ElasticNet<T>(l1, l2));
    }
    case PenaltyConfigView::L2: return PenaltyPtr<T>(new ElasticNet<T>(0, conf.lambda(0)));
    default: CHECK(false) << "unknown penalty";
  }
  return PenaltyPtr<T>();
}

// ---- learning rate ----------------------------------------------------------------------------------
template <typename V>
class LearningRate {
 public:
  explicit LearningRate(const LearningRateConfigView& conf) : conf_(conf) {
    CHECK_GT(alpha(), 0);
    CHECK_GE(beta(), 0);
  }
  V eval(V x = 0) const {
    if (conf_.type() == LearningRateConfigView::CONSTANT) return alpha();  // alpha
    return alpha() / (x + beta());                                          // alpha / (x + beta)
  }
  V alpha() const { return (V)conf_.alpha(); }
  V beta() const { return (V)conf_.beta(); }

 private:
  LearningRateConfig conf_;
};

}  // namespace LM
}  // namespace PS
