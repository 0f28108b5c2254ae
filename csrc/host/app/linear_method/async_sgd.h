// Asynchronous minibatch SGD for linear models: FTRL, AdaGrad and plain SGD server updates.
// ref src/app/linear_method/async_sgd.h:18-301
// Fixed relative to the reference: AdaGrad / SGD entries really update (they are TODO no-ops
// there, :72-99), `ada_grad` selects AdaGrad (the reference's branch is inverted, :136-140),
// and SGDConfig.max_delay bounds the number of minibatches in flight per worker (declared but
// never read in the reference, linear.proto:47).
#pragma once
#include "app/linear_method/darlin.h"
#include "learner/sgd.h"

namespace PS {
namespace LM {

class AsyncSGDScheduler : public ISGDScheduler, public LinearMethod {
 public:
  AsyncSGDScheduler(const string& name, const ConfigView& conf)
      : ISGDScheduler(name), LinearMethod(conf) {}
  virtual ~AsyncSGDScheduler() { stopExecutor(); }
  void run() override {
    updateModel(conf_.training_data(), conf_.async_sgd().report_interval());
    if (conf_.has_model_output()) saveModel();
  }
  double lastLoss() const { return last_loss_; }
  double lastAUC() const { return last_auc_; }
};

template <typename V>
struct SGDState {
  SGDState() {}
  SGDState(const PenaltyConfigView& h_conf, const LearningRateConfigView& lr_conf) {
    lr = shared_ptr<LearningRate<V>>(new LearningRate<V>(lr_conf));
    h = createPenalty<V>(h_conf);
  }
  // called once per applied push: report the model statistics to the scheduler
  void update() {
    ++iter;
    if (!reporter) return;
    SGDProgress prog;
    prog.set_nnz(nnz);
    prog.set_weight_sum(weight_sum);
    prog.set_delta_sum(delta_sum);
    weight_sum = 0;
    delta_sum = 0;
    reporter->report(prog);
  }
  void updateWeight(V new_weight, V old_weight) {
    if (new_weight == 0 && old_weight != 0) --nnz;
    else if (new_weight != 0 && old_weight == 0) ++nnz;
    weight_sum += new_weight * new_weight;
    V delta = new_weight - old_weight;
    delta_sum += delta * delta;
  }
  shared_ptr<LearningRate<V>> lr;
  PenaltyPtr<V> h;
  int iter = 0;
  size_t nnz = 0;
  V weight_sum = 0, delta_sum = 0;
  MonitorSlaver<SGDProgress>* reporter = nullptr;
};

// w <- prox_h(w - eta g, eta), eta = alpha / (beta + sqrt(sum g^2))
template <typename V>
struct AdaGradEntry {
  void get(V const* data, SGDState<V>* state) {
    V grad = *data, w_old = weight;
    sum_sq_grad += grad * grad;
    V eta = state->lr->eval(sqrt(sum_sq_grad));
    weight = state->h->proximal(weight - eta * grad, eta);
    state->updateWeight(weight, w_old);
  }
  void put(V* data, SGDState<V>*) { *data = weight; }
  V weight = 0, sum_sq_grad = 0;
};

// w <- prox_h(w - eta g, eta), eta = alpha / (beta + sqrt(t))
template <typename V>
struct SGDEntry {
  void get(V const* data, SGDState<V>* state) {
    V grad = *data, w_old = weight;
    V eta = state->lr->eval(sqrt((V)(state->iter + 1)));
    weight = state->h->proximal(weight - eta * grad, eta);
    state->updateWeight(weight, w_old);
  }
  void put(V* data, SGDState<V>*) { *data = weight; }
  V weight = 0;
};

// FTRL-proximal (ref async_sgd.h:101-124)
template <typename V>
struct FTRLEntry {
  V w = 0, z = 0, sqrt_n = 0;
  void get(V const* data, SGDState<V>* state) {
    V w_old = w, grad = *data;
    V sqrt_n_new = sqrt(sqrt_n * sqrt_n + grad * grad);
    V sigma = (sqrt_n_new - sqrt_n) / state->lr->alpha();
    z += grad - sigma * w;
    sqrt_n = sqrt_n_new;
    V eta = state->lr->eval(sqrt_n);
    w = state->h->proximal(-z * eta, eta);
    state->updateWeight(w, w_old);
  }
  void put(V* data, SGDState<V>*) { *data = w; }
};

template <typename V>
class AsyncSGDServer : public ISGDCompNode, public LinearMethod {
 public:
  AsyncSGDServer(const string& name, const ConfigView& conf) : ISGDCompNode(name), LinearMethod(conf) {
    SGDConfigView sgd = conf_.async_sgd();
    if (sgd.algo() == SGDConfigView::FTRL)
      model_.reset(new KVStore<Key, V, FTRLEntry<V>, SGDState<V>>(name + "_model", name));
    else if (sgd.ada_grad())
      model_.reset(new KVStore<Key, V, AdaGradEntry<V>, SGDState<V>>(name + "_model", name));
    else
      model_.reset(new KVStore<Key, V, SGDEntry<V>, SGDState<V>>(name + "_model", name));
  }
  virtual ~AsyncSGDServer() { stopExecutor(); }

  void init() override {
    SGDState<V> state(conf_.penalty(), conf_.learning_rate());
    state.reporter = &reporter_;
    model_->setState(state);
    SGDConfigView sgd = conf_.async_sgd();
    if (sgd.tail_feature_freq() > 0)
      model_->setTailFilterSize(0, (int)(sgd.countmin_n() / std::max(1, sys_.yp().num_servers())),
                                sgd.countmin_k());
  }
  void process(const MessagePtr& msg) override {
    if (msg->task.sgd().cmd() == SGDCallView::SAVE_MODEL) saveModel();
  }
  void saveModel() {
    DataConfigView output = conf_.model_output();
    if (output.format() != DataConfigView::TEXT) return;
    CHECK(output.file_size());
    string file = output.file(0) + "_" + myNodeID();
    model_->writeToFile(file);
    LI << myNodeID() << " written the model to " << file;
  }

 protected:
  unique_ptr<KVState<Key, SGDState<V>>> model_;
};

template <typename V>
class AsyncSGDWorker : public ISGDCompNode, public LinearMethod {
 public:
  AsyncSGDWorker(const string& name, const ConfigView& conf)
      : ISGDCompNode(name), LinearMethod(conf), model_(name + "_model", name), processed_batch_(0) {
    loss_ = createLoss<V>(conf_.loss());
  }
  virtual ~AsyncSGDWorker() { stopExecutor(); }

  void process(const MessagePtr& msg) override {
    SGDCallView sgd = msg->task.sgd();
    if (sgd.cmd() == SGDCallView::UPDATE_MODEL) updateModel(sgd);
  }

  void updateModel(const SGDCallView& call) {
    SGDConfigView sgd = conf_.async_sgd();
    MinibatchReader<V> reader;
    const int n = std::max(sgd.num_data_pass(), 1);
    const int m = call.data().file_size();
    std::vector<int> idx(m);
    for (int i = 0; i < m; ++i) idx[i] = i;
    DataConfig data(call.data());
    data.clear_file();
    std::mt19937 rng(1234 + myRank());
    for (int i = 0; i < n; ++i) {  // every pass visits the files in a fresh random order
      std::shuffle(idx.begin(), idx.end(), rng);
      for (int j = 0; j < m; ++j) data.add_file(call.data().file(idx[j]));
    }
    reader.setReader(data, sgd.minibatch(), sgd.data_buf());
    reader.setFilter((size_t)sgd.countmin_n(), sgd.countmin_k(), sgd.tail_feature_freq());
    reader.start();

    // tau: minibatches allowed in flight beyond the current one; unset keeps the reference's
    // behaviour (unbounded, limited only by the prefetch queue)
    const int max_delay = sgd.has_max_delay() ? sgd.max_delay() : -1;
    int id = 0;
    SArray<Key> key;
    while (true) {
      std::pair<MatrixPtr<V>, MatrixPtr<V>> yx;
      if (!reader.read(yx.first, yx.second, key)) break;
      if (max_delay >= 0) {  // bounded delay: wait until minibatch id - tau - 1 has been pushed
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [&] { return id - processed_batch_ <= max_delay; });
      }
      {
        Lock l(mu_);
        data_[id] = yx;
      }
      MessagePtr msg(new Message(kServerGroup));
      msg->setKey(key);
      msg->task.set_key_channel(id);
      // caching the keys on the pull lets the following push travel without them (the
      // reference only tags the push, so its cache never hits, async_sgd.h:228-233 vs :272)
      msg->addFilter(FilterConfigView::KEY_CACHING);
      msg->fin_handle = [this, id]() { computeGradient(id); };
      model_.key(id) = key;
      model_.pull(msg);
      ++id;
    }
    std::unique_lock<std::mutex> l(mu_);
    cv_.wait(l, [&] { return processed_batch_ >= id; });
  }

  // runs on the model customer's executor thread once all servers answered the pull
  void computeGradient(int id) {
    MatrixPtr<V> Y, X;
    {
      Lock l(mu_);
      Y = data_[id].first;
      X = data_[id].second;
      data_.erase(id);
    }
    CHECK_EQ(X->rows(), Y->rows());
    SArray<V> w = model_.value(id);
    if (w.size() != X->cols()) {  // nothing pulled (all keys new on empty servers)
      w.resize(X->cols());
      w.setZero();
    }
    SArray<V> Xw = X->times(w);
    if (getenv("PSB_DEBUG_SGD")) fprintf(stderr, "dbg id=%d rows=%zu cols=%zu nnz=%zu |w|0=%zu wsum=%g Xw0=%g y0=%g keys=%zu\n", id, X->rows(), X->cols(), X->nnz(), w.nnz(), w.sum(), (double)Xw[0], (double)Y->value()[0], model_.key(id).size());
    SGDProgress prog;
    prog.add_objective(loss_->evaluate(Y->value(), Xw));
    prog.add_auc(Evaluation<V>::auc(Y->value(), Xw));
    prog.add_accuracy(Evaluation<V>::accuracy(Y->value(), Xw));
    prog.set_num_examples_processed(Xw.size());
    reporter_.report(prog);

    SArray<V> grad(X->cols());
    grad.setZero();
    loss_->compute(Y->value(), X, Xw, grad);

    MessagePtr msg(new Message(kServerGroup));
    msg->setKey(model_.key(id));
    msg->addValue(grad);
    msg->task.set_key_channel(id);
    msg->addFilter(FilterConfigView::KEY_CACHING).set_clear_cache_if_done(true);
    int nbytes = conf_.async_sgd().fixing_float_by_nbytes();
    if (nbytes) msg->addFilter(FilterConfigView::FIXING_FLOAT).add_fixed_point().set_num_bytes(nbytes);
    model_.push(msg);
    model_.clear(id);
    {
      Lock l(mu_);
      ++processed_batch_;
    }
    cv_.notify_all();
  }

 private:
  KVVector<Key, V> model_;
  LossPtr<V> loss_;
  std::unordered_map<int, std::pair<MatrixPtr<V>, MatrixPtr<V>>> data_;
  std::mutex mu_;
  std::condition_variable cv_;
  int processed_batch_;
};

}  // namespace LM
}  // namespace PS
