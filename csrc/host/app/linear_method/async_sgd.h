// Asynchronous minibatch SGD for linear models: FTRL, AdaGrad and plain SGD on the servers.
// ref src/app/linear_method/async_sgd.h:18-301
//
// Differences from the reference, all towards what it declares but does not do: AdaGrad / SGD
// really update (TODO no-ops there, :72-99), `ada_grad` selects AdaGrad (inverted there,
// :136-140), SGDConfig.max_delay bounds the minibatches a worker has in flight (never read
// there, linear.proto:47), and the pull is tagged for key caching so that the push that follows
// can travel without its keys (only the push is tagged there, so the cache never hits).
//
// Layout: a server-side optimizer is a *rule* (how a gradient changes one coordinate) wrapped
// by Coordinate<Rule>, which adds the bookkeeping every rule shares; the worker is a pipeline
// of two stages (request weights, answer with a gradient) with an InFlight window between them.
#pragma once
#include "app/linear_method/linear.h"
#include "app/linear_method/loss.h"
#include "learner/sgd.h"

namespace PS {
namespace LM {

class AsyncSGDScheduler : public ISGDScheduler {
 public:
  AsyncSGDScheduler(const string& name, const ConfigView& conf) : ISGDScheduler(name), conf_(conf) {}
  virtual ~AsyncSGDScheduler() { stopExecutor(); }
  void run() override {
    updateModel(conf_.training_data(), conf_.async_sgd().report_interval());
    if (conf_.has_model_output()) saveModel();
  }

 private:
  Config conf_;
};

// ---- server side: shared state + per-coordinate rules ------------------------------------------------
// what all coordinates of a shard share: step size schedule, penalty, model statistics
template <typename V>
struct SGDState {
  SGDState() {}
  SGDState(const PenaltyConfigView& penalty, const LearningRateConfigView& rate)
      : lr(new LearningRate<V>(rate)), h(createPenalty<V>(penalty)) {}
  // KVStore calls this after every applied push: the statistics go to the scheduler
  void update() {
    ++iter;
    if (!reporter) return;
    SGDProgress p;
    p.set_nnz(nnz);
    p.set_weight_sum(weight_sum);
    p.set_delta_sum(delta_sum);
    weight_sum = delta_sum = 0;
    reporter->report(p);
  }
  void moved(V from, V to) {
    nnz += (to != 0) - (from != 0);
    weight_sum += to * to;
    delta_sum += (to - from) * (to - from);
  }
  shared_ptr<LearningRate<V>> lr;
  PenaltyPtr<V> h;
  int iter = 0;
  size_t nnz = 0;
  V weight_sum = 0, delta_sum = 0;
  MonitorSlaver<SGDProgress>* reporter = nullptr;
};

// Entry interface of KVStore (get = apply a gradient, put = emit the weight) on top of a rule
// with `V apply(V w, V grad, SGDState<V>&)` returning the new weight
template <typename V, class Rule>
struct Coordinate : Rule {
  V w = 0;
  void get(V const* grad, SGDState<V>* s) {
    const V before = w;
    w = Rule::apply(before, *grad, *s);
    s->moved(before, w);
  }
  void put(V* out, SGDState<V>*) { *out = w; }
};

// w <- prox(w - eta g, eta) with eta = lr(sqrt(sum of squared gradients of this coordinate))
template <typename V>
struct AdaGradRule {
  V sum_sq = 0;
  V apply(V w, V g, SGDState<V>& s) {
    sum_sq += g * g;
    const V eta = s.lr->eval(sqrt(sum_sq));
    return s.h->proximal(w - eta * g, eta);
  }
};
// the same with eta = lr(sqrt(number of pushes applied so far))
template <typename V>
struct PlainRule {
  V apply(V w, V g, SGDState<V>& s) {
    const V eta = s.lr->eval(sqrt((V)(s.iter + 1)));
    return s.h->proximal(w - eta * g, eta);
  }
};
// FTRL-proximal (ref async_sgd.h:101-124): z accumulates g - sigma w with sigma the growth of
// sqrt(n) over alpha; the weight is the proximal point of -z eta
template <typename V>
struct FTRLRule {
  V z = 0, root_n = 0;
  V apply(V w, V g, SGDState<V>& s) {
    const V grown = sqrt(root_n * root_n + g * g);
    z += g - (grown - root_n) / s.lr->alpha() * w;
    root_n = grown;
    const V eta = s.lr->eval(root_n);
    return s.h->proximal(-z * eta, eta);
  }
};
template <typename V> using FTRLEntry = Coordinate<V, FTRLRule<V>>;
template <typename V> using AdaGradEntry = Coordinate<V, AdaGradRule<V>>;
template <typename V> using SGDEntry = Coordinate<V, PlainRule<V>>;

template <typename V>
class AsyncSGDServer : public ISGDCompNode {
 public:
  AsyncSGDServer(const string& name, const ConfigView& conf) : ISGDCompNode(name), conf_(conf) {
    const SGDConfigView sgd = conf_.async_sgd();
    const string store = name + "_model";
    if (sgd.algo() == SGDConfigView::FTRL) model_.reset(new KVStore<Key, V, FTRLEntry<V>, SGDState<V>>(store, name));
    else if (sgd.ada_grad()) model_.reset(new KVStore<Key, V, AdaGradEntry<V>, SGDState<V>>(store, name));
    else model_.reset(new KVStore<Key, V, SGDEntry<V>, SGDState<V>>(store, name));
  }
  virtual ~AsyncSGDServer() { stopExecutor(); }

  void init() override {
    SGDState<V> shared(conf_.penalty(), conf_.learning_rate());
    shared.reporter = &reporter_;
    model_->setState(shared);
    const SGDConfigView sgd = conf_.async_sgd();
    if (sgd.tail_feature_freq() > 0)
      model_->setTailFilterSize(0, (int)(sgd.countmin_n() / std::max(1, sys_.yp().num_servers())), sgd.countmin_k());
  }
  void process(const MessagePtr& msg) override {
    if (msg->task.sgd().cmd() != SGDCallView::SAVE_MODEL) return;
    const DataConfigView out = conf_.model_output();
    if (out.format() != DataConfigView::TEXT) return;
    CHECK(out.file_size()) << "model_output has no file";
    const string file = out.file(0) + "_" + myNodeID();
    model_->writeToFile(file);
    LI << myNodeID() << " written the model to " << file;
  }

 private:
  Config conf_;
  unique_ptr<KVState<Key, SGDState<V>>> model_;
};

// ---- worker -------------------------------------------------------------------------------------------------
// minibatches between "weights requested" and "gradient pushed"; enter() blocks while more than
// `limit` are in flight (limit < 0: never), drain() waits for all of them
class InFlight {
 public:
  explicit InFlight(int limit) : limit_(limit) {}
  void enter() {
    std::unique_lock<std::mutex> l(mu_);
    if (limit_ >= 0) cv_.wait(l, [&] { return started_ - done_ <= limit_; });
    ++started_;
  }
  void leave() {
    { Lock l(mu_); ++done_; }
    cv_.notify_all();
  }
  void drain() {
    std::unique_lock<std::mutex> l(mu_);
    cv_.wait(l, [&] { return done_ >= started_; });
  }

 private:
  std::mutex mu_;
  std::condition_variable cv_;
  int limit_, started_ = 0, done_ = 0;
};

template <typename V>
class AsyncSGDWorker : public ISGDCompNode {
 public:
  AsyncSGDWorker(const string& name, const ConfigView& conf)
      : ISGDCompNode(name), conf_(conf), model_(name + "_model", name), loss_(createLoss<V>(conf_.loss())) {}
  virtual ~AsyncSGDWorker() { stopExecutor(); }

  void process(const MessagePtr& msg) override {
    const SGDCallView call = msg->task.sgd();
    if (call.cmd() == SGDCallView::UPDATE_MODEL) train(call.data());
  }

 private:
  // ref async_sgd.h:197-239
  void train(const DataConfigView& my_files) {
    const SGDConfigView sgd = conf_.async_sgd();
    MinibatchSource<V> source;
    source.open(epochs(my_files, std::max(sgd.num_data_pass(), 1)), sgd.minibatch(), sgd.data_buf());
    source.dropRareKeys((size_t)sgd.countmin_n(), sgd.countmin_k(), sgd.tail_feature_freq());
    source.start();
    // an unset max_delay keeps the reference's behaviour: bounded by the prefetch queue only
    InFlight window(sgd.has_max_delay() ? sgd.max_delay() : -1);
    window_ = &window;
    Minibatch<V> mb;
    for (int id = 0; source.next(&mb); ++id) {
      window.enter();
      { Lock l(mu_); waiting_[id] = mb; }
      requestWeights(id, mb.keys);
    }
    window.drain();
    window_ = nullptr;
  }

  // the files of one pass, visited in a fresh random order in every pass
  DataConfig epochs(const DataConfigView& files, int passes) {
    DataConfig all(files);
    all.clear_file();
    std::vector<int> order(files.file_size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
    std::mt19937 rng(1234 + myRank());
    for (int p = 0; p < passes; ++p) {
      std::shuffle(order.begin(), order.end(), rng);
      for (int i : order) all.add_file(files.file(i));
    }
    return all;
  }

  // stage 1: pull the weights of the minibatch's keys on channel `id`
  void requestWeights(int id, const SArray<Key>& keys) {
    MessagePtr ask(new Message(kServerGroup));
    ask->setKey(keys);
    ask->task.set_key_channel(id);
    ask->addFilter(FilterConfigView::KEY_CACHING);
    ask->fin_handle = [this, id]() { answerWithGradient(id); };
    model_.key(id) = keys;
    model_.pull(ask);
  }

  // stage 2, on the model customer's executor thread once every server answered: margins,
  // progress report, gradient, push.  ref async_sgd.h:241-289
  void answerWithGradient(int id) {
    Minibatch<V> mb;
    {
      Lock l(mu_);
      auto it = waiting_.find(id);
      mb = it->second;
      waiting_.erase(it);
    }
    CHECK_EQ(mb.X->rows(), mb.Y->rows());
    SArray<V> w = model_.value(id);
    if (w.size() != mb.X->cols()) w = SArray<V>(mb.X->cols(), 0);  // nothing came back: all keys are new
    const SArray<V> margin = mb.X->times(w);

    SGDProgress seen;
    seen.add_objective(loss_->evaluate(mb.Y->value(), margin));
    seen.add_auc(Evaluation<V>::auc(mb.Y->value(), margin));
    seen.add_accuracy(Evaluation<V>::accuracy(mb.Y->value(), margin));
    seen.set_num_examples_processed(margin.size());
    reporter_.report(seen);

    SArray<V> grad(mb.X->cols(), 0);
    loss_->compute(mb.Y->value(), mb.X, margin, grad);
    MessagePtr give(new Message(kServerGroup));
    give->setKey(model_.key(id));
    give->addValue(grad);
    give->task.set_key_channel(id);
    give->addFilter(FilterConfigView::KEY_CACHING).set_clear_cache_if_done(true);
    if (const int nbytes = conf_.async_sgd().fixing_float_by_nbytes())
      give->addFilter(FilterConfigView::FIXING_FLOAT).add_fixed_point().set_num_bytes(nbytes);
    model_.push(give);
    model_.clear(id);
    window_->leave();
  }

  Config conf_;
  KVVector<Key, V> model_;
  LossPtr<V> loss_;
  std::mutex mu_;
  std::unordered_map<int, Minibatch<V>> waiting_;
  InFlight* window_ = nullptr;
};

}  // namespace LM
}  // namespace PS
