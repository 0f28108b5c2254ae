// Minibatch SGD over the parameter server: the objective-independent parts.
// ref src/learner/sgd.h:14-157, src/learner/proto/sgd.proto
//
//   SGDReportTable    accumulates the SGDProgress reports of workers and servers, prints one
//                     table row per reporting interval
//   ISGDScheduler     hands the files to the workers, owns the monitor
//   ISGDCompNode      a worker or server with a reporter to the scheduler
//   MinibatchSource   a producer thread parses minibatches ahead; next() localises one:
//                     unique keys -> tail filter (persistent CountMin) -> dense local columns
// The device engine keeps the same split (parameter_server_b200/models/async_sgd.py + sgd.cu).
#pragma once
#include "data/slot_reader.h"
#include "parameter/kv.h"
#include "system/monitor.h"
#include "util/localizer.h"

namespace PS {

class SGDReportTable {
 public:
  // reports of one sender between two rows are summed (the reference keeps only the newest
  // report's loss / auc lists, sgd.h:82-86)
  static void accumulate(const SGDProgress& in, SGDProgress* sum) {
    for (int i = 0; i < in.objective_size(); ++i) sum->add_objective(in.objective(i));
    for (int i = 0; i < in.auc_size(); ++i) sum->add_auc(in.auc(i));
    for (int i = 0; i < in.accuracy_size(); ++i) sum->add_accuracy(in.accuracy(i));
    sum->set_num_examples_processed(sum->num_examples_processed() + in.num_examples_processed());
    sum->set_weight_sum(sum->weight_sum() + in.weight_sum());
    sum->set_delta_sum(sum->delta_sum() + in.delta_sum());
    if (in.has_nnz()) sum->set_nnz(in.nnz());
  }
  // one row from everything that arrived since the last one; clears `per_node`
  void row(double seconds, std::unordered_map<NodeID, SGDProgress>* per_node) {
    Stats s;
    for (const auto& kv : *per_node) s.add(kv.second);
    per_node->clear();
    examples_ += s.examples;
    if (!header_done_) {
      fprintf(stderr, " sec  examples    loss      auc   accuracy   |w|_0  updt ratio\n");
      header_done_ = true;
    }
    if (s.examples) {  // rows fed by servers only carry model statistics, no loss
      last_loss_ = s.loss / (double)s.examples;
      last_auc_ = s.mean(s.auc, s.n_auc);
    }
    fprintf(stderr, "%4d  %.2e  %.3e  %.4f  %.4f  %.2e  %.2e\n", (int)seconds, (double)examples_,
            s.examples ? s.loss / (double)s.examples : 0.0, s.mean(s.auc, s.n_auc), s.mean(s.acc, s.n_acc),
            (double)s.nnz, sqrt(s.delta_sq) / sqrt(std::max(s.weight_sq, 1e-300)));
  }
  double lastLoss() const { return last_loss_; }
  double lastAUC() const { return last_auc_; }

 private:
  struct Stats {
    uint64 examples = 0, nnz = 0;
    double loss = 0, auc = 0, acc = 0, weight_sq = 0, delta_sq = 1e-20;
    int n_auc = 0, n_acc = 0;
    void add(const SGDProgress& p) {
      examples += p.num_examples_processed();
      nnz += p.nnz();
      for (int i = 0; i < p.objective_size(); ++i) loss += p.objective(i);
      for (int i = 0; i < p.auc_size(); ++i, ++n_auc) auc += p.auc(i);
      for (int i = 0; i < p.accuracy_size(); ++i, ++n_acc) acc += p.accuracy(i);
      weight_sq += p.weight_sum();
      delta_sq += p.delta_sum();
    }
    static double mean(double sum, int n) { return n ? sum / n : 0; }
  };
  size_t examples_ = 0;
  bool header_done_ = false;
  double last_loss_ = 0, last_auc_ = 0;
};

class ISGDScheduler : public App {
 public:
  explicit ISGDScheduler(const string& name) : App(name), monitor_(this) {}
  virtual ~ISGDScheduler() { stopExecutor(); }
  double lastLoss() const { return table_.lastLoss(); }
  double lastAUC() const { return table_.lastAUC(); }

 protected:
  // every worker gets its share of the files and trains until it has consumed them
  void updateModel(const DataConfigView& data, int report_interval) {
    monitor_.setMerger([](const SGDProgress& in, SGDProgress* sum) { SGDReportTable::accumulate(in, sum); });
    monitor_.setPrinter(report_interval, [this](double sec, std::unordered_map<NodeID, SGDProgress>* got) {
      table_.row(sec, got);
    });
    const std::vector<DataConfig> shares = Postmaster::partitionData(data, sys_.yp().num_workers());
    std::vector<Task> jobs(shares.size());
    for (size_t w = 0; w < shares.size(); ++w) {
      SGDCallView call = jobs[w].mutable_sgd();
      call.set_cmd(SGDCallView::UPDATE_MODEL);
      call.mutable_data().raw()->copyFrom(*shares[w].raw());
    }
    port(kWorkerGroup)->submitAndWait(jobs);
    monitor_.flush();
  }
  void saveModel() {
    Task save;
    save.mutable_sgd().set_cmd(SGDCallView::SAVE_MODEL);
    port(kServerGroup)->submitAndWait(save);
  }

  MonitorMaster<SGDProgress> monitor_;
  SGDReportTable table_;
};

class ISGDCompNode : public App {
 public:
  explicit ISGDCompNode(const string& name) : App(name), reporter_(schedulerID(), this) {}
  virtual ~ISGDCompNode() { stopExecutor(); }

 protected:
  MonitorSlaver<SGDProgress> reporter_;
};

// one localised minibatch: labels, the matrix over local columns, the keys of those columns
template <typename V>
struct Minibatch {
  MatrixPtr<V> Y, X;
  SArray<Key> keys;
};

// ref MinibatchReader (sgd.h:104-157)
template <typename V>
class MinibatchSource {
 public:
  void open(const DataConfigView& files, int rows_per_batch, int buffered_mb = 1000) {
    reader_.init(files);
    rows_ = rows_per_batch;
    ahead_.setCapacity(buffered_mb);
  }
  // keys seen at most `freq` times so far (over all minibatches of this worker) are dropped
  void dropRareKeys(size_t sketch_cells, int probes, int freq) {
    if (freq > 0) seen_.resize((int)std::min<size_t>(sketch_cells, 1u << 30), probes);
    min_freq_ = freq;
  }
  void start() {
    ahead_.startProducer([this](MatrixPtrList<V>* parsed, size_t* bytes) -> bool {
      const bool more = reader_.readMatrices(rows_, parsed);
      for (const auto& m : *parsed) *bytes += m->value().memSize() + m->nnz() * sizeof(uint64);
      return more && !parsed->empty();
    });
  }
  bool next(Minibatch<V>* out) {
    MatrixPtrList<V> parsed;
    if (!ahead_.pop(&parsed)) return false;
    CHECK_GE(parsed.size(), (size_t)2) << "expect a label and one feature group per minibatch";
    out->Y = parsed[0];
    const auto raw = std::static_pointer_cast<SparseMatrix<Key, V>>(parsed[1]);
    Localizer<Key, V> loc;
    SArray<Key> uniq;
    SArray<uint8> count;
    loc.countUniqIndex(raw->index(), &uniq, &count);
    if (min_freq_ > 0) {
      seen_.insertKeys(uniq, count);
      out->keys = seen_.queryKeys(uniq, min_freq_);
    } else {
      out->keys = uniq;
    }
    out->X = loc.remapIndex(raw->info(), raw->offset(), raw->index(), raw->value(), out->keys);
    if (!out->X) out->X = emptyLike(*raw, out->Y->rows(), out->keys.size());
    return true;
  }

 private:
  // all keys filtered: a matrix with the right shape and no entries
  static MatrixPtr<V> emptyLike(const SparseMatrix<Key, V>& raw, size_t rows, size_t cols) {
    MatrixInfo info(raw.info());
    info.set_nnz(0);
    info.mutable_col().set_begin(0);
    info.mutable_col().set_end(cols);
    SArray<size_t> offsets(rows + 1, 0);
    return MatrixPtr<V>(new SparseMatrix<uint32, V>(info, offsets, SArray<uint32>(), SArray<V>()));
  }
  int rows_ = 1000, min_freq_ = 0;
  StreamReader<V> reader_;
  FreqencyFilter<Key, uint8> seen_;
  ProducerConsumer<MatrixPtrList<V>> ahead_;
};

}  // namespace PS
