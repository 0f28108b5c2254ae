// Minibatch SGD skeletons: scheduler, computation node, prefetching minibatch reader.
// ref src/learner/sgd.h:14-157, src/learner/proto/sgd.proto
#pragma once
#include "data/slot_reader.h"
#include "parameter/kv.h"
#include "system/monitor.h"
#include "util/localizer.h"

namespace PS {

class ISGDScheduler : public App {
 public:
  explicit ISGDScheduler(const string& name) : App(name), monitor_(this) {}
  virtual ~ISGDScheduler() { stopExecutor(); }

 protected:
  void saveModel() {
    Task task;
    task.mutable_sgd().set_cmd(SGDCallView::SAVE_MODEL);
    port(kServerGroup)->submitAndWait(task);
  }
  void updateModel(const DataConfigView& data, int report_interval) {
    using namespace std::placeholders;
    monitor_.setMerger(std::bind(&ISGDScheduler::mergeProgress, this, _1, _2));
    monitor_.setPrinter(report_interval, std::bind(&ISGDScheduler::showProgress, this, _1, _2));
    auto conf = Postmaster::partitionData(data, sys_.yp().num_workers());
    std::vector<Task> tasks(conf.size());
    for (size_t i = 0; i < conf.size(); ++i) {
      SGDCallView sgd = tasks[i].mutable_sgd();
      sgd.set_cmd(SGDCallView::UPDATE_MODEL);
      sgd.mutable_data().raw()->copyFrom(*conf[i].raw());
    }
    port(kWorkerGroup)->submitAndWait(tasks);
    monitor_.flush();
  }

  virtual void showProgress(double time, std::unordered_map<NodeID, SGDProgress>* progress) {
    uint64 num_ex = 0, nnz_w = 0;
    SArray<double> objv, auc, acc;
    double weight_sum = 0, delta_sum = 1e-20;
    for (const auto& it : *progress) {
      const SGDProgress& prog = it.second;
      num_ex += prog.num_examples_processed();
      nnz_w += prog.nnz();
      for (int i = 0; i < prog.objective_size(); ++i) objv.pushBack(prog.objective(i));
      for (int i = 0; i < prog.auc_size(); ++i) auc.pushBack(prog.auc(i));
      for (int i = 0; i < prog.accuracy_size(); ++i) acc.pushBack(prog.accuracy(i));
      weight_sum += prog.weight_sum();
      delta_sum += prog.delta_sum();
    }
    progress->clear();
    num_ex_processed_ += num_ex;
    if (show_prog_head_) {
      fprintf(stderr, " sec  examples    loss      auc   accuracy   |w|_0  updt ratio\n");
      show_prog_head_ = false;
    }
    if (num_ex) {  // server-only reports carry model statistics but no loss
      last_loss_ = objv.sum() / (double)num_ex;
      last_auc_ = auc.mean();
    }
    fprintf(stderr, "%4d  %.2e  %.3e  %.4f  %.4f  %.2e  %.2e\n", (int)time,
            (double)num_ex_processed_, num_ex ? objv.sum() / (double)num_ex : 0.0, auc.mean(),
            acc.mean(), (double)nnz_w,
            sqrt(delta_sum) / sqrt(std::max(weight_sum, 1e-300)));
  }

  // reports of one sender between two prints are accumulated (the reference keeps only the
  // newest report's loss / auc lists, sgd.h:82-86)
  virtual void mergeProgress(const SGDProgress& src, SGDProgress* dst) {
    for (int i = 0; i < src.objective_size(); ++i) dst->add_objective(src.objective(i));
    for (int i = 0; i < src.auc_size(); ++i) dst->add_auc(src.auc(i));
    for (int i = 0; i < src.accuracy_size(); ++i) dst->add_accuracy(src.accuracy(i));
    dst->set_num_examples_processed(dst->num_examples_processed() + src.num_examples_processed());
    if (src.has_nnz()) dst->set_nnz(src.nnz());
    dst->set_weight_sum(dst->weight_sum() + src.weight_sum());
    dst->set_delta_sum(dst->delta_sum() + src.delta_sum());
  }

  MonitorMaster<SGDProgress> monitor_;
  size_t num_ex_processed_ = 0;
  bool show_prog_head_ = true;
  double last_loss_ = 0, last_auc_ = 0;
};

class ISGDCompNode : public App {
 public:
  explicit ISGDCompNode(const string& name) : App(name), reporter_(schedulerID(), this) {}
  virtual ~ISGDCompNode() { stopExecutor(); }

 protected:
  MonitorSlaver<SGDProgress> reporter_;
};

// a producer thread parses minibatches ahead of the consumer; read() localizes one minibatch:
// unique keys -> (tail filter) -> dense local columns
template <typename V>
class MinibatchReader {
 public:
  void setReader(const DataConfigView& file, int minibatch_size, int data_buf = 1000) {
    reader_.init(file);
    minibatch_size_ = minibatch_size;
    data_prefetcher_.setCapacity(data_buf);
  }
  void setFilter(size_t n, int k, int freq) {
    if (freq > 0) filter_.resize((int)std::min<size_t>(n, 1u << 30), k);
    key_freq_ = freq;
  }
  void start() {
    data_prefetcher_.startProducer([this](MatrixPtrList<V>* data, size_t* size) -> bool {
      bool ret = reader_.readMatrices(minibatch_size_, data);
      for (const auto& mat : *data) *size += mat->value().memSize() + mat->nnz() * sizeof(uint64);
      return ret && !data->empty();
    });
  }
  bool read(MatrixPtr<V>& Y, MatrixPtr<V>& X, SArray<Key>& key) {
    MatrixPtrList<V> data;
    if (!data_prefetcher_.pop(&data)) return false;
    CHECK_GE(data.size(), (size_t)2) << "expect a label and one feature group per minibatch";
    Y = data[0];
    auto Xs = std::static_pointer_cast<SparseMatrix<Key, V>>(data[1]);
    SArray<Key> uniq_key;
    SArray<uint8> key_cnt;
    Localizer<Key, V> localizer;
    localizer.countUniqIndex(Xs->index(), &uniq_key, &key_cnt);
    if (key_freq_ > 0) {
      filter_.insertKeys(uniq_key, key_cnt);
      key = filter_.queryKeys(uniq_key, key_freq_);
    } else {
      key = uniq_key;
    }
    X = localizer.remapIndex(Xs->info(), Xs->offset(), Xs->index(), Xs->value(), key);
    if (!X) {  // every key was filtered: an empty matrix with the right number of rows
      MatrixInfo info(Xs->info());
      info.set_nnz(0);
      info.mutable_col().set_begin(0);
      info.mutable_col().set_end(key.size());
      SArray<size_t> off(Y->rows() + 1);
      off.setZero();
      X = MatrixPtr<V>(new SparseMatrix<uint32, V>(info, off, SArray<uint32>(), SArray<V>()));
    }
    return true;
  }

 private:
  int minibatch_size_ = 1000;
  StreamReader<V> reader_;
  FreqencyFilter<Key, uint8> filter_;
  int key_freq_ = 0;
  ProducerConsumer<MatrixPtrList<V>> data_prefetcher_;
};

}  // namespace PS
