// Model evaluation: load a text model, stream the validation data, report AUC / accuracy.
// ref src/app/linear_method/model_evaluation.h:9-74 (runs on the scheduler alone,
// -num_servers 0 -num_workers 0)
#pragma once
#include "app/linear_method/darlin.h"
#include "data/slot_reader.h"
#include "util/localizer.h"

namespace PS {
namespace LM {

class ModelEvaluation : public App {
 public:
  ModelEvaluation(const string& name, const ConfigView& conf) : App(name), conf_(conf) {}
  virtual ~ModelEvaluation() { stopExecutor(); }
  void init() override {}
  void run() override {
    typedef double Real;
    // model: whitespace separated "key value" pairs from every file matching model_input
    DataConfig model = searchFiles(conf_.model_input());
    std::unordered_map<Key, Real> weight;
    for (int i = 0; i < model.file_size(); ++i) {
      std::ifstream in(model.file(i));
      if (!in.good()) continue;
      Key k;
      Real v;
      while (in >> k >> v) weight[k] = v;
    }
    LI << "load " << weight.size() << " model entries";

    DataConfig data = searchFiles(conf_.validation_data());
    data.set_ignore_feature_group(true);
    LI << "find " << data.file_size() << " data files";
    StreamReader<Real> reader(data);
    SArray<Real> label, predict;
    MatrixPtrList<Real> mats;
    while (reader.readMatrices(100000, &mats)) {
      if (mats.size() < 2) break;
      label.append(mats[0]->value());
      auto X = std::static_pointer_cast<SparseMatrix<Key, Real>>(mats[1]);
      const size_t* off = X->offset().data();
      const Key* idx = X->index().data();
      const bool bin = X->binary();
      SArray<Real> Xw(X->rows());
      for (size_t r = 0; r < X->rows(); ++r) {
        Real s = 0;
        for (size_t o = off[r]; o < off[r + 1]; ++o) {
          auto it = weight.find(idx[o]);
          if (it != weight.end()) s += bin ? it->second : it->second * X->value()[o];
        }
        Xw[r] = s;
      }
      predict.append(Xw);
    }
    auc_ = Evaluation<Real>::auc(label, predict);
    accuracy_ = Evaluation<Real>::accuracy(label, predict);
    num_ex_ = label.size();
    LI << "evaluated " << num_ex_ << " examples";
    LI << "auc: " << auc_;
    LI << "accuracy: " << accuracy_;
    LI << "logloss: " << (num_ex_ ? Evaluation<Real>::logloss(label, predict) / num_ex_ : 0);
  }
  double auc() const { return auc_; }
  double accuracy() const { return accuracy_; }
  size_t numExamples() const { return num_ex_; }

 private:
  Config conf_;
  double auc_ = 0, accuracy_ = 0;
  size_t num_ex_ = 0;
};

}  // namespace LM
}  // namespace PS
