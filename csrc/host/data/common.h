// File search / partition helpers and ExampleInfo bookkeeping.
// ref src/data/common.{h,cc}: searchFiles (regex on the basename inside the directory),
// divideFiles (round robin with max_num_files_per_worker), mergeExampleInfo, readMatrixInfo.
#pragma once
#include <random>
#include <regex>

#include "proto/messages.h"
#include "util/file.h"

namespace PS {

inline DataConfig ithFile(const DataConfigView& conf, int i, const string& suffix = "") {
  CHECK_GE(i, 0);
  CHECK_LT(i, conf.file_size());
  DataConfig f(conf);
  f.clear_file();
  f.add_file(conf.file(i) + suffix);
  return f;
}

inline MatrixInfo readMatrixInfo(const ExampleInfoView& info, int slot_id, int sizeof_idx,
                                 int sizeof_val) {
  MatrixInfo f;
  int i = 0;
  for (; i < info.slot_size(); ++i)
    if (info.slot(i).id() == slot_id) break;
  if (i == info.slot_size()) return f;
  SlotInfoView slot = info.slot(i);
  switch (slot.format()) {
    case SlotInfoView::DENSE: f.set_type(MatrixInfoView::DENSE); break;
    case SlotInfoView::SPARSE: f.set_type(MatrixInfoView::SPARSE); break;
    case SlotInfoView::SPARSE_BINARY: f.set_type(MatrixInfoView::SPARSE_BINARY); break;
  }
  f.set_row_major(true);
  f.set_id(slot.id());
  f.mutable_row().set_begin(0);
  f.mutable_row().set_end(info.num_ex());
  f.mutable_col().set_begin(slot.min_key());
  f.mutable_col().set_end(slot.max_key());
  f.set_nnz(slot.nnz_ele());
  f.set_sizeof_index(sizeof_idx);
  f.set_sizeof_value(sizeof_val);
  return f;
}

inline ExampleInfo mergeExampleInfo(const ExampleInfoView& A, const ExampleInfoView& B) {
  std::map<int, SlotInfo> slots;
  for (int i = 0; i < A.slot_size(); ++i) slots[A.slot(i).id()] = A.slot(i);
  for (int i = 0; i < B.slot_size(); ++i) {
    SlotInfoView b = B.slot(i);
    auto it = slots.find(b.id());
    if (it == slots.end()) {
      slots[b.id()] = b;
      continue;
    }
    SlotInfo& a = it->second;
    CHECK_EQ((int)a.format(), (int)b.format());
    a.set_min_key(std::min(a.min_key(), b.min_key()));
    a.set_max_key(std::max(a.max_key(), b.max_key()));
    a.set_nnz_ele(a.nnz_ele() + b.nnz_ele());
    a.set_nnz_ex(a.nnz_ex() + b.nnz_ex());
  }
  ExampleInfo C;
  C.set_num_ex(A.num_ex() + B.num_ex());
  for (const auto& kv : slots) C.add_slot().raw()->copyFrom(*kv.second.raw());
  return C;
}

// expand the regexes in conf.file into the list of matching files
inline DataConfig searchFiles(const DataConfigView& config) {
  int n = config.file_size();
  CHECK_GE(n, 1) << "empty files: " << config.DebugString();
  std::vector<string> matched;
  for (int i = 0; i < n; ++i) {
    const string base = getFilename(config.file(i));
    std::regex pattern;
    try {
      // libstdc++ fills a per-locale narrow() cache while compiling: one compilation at a time
      static std::mutex regex_mu;
      Lock l(regex_mu);
      pattern = std::regex(base);
    } catch (const std::regex_error& e) {
      CHECK(false) << base << " is not a valid regex: " << e.what();
    }
    DataConfig dir(config);
    dir.clear_file();
    dir.add_file(getPath(config.file(i)));
    for (const auto& f : readFilenamesInDirectory(dir)) {
      const string name = getFilename(f);
      if (std::regex_match(name, pattern) || name == base) {
        // BIN matrices are several files sharing a stem; TEXT and PROTO (RecordIO) are one file each
        const bool whole = config.format() == DataConfigView::TEXT || config.format() == DataConfigView::PROTO;
        string l = whole ? name : removeExtension(name);
        matched.push_back(dir.file(0) + "/" + l);
      }
    }
  }
  std::sort(matched.begin(), matched.end());
  matched.erase(std::unique(matched.begin(), matched.end()), matched.end());
  DataConfig ret(config);
  ret.clear_file();
  for (const auto& f : matched) ret.add_file(f);
  return ret;
}

inline std::vector<DataConfig> divideFiles(const DataConfigView& data, int num) {
  std::vector<DataConfig> parts;
  const int limit = data.max_num_files_per_worker();
  for (int i = 0; i < num; ++i) {
    DataConfig dc(data);
    dc.clear_file();
    for (int j = 0; j < data.file_size(); ++j) {
      if (limit >= 0 && dc.file_size() >= limit) break;
      if (j % num == i) dc.add_file(data.file(j));
    }
    parts.push_back(dc);
  }
  return parts;
}

inline DataConfig shuffleFiles(const DataConfigView& data, uint64 seed = 0) {
  DataConfig ret(data);
  ret.clear_file();
  std::vector<int> idx(data.file_size());
  for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int)i;
  std::mt19937_64 gen(seed);
  std::shuffle(idx.begin(), idx.end(), gen);
  for (int i : idx) ret.add_file(data.file(i));
  return ret;
}

}  // namespace PS
