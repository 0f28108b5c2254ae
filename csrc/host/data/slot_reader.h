// SlotReader (batch, columnar, cached) and StreamReader (streaming minibatches).
// ref src/data/slot_reader.{h,cc}, src/data/stream_reader.h
//
// SlotReader parses every input file once into per-slot column arrays
//   value (float32) / colidx (uint64) / rowsiz (uint16 entries per row)
// and stores them snappy-compressed next to a text `.info`, exactly the cache layout of the
// reference (slot_reader.cc:18-21,78-87,136-143), so a second run skips the text parsing.
// On the GPU path these arrays are what gets copied (pinned, H2D) into the device localizer.
#pragma once
#include "data/common.h"
#include "data/text_parser.h"
#include "util/concurrent.h"
#include "util/matrix.h"

namespace PS {

DECLARE_int32(num_threads);
DECLARE_bool(verbose);

class SlotReader {
 public:
  SlotReader() {}
  SlotReader(const DataConfigView& data, const DataConfigView& cache) { init(data, cache); }

  void init(const DataConfigView& data, const DataConfigView& cache) {
    // TEXT as in the reference (slot_reader.cc:14), plus PROTO: RecordIO files of Example protos
    CHECK(data.format() == DataConfigView::TEXT || data.format() == DataConfigView::PROTO)
        << "SlotReader reads TEXT or PROTO data";
    if (cache.file_size()) {
      cache_ = cache.file(0);
      string dir = cache_.back() == '/' ? cache_ : getPath(cache_);
      createDir(dir);
    }
    data_ = searchFiles(data);
    num_ex_.assign(data_.file_size(), 0);
    loaded_file_count_ = 0;
    info_ = ExampleInfo();
    slot_info_.clear();
  }

  // parse (or find the cache of) every file; *info receives the merged statistics
  int read(ExampleInfo* info = nullptr) {
    CHECK_GT(FLAGS_num_threads, 0);
    {
      ThreadPool pool(FLAGS_num_threads);
      for (int i = 0; i < data_.file_size(); ++i) {
        DataConfig one = ithFile(data_, i);
        pool.add([this, one, i] { readOneFile(one, i); });
      }
      pool.startWorkers();
    }
    if (info) *info = info_;
    for (int i = 0; i < info_.slot_size(); ++i) slot_info_[info_.slot(i).id()] = info_.slot(i);
    return 0;
  }

  // concatenated arrays of one slot over all files
  SArray<uint64> index(int slot_id) {
    SArray<uint64> idx = index_cache_[slot_id];
    if (!idx.empty()) return idx;
    if (nnzEle(slot_id) == 0 || isDense(slot_id)) return idx;  // dense slots (the label) have no keys
    for (int i = 0; i < data_.file_size(); ++i) {
      SArray<char> comp;
      if (!comp.readFromFile(cacheName(ithFile(data_, i), slot_id) + ".colidx")) continue;
      SArray<uint64> part;
      part.uncompressFrom(comp);
      idx.append(part);
    }
    CHECK_EQ(idx.size(), nnzEle(slot_id));
    index_cache_[slot_id] = idx;
    return idx;
  }
  SArray<size_t> offset(int slot_id) {
    SArray<size_t> os = offset_cache_[slot_id];
    if (!os.empty()) return os;
    if (nnzEle(slot_id) == 0) return os;
    os.pushBack(0);
    for (int i = 0; i < data_.file_size(); ++i) {
      SArray<char> comp;
      if (!comp.readFromFile(cacheName(ithFile(data_, i), slot_id) + ".rowsiz")) continue;
      SArray<uint16> rowsiz;
      rowsiz.uncompressFrom(comp);
      size_t n = os.size();
      os.resize(n + rowsiz.size());
      for (size_t j = 0; j < rowsiz.size(); ++j) os[n + j] = os[n + j - 1] + rowsiz[j];
    }
    CHECK_EQ(os.back(), nnzEle(slot_id));
    offset_cache_[slot_id] = os;
    return os;
  }
  template <typename V>
  SArray<V> value(int slot_id) {
    SArray<V> val;
    if (nnzEle(slot_id) == 0) return val;
    for (int i = 0; i < data_.file_size(); ++i) {
      SArray<char> comp;
      if (!comp.readFromFile(cacheName(ithFile(data_, i), slot_id) + ".value")) continue;
      SArray<float> part;
      part.uncompressFrom(comp);
      size_t n = val.size();
      val.resize(n + part.size());
      for (size_t j = 0; j < part.size(); ++j) val[n + j] = (V)part[j];
    }
    CHECK(val.empty() || val.size() == nnzEle(slot_id)) << val.size() << " vs " << nnzEle(slot_id);
    return val;
  }
  void clear(int slot_id) {
    index_cache_.erase(slot_id);
    offset_cache_.erase(slot_id);
  }
  template <typename V>
  MatrixInfo info(int slot_id) const {
    return readMatrixInfo(info_, slot_id, sizeof(uint64), sizeof(V));
  }
  const ExampleInfo& info() const { return info_; }
  bool isDense(int slot_id) const {
    auto it = slot_info_.find(slot_id);
    return it != slot_info_.end() && it->second.format() == SlotInfoView::DENSE;
  }
  size_t nnzEle(int slot_id) const {
    auto it = slot_info_.find(slot_id);
    return it == slot_info_.end() ? 0 : (size_t)it->second.nnz_ele();
  }

 private:
  string cacheName(const DataConfigView& data, int slot_id) const {
    CHECK_GT(data.file_size(), 0);
    return cache_ + getFilename(data.file(0)) + "_slot_" + to_string(slot_id);
  }

  struct VSlot {
    SArray<float> val;
    SArray<uint64> col_idx;
    SArray<uint16> row_siz;
    bool writeToFile(const string& name) const {
      return val.compressTo().writeToFile(name + ".value") &&
             col_idx.compressTo().writeToFile(name + ".colidx") &&
             row_siz.compressTo().writeToFile(name + ".rowsiz");
    }
  };

  bool readOneFile(const DataConfigView& data, int ith_file) {
    const string info_name = cache_ + getFilename(data.file(0)) + ".info";
    ExampleInfo info;
    if (readFileToProto(info_name, &info)) {  // cache hit
      Lock l(mu_);
      info_ = mergeExampleInfo(info_, info);
      num_ex_[ith_file] = info.num_ex();
      ++loaded_file_count_;
      return true;
    }
    InfoParser info_parser;
    ExampleParser text_parser;
    text_parser.init(data.text(), data.ignore_feature_group());
    std::vector<VSlot> vslots(kSlotIDmax);
    uint32 num_ex = 0;
    Example ex;
    auto take = [&]() {
      if (!info_parser.add(ex)) return;
      for (const Slot& slot : ex.slot) {
        VSlot& vs = vslots[slot.id];
        for (float v : slot.val) vs.val.pushBack(v);
        for (uint64 k : slot.key) vs.col_idx.pushBack(k);
        while (vs.row_siz.size() < num_ex) vs.row_siz.pushBack(0);
        vs.row_siz.pushBack((uint16)std::max(slot.key.size(), slot.val.size()));
      }
      ++num_ex;
    };
    if (data.format() == DataConfigView::PROTO) {
      File* f = File::open(data.file(0), "r");
      if (!f) {
        LL << "cannot read " << data.file(0);
        return false;
      }
      RecordReader records(f);
      while (records.ReadProtocolMessage(&ex)) take();
      f->close();
    } else {
      FileLineReader reader(data);
      reader.set_line_callback([&](char* line) {
        if (text_parser.toProto(line, &ex)) take();
      });
      reader.Reload();
      if (!reader.loaded_successfully()) {
        LL << "cannot read " << data.file(0);
        return false;
      }
    }
    info = info_parser.info();
    for (int i = 0; i < kSlotIDmax; ++i) {
      VSlot& vs = vslots[i];
      if (vs.row_siz.empty() && vs.val.empty()) continue;
      while (vs.row_siz.size() < num_ex) vs.row_siz.pushBack(0);
      CHECK(vs.writeToFile(cacheName(data, i))) << "cannot write the slot cache " << cacheName(data, i);
    }
    writeProtoToASCIIFile(info, info_name);
    Lock l(mu_);
    info_ = mergeExampleInfo(info_, info);
    num_ex_[ith_file] = num_ex;
    ++loaded_file_count_;
    if (FLAGS_verbose) LI << "loaded " << data.file(0) << " (" << num_ex << " examples)";
    return true;
  }

  string cache_;
  DataConfig data_;
  std::mutex mu_;
  size_t loaded_file_count_ = 0;
  std::vector<uint32> num_ex_;
  ExampleInfo info_;
  std::unordered_map<int, SlotInfo> slot_info_;
  std::unordered_map<int, SArray<uint64>> index_cache_;
  std::unordered_map<int, SArray<size_t>> offset_cache_;
};

// ---- StreamReader -----------------------------------------------------------------------------------
// reads `num_examples` at a time from TEXT or PROTO (RecordIO) files and returns one matrix per
// slot: slot 0 (label) dense, the others row-major sparse with uint64 column indices
template <typename V>
class StreamReader {
 public:
  StreamReader() {}
  explicit StreamReader(const DataConfigView& data) { init(data); }
  ~StreamReader() { if (data_file_) data_file_->close(); }

  void init(const DataConfigView& data) {
    data_ = searchFiles(data);
    if (data_.format() == DataConfigView::TEXT)
      text_parser_.init(data_.text(), data_.ignore_feature_group());
    next_file_ = 0;
    max_num_files_ = data_.file_size();
    if (data_.max_num_files_per_worker() >= 0)
      max_num_files_ = std::min(max_num_files_, data_.max_num_files_per_worker());
    reach_data_end_ = false;
    openNextFile();
  }

  // false once the data is exhausted and nothing was read
  bool readMatrices(uint32 num_examples, MatrixPtrList<V>* matrices,
                    std::vector<Example>* examples = nullptr) {
    if (reach_data_end_) return false;
    parser_.clear();
    vslots_.assign(kSlotIDmax, VSlot());
    num_ex_ = 0;
    if (examples) examples->clear();
    bool got_all = readExamples(num_examples, examples);
    if (matrices) fillMatrices(matrices);
    return got_all || num_ex_ > 0;
  }

  struct LabeledBatch {  // convenience for SGD: label vector + one feature matrix (first group)
    MatrixPtr<V> Y, X;
  };

 private:
  struct VSlot {
    SArray<V> val;
    SArray<uint64> col_idx;
    SArray<uint16> row_siz;
  };

  bool openNextFile() {
    if (data_file_) { data_file_->close(); data_file_ = nullptr; }
    while (true) {
      if (next_file_ >= max_num_files_) {
        reach_data_end_ = true;
        return false;
      }
      data_file_ = File::open(ithFile(data_, next_file_++), "r");
      if (data_file_) return true;
      LL << "cannot open " << data_.file(next_file_ - 1) << ", skipped";
    }
  }

  void parseExample(const Example& ex, int num_read) {
    if (!parser_.add(ex)) return;
    for (const Slot& slot : ex.slot) {
      VSlot& vs = vslots_[slot.id];
      for (float v : slot.val) vs.val.pushBack((V)v);
      for (uint64 k : slot.key) vs.col_idx.pushBack(k);
      while ((int)vs.row_siz.size() < num_read) vs.row_siz.pushBack(0);
      vs.row_siz.pushBack((uint16)std::max(slot.key.size(), slot.val.size()));
    }
  }

  bool readExamples(uint32 num_ex, std::vector<Example>* examples) {
    const int kMaxLine = 60 * 1024;
    std::vector<char> line(kMaxLine);
    Example ex;
    const int line_limit = data_.max_num_lines_per_file();
    while (num_ex_ < num_ex) {
      if (!data_file_) return false;
      bool ok = false;
      if (data_.format() == DataConfigView::TEXT) {
        if (line_limit >= 0 && lines_in_file_ >= line_limit) {
          ok = false;
        } else if (data_file_->readLine(line.data(), kMaxLine)) {
          ++lines_in_file_;
          size_t n = strlen(line.data());
          while (n && (line[n - 1] == '\n' || line[n - 1] == '\r')) line[--n] = 0;
          if (!text_parser_.toProto(line.data(), &ex)) continue;
          ok = true;
        }
      } else if (data_.format() == DataConfigView::PROTO) {
        RecordReader rr(data_file_);
        ok = rr.ReadProtocolMessage(&ex);
      } else {
        CHECK(false) << "unsupported data format for streaming";
      }
      if (!ok) {
        lines_in_file_ = 0;
        if (!openNextFile()) return false;
        continue;
      }
      parseExample(ex, num_ex_);
      if (examples) examples->push_back(ex);
      ++num_ex_;
    }
    return true;
  }

  void fillMatrices(MatrixPtrList<V>* mats) {
    mats->clear();
    ExampleInfo info = parser_.info();
    for (int i = 0; i < info.slot_size(); ++i) {
      SlotInfoView sinfo = info.slot(i);
      int id = sinfo.id();
      VSlot& vs = vslots_[id];
      MatrixInfo minfo = readMatrixInfo(info, id, sizeof(uint64), sizeof(V));
      if (sinfo.format() == SlotInfoView::DENSE) {
        mats->push_back(MatrixPtr<V>(new DenseMatrix<V>(minfo, vs.val)));
        continue;
      }
      while (vs.row_siz.size() < num_ex_) vs.row_siz.pushBack(0);
      SArray<size_t> offset(vs.row_siz.size() + 1);
      offset[0] = 0;
      for (size_t j = 0; j < vs.row_siz.size(); ++j) offset[j + 1] = offset[j] + vs.row_siz[j];
      mats->push_back(MatrixPtr<V>(new SparseMatrix<uint64, V>(minfo, offset, vs.col_idx, vs.val)));
    }
  }

  DataConfig data_;
  int next_file_ = 0, max_num_files_ = 0, lines_in_file_ = 0;
  File* data_file_ = nullptr;
  bool reach_data_end_ = false;
  ExampleParser text_parser_;
  InfoParser parser_;
  std::vector<VSlot> vslots_;
  uint32 num_ex_ = 0;
};

}  // namespace PS
