// Message filters: key caching, lossless compression, lossy fixed-point values, and the
// NaN "sparse" marker used by Darlin's KKT filter.
// ref src/filter/filter.{h,cc}, key_caching.h, compressing.h, fixing_float.h, sparse_filter.h
//
// Encode order on send is the order of task.filter with COMPRESSING appended last when
// -message_compression is on; decode runs in reverse (ref src/system/remote_node.cc:177-191).
// On the device path these become: a 64-bit key-set signature computed once (K16), and
// block-scaled int8 quantisation fused into the push kernel (K9); snappy stays host-only.
#pragma once
#include "system/message.h"
#include "util/hash.h"

namespace PS {

class Filter {
 public:
  virtual ~Filter() {}
  static Filter* create(const FilterConfigView& conf);
  virtual void encode(const MessagePtr& msg) {}
  virtual void decode(const MessagePtr& msg) {}

  // the config of `type` inside the task header, or an invalid view
  static FilterConfigView find(FilterConfigView::Type type, const MessagePtr& msg) {
    return find(type, &msg->task);
  }
  static FilterConfigView find(FilterConfigView::Type type, Task* task) {
    for (int i = 0; i < task->filter_size(); ++i)
      if (task->filter(i).type() == type) return task->mutable_filter(i);
    return FilterConfigView();
  }
};

// ---- KEY_CACHING -------------------------------------------------------------------------------
// Both ends remember the last key array per (channel, key range).  If the sender's array has
// the same signature and size as the remembered one, only the signature travels.
// Deviation: the signature covers the whole array, not just its first 2 KB
// (ref key_caching.h:18,74), so two key sets sharing a prefix can never alias.
class KeyCachingFilter : public Filter {
 public:
  void encode(const MessagePtr& msg) override {
    FilterConfigView conf = find(FilterConfigView::KEY_CACHING, msg);
    if (!conf.valid()) return;
    if (!msg->task.has_key()) {
      conf.clear_signature();
      return;
    }
    const SArray<char>& key = msg->key;
    uint32 sig = crc32c::Value(key.data(), key.size());
    conf.set_signature(sig);
    auto ck = cacheKey(msg->task);
    Lock l(mu_);
    auto& cache = cache_[ck];
    bool hit = cache.first == sig && cache.second.size() == key.size() && !key.empty();
    if (hit) {
      msg->clearKey();
    } else {
      cache.first = sig;
      cache.second = key;
    }
    if (conf.clear_cache_if_done() && isDone(msg->task)) cache_.erase(ck);
  }

  void decode(const MessagePtr& msg) override {
    FilterConfigView conf = find(FilterConfigView::KEY_CACHING, msg);
    if (!conf.valid() || !conf.has_signature()) return;
    uint32 sig = (uint32)conf.signature();
    if (msg->task.has_key())
      CHECK_EQ(crc32c::Value(msg->key.data(), msg->key.size()), sig) << "corrupt key array";
    auto ck = cacheKey(msg->task);
    Lock l(mu_);
    auto& cache = cache_[ck];
    if (msg->task.has_key()) {
      cache.first = sig;
      cache.second = msg->key;
    } else {
      CHECK_EQ(sig, cache.first) << "key cache miss: " << msg->debugString();
      msg->key = cache.second;
      msg->task.set_has_key(true);
    }
    if (conf.clear_cache_if_done() && isDone(msg->task)) cache_.erase(ck);
  }

  size_t cachedArrays() {
    Lock l(mu_);
    return cache_.size();
  }

 private:
  typedef std::pair<int, Range<Key>> CacheKey;
  static CacheKey cacheKey(const TaskView& t) {
    return std::make_pair(t.key_channel(), t.has_key_range() ? t.key_range().to<Key>()
                                                             : Range<Key>::all());
  }
  static bool isDone(const TaskView& t) {
    return !t.request() ||
           (t.has_shared_para() && t.shared_para().cmd() == CallSharedParaView::PUSH);
  }
  std::unordered_map<CacheKey, std::pair<uint32, SArray<char>>> cache_;
  std::mutex mu_;
};

// ---- COMPRESSING ---------------------------------------------------------------------------------
// Snappy block format on the key and on every value array; original sizes ride in the header.
class CompressingFilter : public Filter {
 public:
  void encode(const MessagePtr& msg) override {
    FilterConfigView conf = find(FilterConfigView::COMPRESSING, msg);
    if (!conf.valid()) return;
    conf.clear_uncompressed_size();
    if (msg->task.has_key()) {
      conf.add_uncompressed_size(msg->key.size());
      msg->key = msg->key.compressTo();
    }
    for (auto& v : msg->value) {
      conf.add_uncompressed_size(v.size());
      v = v.compressTo();
    }
  }
  void decode(const MessagePtr& msg) override {
    FilterConfigView conf = find(FilterConfigView::COMPRESSING, msg);
    if (!conf.valid()) return;
    int has_key = msg->task.has_key();
    CHECK_EQ((size_t)conf.uncompressed_size_size(), msg->value.size() + has_key);
    if (has_key) {
      SArray<char> raw;
      raw.uncompressFrom(msg->key);
      CHECK_EQ(raw.size(), (size_t)conf.uncompressed_size(0));
      msg->key = raw;
    }
    for (size_t i = 0; i < msg->value.size(); ++i) {
      SArray<char> raw;
      raw.uncompressFrom(msg->value[i]);
      CHECK_EQ(raw.size(), (size_t)conf.uncompressed_size((int)i + has_key));
      msg->value[i] = raw;
    }
  }
};

// ---- FIXING_FLOAT ----------------------------------------------------------------------------------
// Lossy fixed point: x in [min,max] -> nbytes-wide integer code, little endian.
// ratio = 2^(8 nbytes) - 2 as in the reference (fixing_float.h:48-64); min/max default to the
// array's range (max + 1e-6) and travel in the header, so the receiver is stateless.
// Deviation: rounding is true stochastic rounding floor(t + u), u ~ U[0,1), which is unbiased;
// the reference adds a fair coin to floor(t) whatever the fraction (fixing_float.h:75).
class FixingFloatFilter : public Filter {
 public:
  void encode(const MessagePtr& msg) override { convert(msg, true); }
  void decode(const MessagePtr& msg) override { convert(msg, false); }

  template <typename V>
  static SArray<char> encodeArray(const SArray<V>& orig, int nbytes, double min_v, double max_v,
                                  uint64* rng_state) {
    CHECK_GT(nbytes, 0);
    CHECK_LT(nbytes, 8);
    const double ratio = (double)(1ULL << (nbytes * 8)) - 2;
    const double bin = max_v - min_v;
    CHECK_GT(bin, 0);
    SArray<uint8> code(orig.size() * nbytes);
    uint8* p = code.data();
    uint64 s = *rng_state ? *rng_state : 0x853c49e6748fea9bULL;
    for (size_t i = 0; i < orig.size(); ++i) {
      double x = orig[i];
      double proj = x > max_v ? max_v : (x < min_v ? min_v : x);
      double t = (proj - min_v) / bin * ratio;
      s ^= s << 13; s ^= s >> 7; s ^= s << 17;  // xorshift64
      double u = (double)(s >> 11) * (1.0 / 9007199254740992.0);
      uint64 r = (uint64)floor(t + u);
      for (int j = 0; j < nbytes; ++j) {
        *p++ = (uint8)(r & 0xff);
        r >>= 8;
      }
    }
    *rng_state = s;
    return SArray<char>(code);
  }
  template <typename V>
  static SArray<char> decodeArray(const SArray<char>& array, int nbytes, double min_v,
                                  double max_v) {
    const double ratio = (double)(1ULL << (nbytes * 8)) - 2;
    const double bin = max_v - min_v;
    const uint8* p = reinterpret_cast<const uint8*>(array.data());
    SArray<V> orig(array.size() / nbytes);
    for (size_t i = 0; i < orig.size(); ++i) {
      uint64 r = 0;
      for (int j = 0; j < nbytes; ++j) r |= (uint64)(*p++) << (8 * j);
      orig[i] = (V)((double)r / ratio * bin + min_v);
    }
    return SArray<char>(orig);
  }

 private:
  void convert(const MessagePtr& msg, bool enc) {
    FilterConfigView conf = find(FilterConfigView::FIXING_FLOAT, msg);
    if (!conf.valid()) return;
    const int n = (int)msg->value.size();
    CHECK_EQ(n, msg->task.value_type_size());
    int k = 0;
    for (int i = 0; i < n; ++i) {
      DataType type = msg->task.value_type(i);
      if (type != FLOAT && type != DOUBLE) continue;
      if (msg->value[i].empty()) continue;
      while (conf.fixed_point_size() <= k) {
        CHECK(enc) << "missing fixed_point config for value " << i;
        conf.add_fixed_point();
      }
      FixedFloatConfigView fc = conf.mutable_fixed_point(k++);
      if (type == FLOAT) msg->value[i] = convertT<float>(msg->value[i], enc, fc);
      else msg->value[i] = convertT<double>(msg->value[i], enc, fc);
    }
  }
  template <typename V>
  SArray<char> convertT(const SArray<char>& array, bool enc, FixedFloatConfigView fc) {
    const int nbytes = fc.num_bytes();
    if (enc) {
      SArray<V> orig(array);
      if (!fc.has_min_value()) fc.set_min_value(*std::min_element(orig.begin(), orig.end()));
      if (!fc.has_max_value()) fc.set_max_value(*std::max_element(orig.begin(), orig.end()) + 1e-6);
      return encodeArray<V>(orig, nbytes, fc.min_value(), fc.max_value(), &rng_);
    }
    CHECK(fc.has_min_value() && fc.has_max_value());
    return decodeArray<V>(array, nbytes, fc.min_value(), fc.max_value());
  }
  uint64 rng_ = 0x9e3779b97f4a7c15ULL;
};

// ---- sparse marker -----------------------------------------------------------------------------------
// mark(v) writes the all-ones bit pattern (a NaN); marked(v) <=> v != v.
class SparseFilter {
 public:
  // the all-ones bit pattern: a NaN no arithmetic produces
  static void mark(float* v) { memset(v, 0xff, sizeof(float)); }
  static void mark(double* v) { memset(v, 0xff, sizeof(double)); }
  static bool marked(double v) { return v != v; }
  static bool marked(float v) { return v != v; }
};

inline Filter* Filter::create(const FilterConfigView& conf) {
  switch (conf.type()) {
    case FilterConfigView::KEY_CACHING: return new KeyCachingFilter();
    case FilterConfigView::COMPRESSING: return new CompressingFilter();
    case FilterConfigView::FIXING_FLOAT: return new FixingFloatFilter();
    default: CHECK(false) << "unknown filter type"; return nullptr;
  }
}

}  // namespace PS
