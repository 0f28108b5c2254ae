// Bitmap, CountMin sketch, Bloom filters.
// ref src/util/bitmap.h:13-97 (Darlin active set), src/util/sketch.h:7-34 +
// src/util/countmin.h:7-48 (tail-feature filter), src/util/bloom_filter.h:5-47,
// src/util/block_bloom_filter.h:10-69 (cache-line blocked variant).
// The CountMin hash and probe sequence are bit-identical to the device kernels in
// csrc/cuda/sketch.cu so host and device sketches can be merged.
#pragma once
#include "util/hash.h"
#include "util/shared_array.h"

namespace PS {

class Bitmap {
 public:
  Bitmap() {}
  explicit Bitmap(uint32 size, bool value = false) { resize(size, value); }
  void resize(uint32 size, bool value = false) {
    size_ = size;
    map_.assign((size + 31) / 32, 0);
    fill(value);
  }
  void clear() { map_.clear(); size_ = 0; }
  void set(uint32 i) { map_[i >> 5] |= 1u << (i & 31); }
  void clear(uint32 i) { map_[i >> 5] &= ~(1u << (i & 31)); }
  bool test(uint32 i) const { return (map_[i >> 5] >> (i & 31)) & 1u; }
  bool operator[](uint32 i) const { return test(i); }
  void fill(bool value) {
    std::fill(map_.begin(), map_.end(), value ? 0xffffffffu : 0u);
    if (value && (size_ & 31) && !map_.empty()) map_.back() = (1u << (size_ & 31)) - 1u;
  }
  uint32 size() const { return size_; }
  size_t memSize() const { return map_.size() * 4; }
  uint32 nnz() const {
    uint32 n = 0;
    for (uint32 w : map_) n += __builtin_popcount(w);
    return n;
  }
  const uint32* words() const { return map_.data(); }

 private:
  std::vector<uint32> map_;
  uint32 size_ = 0;
};

class Sketch {
 public:
  // 32-bit mix of a 64-bit key: two rounds of (add word, multiply, xor-shift) over the low and
  // high halves.  Kept value-compatible with the reference's Sketch::hash (sketch.h:20-33) so
  // a CountMin of the same (n,k) keeps exactly the same keys under the tail filter.
  static uint32 hash(uint64 key) {
    const uint32 m = 0xc6a4a793u;
    uint32 h = 0xbc9f1d34u ^ (8u * m);
    h += (uint32)key;
    h *= m;
    h ^= h >> 16;
    h += (uint32)(key >> 32);
    h *= m;
    h ^= h >> 16;
    return h;
  }
};

// k probes by double hashing h += rot17(h); counters saturate at v_max
template <typename K, typename V>
class CountMin : public Sketch {
 public:
  bool empty() const { return n_ == 0; }
  void clear() { data_.clear(); n_ = 0; }
  void resize(int n, int k, V v_max) {
    n_ = std::max(n, 64);
    k_ = std::min(30, std::max(1, k));
    v_max_ = v_max;
    data_.resize(n_);
    data_.setZero();
  }
  void bulkInsert(const SArray<K>& key, const SArray<V>& count) {
    CHECK_EQ(key.size(), count.size());
    for (size_t i = 0; i < key.size(); ++i) insert(key[i], count[i]);
  }
  void insert(const K& key, const V& count) {
    CHECK_GT(n_, 0) << "CountMin::insert before resize()";
    uint32 h = hash(key);
    const uint32 delta = (h >> 17) | (h << 15);
    for (int j = 0; j < k_; ++j) {
      V& c = data_[h % n_];
      c = count > v_max_ - c ? v_max_ : c + count;  // saturating add
      h += delta;
    }
  }
  V query(const K& key) const {
    V res = 0;
    if (n_ == 0) return res;  // nothing was ever counted (a query may overtake every insert)
    uint32 h = hash(key);
    const uint32 delta = (h >> 17) | (h << 15);
    for (int j = 0; j < k_; ++j) {
      V v = data_[h % n_];
      res = j == 0 ? v : std::min(res, v);
      h += delta;
    }
    return res;
  }
  SArray<V>& data() { return data_; }
  int n() const { return n_; }
  int k() const { return k_; }

 private:
  SArray<V> data_;
  int n_ = 0;
  int k_ = 1;
  V v_max_ = 0;
};

template <typename K>
class BloomFilter : public Sketch {
 public:
  BloomFilter() {}
  BloomFilter(int m, int k) { resize(m, k); }
  void resize(int m, int k) {
    m_ = std::max(m, 64);
    k_ = std::max(1, k);
    data_.assign((m_ + 7) / 8, 0);
  }
  void insert(K key) {
    uint32 h = hash(key);
    const uint32 delta = (h >> 17) | (h << 15);
    for (int j = 0; j < k_; ++j) {
      uint32 b = h % m_;
      data_[b >> 3] |= 1u << (b & 7);
      h += delta;
    }
  }
  bool query(K key) const {
    uint32 h = hash(key);
    const uint32 delta = (h >> 17) | (h << 15);
    for (int j = 0; j < k_; ++j) {
      uint32 b = h % m_;
      if (!(data_[b >> 3] & (1u << (b & 7)))) return false;
      h += delta;
    }
    return true;
  }
  bool operator[](K key) const { return query(key); }

 private:
  std::vector<uint8> data_;
  uint32 m_ = 0;
  int k_ = 0;
};

// all k probes of a key land in one 64-byte block: one cache line per query
template <typename K>
class BlockBloomFilter : public Sketch {
 public:
  BlockBloomFilter() {}
  BlockBloomFilter(size_t n_bits, int k) { resize(n_bits, k); }
  void resize(size_t n_bits, int k) {
    k_ = std::max(1, k);
    n_blocks_ = std::max<size_t>(1, (n_bits + 511) / 512);
    void* p = nullptr;
    CHECK_EQ(posix_memalign(&p, 64, n_blocks_ * 64), 0);
    memset(p, 0, n_blocks_ * 64);
    data_.reset(static_cast<uint8*>(p), [](uint8* q) { free(q); });
  }
  void insert(K key) {
    uint32 h = hash(key);
    uint8* blk = data_.get() + (size_t)(h % n_blocks_) * 64;
    uint32 g = (h >> 17) | (h << 15);
    const uint32 delta = fmix32(h) | 1u;
    for (int j = 0; j < k_; ++j) {
      uint32 b = g & 511;
      blk[b >> 3] |= 1u << (b & 7);
      g += delta;
    }
  }
  bool query(K key) const {
    uint32 h = hash(key);
    const uint8* blk = data_.get() + (size_t)(h % n_blocks_) * 64;
    uint32 g = (h >> 17) | (h << 15);
    const uint32 delta = fmix32(h) | 1u;
    for (int j = 0; j < k_; ++j) {
      uint32 b = g & 511;
      if (!(blk[b >> 3] & (1u << (b & 7)))) return false;
      g += delta;
    }
    return true;
  }
  bool operator[](K key) const { return query(key); }

 private:
  shared_ptr<uint8> data_;
  size_t n_blocks_ = 0;
  int k_ = 0;
};

}  // namespace PS
