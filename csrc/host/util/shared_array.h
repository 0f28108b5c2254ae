// SArray<V>: a reference-counted, typed, zero-copy view of a byte buffer.
//
// Same role and method names as the reference's SArray (ref src/util/shared_array.h:28-184,
// shared_array_inl.h): message payloads, KV containers and matrices all hold SArrays, slices
// (`segment`) and reinterpretations (`SArray<char>(arr)`) never copy.  The buffer is a
// shared_ptr<char> with a custom deleter, so memory handed in from elsewhere (e.g. a pinned
// host staging buffer) can be wrapped without ownership.
#pragma once
#include <random>

#include "proto/messages.h"
#include "util/range.h"
#include "util/snappy.h"

namespace PS {

template <typename V>
class SArray {
 public:
  SArray() {}
  explicit SArray(size_t n) { resize(n); }
  SArray(size_t n, V val) { resize(n, val); }
  // zero-copy reinterpretation of another element type
  template <typename W>
  explicit SArray(const SArray<W>& arr) { *this = arr; }
  template <typename W>
  void operator=(const SArray<W>& arr) {
    size_ = arr.size() * sizeof(W) / sizeof(V);
    capacity_ = arr.capacity() * sizeof(W) / sizeof(V);
    ptr_ = arr.pointer();
    data_ = reinterpret_cast<V*>(arr.data());
    dev_ = arr.onDevice();
  }
  // wrap external memory; deletable=false leaves ownership with the caller
  SArray(V* data, size_t size, bool deletable = true) { reset(data, size, deletable); }
  template <typename W>
  SArray(const std::initializer_list<W>& list) { copyFrom(list.begin(), list.end()); }
  template <typename W>
  void operator=(const std::initializer_list<W>& list) { copyFrom(list.begin(), list.end()); }

  void copyFrom(const V* src, size_t size) {
    resize(size);
    if (size) memcpy(data_, src, size * sizeof(V));
  }
  void copyFrom(const SArray<V>& arr) { if (this != &arr) copyFrom(arr.data(), arr.size()); }
  template <typename It>
  void copyFrom(It first, It last) {
    size_t n = std::distance(first, last);
    resize(n);
    V* p = data_;
    for (It it = first; it != last; ++it) *p++ = static_cast<V>(*it);
  }

  // zero-copy slice
  SArray<V> segment(const Range<size_t>& range) const {
    CHECK(range.valid());
    CHECK_LE(range.end(), size_);
    SArray<V> r;
    r.ptr_ = ptr_;
    r.data_ = data_ + range.begin();
    r.size_ = range.size();
    r.capacity_ = r.size_;
    r.dev_ = dev_;
    return r;
  }

  // A view of DEVICE memory (HBM, possibly another process' buffer mapped with CUDA IPC): only
  // pointer arithmetic (segment, reinterpretation, size) is valid on the host, the bytes are
  // touched by kernels.  `owner` keeps the memory alive; its deleter runs when the last view dies.
  static SArray<V> wrapDevice(const shared_ptr<char>& owner, V* data, size_t n) {
    SArray<V> r;
    r.ptr_ = owner;
    r.data_ = data;
    r.size_ = r.capacity_ = n;
    r.dev_ = true;
    return r;
  }
  bool onDevice() const { return dev_; }

  // both arrays must be sorted and duplicate-free
  SArray<V> setIntersection(const SArray<V>& other) const {
    SArray<V> r(std::min(size_, other.size_));
    V* e = std::set_intersection(begin(), end(), other.begin(), other.end(), r.begin());
    r.size_ = e - r.begin();
    return r;
  }
  SArray<V> setUnion(const SArray<V>& other) const {
    SArray<V> r(size_ + other.size_);
    V* e = std::set_union(begin(), end(), other.begin(), other.end(), r.begin());
    r.size_ = e - r.begin();
    return r;
  }
  // positions [lo,hi) of the sorted entries that fall inside `bound`
  SizeR findRange(const Range<V>& bound) const {
    if (empty()) return SizeR(0, 0);
    CHECK(bound.valid());
    const V* lb = std::lower_bound(begin(), end(), bound.begin());
    const V* ub = std::lower_bound(begin(), end(), bound.end());
    return SizeR(lb - begin(), ub - begin());
  }

  void reset(V* data, size_t size, bool deletable = true) {
    size_ = capacity_ = size;
    data_ = data;
    if (deletable) {
      ptr_ = shared_ptr<char>(reinterpret_cast<char*>(data), [](char* p) { free(p); });
    } else {
      ptr_ = shared_ptr<char>(reinterpret_cast<char*>(data), [](char*) {});
    }
  }
  // keeps the old content, new entries are NOT initialised (like the reference)
  void resize(size_t n) {
    if (n <= capacity_) { size_ = n; return; }
    reserve(std::max(n, capacity_ + capacity_ / 2));
    size_ = n;
  }
  void resize(size_t n, V val) {
    size_t old = size_;
    resize(n);
    for (size_t i = old; i < n; ++i) data_[i] = val;
  }
  void reserve(size_t n) {
    if (n <= capacity_) return;
    void* p = nullptr;
    CHECK_EQ(posix_memalign(&p, 64, std::max<size_t>(n * sizeof(V), 64)), 0);
    if (size_) memcpy(p, data_, size_ * sizeof(V));
    ptr_ = shared_ptr<char>(reinterpret_cast<char*>(p), [](char* q) { free(q); });
    data_ = reinterpret_cast<V*>(p);
    capacity_ = n;
  }
  void clear() {
    ptr_.reset();
    data_ = nullptr;
    size_ = capacity_ = 0;
    dev_ = false;
  }
  void append(const SArray<V>& tail) {
    if (tail.empty()) return;
    size_t old = size_;
    resize(old + tail.size());
    memcpy(data_ + old, tail.data(), tail.size() * sizeof(V));
  }
  void pushBack(const V& val) {
    if (size_ == capacity_) reserve(size_ * 2 + 5);
    data_[size_++] = val;
  }
  void setValue(V value) { std::fill(begin(), end(), value); }
  void setZero() { if (size_) memset(data_, 0, size_ * sizeof(V)); }

  // ZERO / CONSTANT / GAUSSIAN / FILE (ref shared_array_inl.h:103-130; the reference aborts on
  // FILE and, by re-seeding per element, degenerates GAUSSIAN to one value — both fixed here)
  void setValue(const ParameterInitConfigView& cf, uint64 seed = 0) {
    typedef ParameterInitConfigView Cf;
    switch (cf.type()) {
      case Cf::ZERO: setZero(); break;
      case Cf::CONSTANT: setValue((V)cf.constant()); break;
      case Cf::GAUSSIAN: {
        std::mt19937_64 gen(seed ? seed : 0x9e3779b97f4a7c15ULL);
        std::normal_distribution<double> d(cf.mean(), cf.std());
        for (size_t i = 0; i < size_; ++i) data_[i] = (V)d(gen);
        break;
      }
      case Cf::FILE:
        CHECK(readFromFile(SizeR(0, size_), cf.file_name())) << "cannot read " << cf.file_name();
        break;
      default: setZero(); break;
    }
  }

  size_t nnz() const {
    size_t n = 0;
    for (size_t i = 0; i < size_; ++i) n += data_[i] != 0;
    return n;
  }
  template <typename W>
  bool operator==(const SArray<W>& rhs) const {
    if (rhs.size() * sizeof(W) != size_ * sizeof(V)) return false;
    if (size_ == 0) return true;
    return memcmp(data_, rhs.data(), size_ * sizeof(V)) == 0;
  }

  V* data() const { return data_; }
  V* begin() const { return data_; }
  V* end() const { return data_ + size_; }
  V& operator[](size_t i) const { return data_[i]; }
  V front() const { CHECK(!empty()); return data_[0]; }
  V back() const { CHECK(!empty()); return data_[size_ - 1]; }
  size_t size() const { return size_; }
  size_t capacity() const { return capacity_; }
  size_t memSize() const { return size_ * sizeof(V); }
  bool empty() const { return size_ == 0; }
  const shared_ptr<char>& pointer() const { return ptr_; }
  long useCount() const { return ptr_.use_count(); }

  double sum() const { double s = 0; for (size_t i = 0; i < size_; ++i) s += data_[i]; return s; }
  double mean() const { return size_ ? sum() / size_ : 0; }
  double std() const {
    if (!size_) return 0;
    double m = mean(), s = 0;
    for (size_t i = 0; i < size_; ++i) s += (data_[i] - m) * (data_[i] - m);
    return sqrt(s / size_);
  }

  // Snappy block format (ref shared_array_inl.h:225-248)
  SArray<char> compressTo() const {
    SArray<char> dest(snappy::MaxCompressedLength(memSize()));
    size_t n = 0;
    snappy::RawCompress(reinterpret_cast<const char*>(data_), memSize(), dest.data(), &n);
    dest.resize(n);
    return dest;
  }
  void uncompressFrom(const char* src, size_t src_size) {
    size_t n = 0;
    CHECK(snappy::GetUncompressedLength(src, src_size, &n));
    CHECK_EQ(n % sizeof(V), (size_t)0);
    resize(n / sizeof(V));
    CHECK(snappy::RawUncompress(src, src_size, reinterpret_cast<char*>(data_)));
  }
  void uncompressFrom(const SArray<char>& src) { uncompressFrom(src.data(), src.size()); }

  // raw binary file IO (ref shared_array_inl.h:198-223); range is in elements
  bool readFromFile(SizeR range, const string& file_name) {
    FILE* f = fopen(file_name.c_str(), "rb");
    if (!f) return false;
    if (range.empty()) {  // whole file
      fseek(f, 0, SEEK_END);
      range = SizeR(0, ftell(f) / sizeof(V));
    }
    resize(range.size());
    bool ok = fseek(f, range.begin() * sizeof(V), SEEK_SET) == 0 &&
              fread(data_, sizeof(V), size_, f) == size_;
    fclose(f);
    return ok;
  }
  bool readFromFile(const string& file_name) { return readFromFile(SizeR(0, 0), file_name); }
  bool writeToFile(SizeR range, const string& file_name) const {
    if (range.empty()) range = SizeR(0, size_);
    CHECK_LE(range.end(), size_);
    FILE* f = fopen(file_name.c_str(), "wb");
    if (!f) return false;
    bool ok = fwrite(data_ + range.begin(), sizeof(V), range.size(), f) == range.size();
    return fclose(f) == 0 && ok;
  }
  bool writeToFile(const string& file_name) const { return writeToFile(SizeR(0, size_), file_name); }

 private:
  template <typename W> friend class SArray;
  size_t size_ = 0;
  size_t capacity_ = 0;
  V* data_ = nullptr;
  shared_ptr<char> ptr_;
  bool dev_ = false;
};

template <typename V>
std::ostream& operator<<(std::ostream& os, const SArray<V>& a) {
  return os << dbstr(a.data(), (int)a.size(), 10);
}

}  // namespace PS
