// Half-open interval [begin, end) over keys or array positions.
// ref src/util/range.h:10-131.  evenDivide uses long double like the reference so that
// server shards / feature blocks land on the same boundaries; unlike the reference the last
// part is closed on `end` (long-double truncation can otherwise drop the largest key).
#pragma once
#include "util/common.h"

namespace PS {

template <class T>
class Range {
 public:
  Range() : begin_(0), end_(0) {}
  Range(T b, T e) : begin_(b), end_(e) {}
  template <typename V>
  Range(const Range<V>& o) : begin_((T)o.begin()), end_((T)o.end()) {}

  static Range all() { return Range(0, (T)-1); }  // [0, 2^64-1) for uint64 keys

  T begin() const { return begin_; }
  T end() const { return end_; }
  T& begin() { return begin_; }
  T& end() { return end_; }
  void set(T b, T e) { begin_ = b; end_ = e; }

  size_t size() const { return end_ > begin_ ? (size_t)(end_ - begin_) : 0; }
  bool valid() const { return end_ >= begin_; }
  bool empty() const { return begin_ >= end_; }
  bool contains(T v) const { return begin_ <= v && v < end_; }
  bool inLeft(const Range& o) const { return begin_ <= o.begin_ || (begin_ == o.begin_ && end_ <= o.end_); }
  bool inRight(const Range& o) const { return !inLeft(o); }

  bool operator==(const Range& r) const { return begin_ == r.begin_ && end_ == r.end_; }
  bool operator!=(const Range& r) const { return !(*this == r); }
  bool operator<(const Range& r) const {
    return begin_ < r.begin_ || (begin_ == r.begin_ && end_ < r.end_);
  }
  Range operator+(T v) const { return Range(begin_ + v, end_ + v); }
  Range operator-(T v) const { return Range(begin_ - v, end_ - v); }
  Range operator*(T v) const { return Range(begin_ * v, end_ * v); }

  Range setIntersection(const Range& o) const {
    return Range(std::max(begin_, o.begin_), std::min(end_, o.end_));
  }
  Range setUnion(const Range& o) const {
    return Range(std::min(begin_, o.begin_), std::max(end_, o.end_));
  }
  // position of this range inside `o` (clipped), relative to o.begin
  Range project(const Range& o) const {
    Range r = setIntersection(o);
    return r.empty() ? Range(0, 0) : r - o.begin_;
  }

  // the i-th of n nearly equal parts
  Range evenDivide(size_t n, size_t i) const {
    CHECK(valid());
    CHECK_GT(n, (size_t)0);
    CHECK_LT(i, n);
    long double itv = static_cast<long double>(end_ - begin_) / static_cast<long double>(n);
    T b = static_cast<T>(begin_ + itv * i);
    T e = (i + 1 == n) ? end_ : static_cast<T>(begin_ + itv * (i + 1));
    return Range(b, e);
  }

  string toString() const { return "[" + to_string(begin_) + "," + to_string(end_) + ")"; }

 private:
  T begin_;
  T end_;
};

typedef Range<size_t> SizeR;

template <typename T>
std::ostream& operator<<(std::ostream& os, const Range<T>& r) {
  return os << r.toString();
}

}  // namespace PS

namespace std {
template <typename T>
struct hash<PS::Range<T>> {
  size_t operator()(const PS::Range<T>& r) const {
    return std::hash<T>()(r.begin()) * 1000003u ^ std::hash<T>()(r.end());
  }
};
template <typename T>
struct hash<std::pair<int, PS::Range<T>>> {
  size_t operator()(const std::pair<int, PS::Range<T>>& p) const {
    return std::hash<PS::Range<T>>()(p.second) * 31u + (size_t)p.first;
  }
};
}  // namespace std
