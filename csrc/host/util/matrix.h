// Host matrices: CSR/CSC sparse (optionally binary) and dense, with threaded SpMV.
// ref src/util/matrix.h:29-125, src/util/sparse_matrix.h:23-241, src/util/dense_matrix.h:7-104.
// The device path keeps the same data as raw CSR/CSC arrays (csrc/cuda/sparse.cu); these
// classes serve the CPU plumbing path, the oracle and data conversion.
#pragma once
#include <math.h>

#include "proto/messages.h"
#include "util/concurrent.h"
#include "util/file.h"
#include "util/flags.h"
#include "util/shared_array.h"

namespace PS {

DECLARE_int32(num_threads);

template <typename V> class Matrix;
template <typename V> using MatrixPtr = shared_ptr<Matrix<V>>;
template <typename V> using MatrixPtrList = std::vector<MatrixPtr<V>>;

template <typename V>
class Matrix {
 public:
  Matrix() {}
  Matrix(const MatrixInfoView& info, const SArray<V>& value) : info_(info), value_(value) {}
  virtual ~Matrix() {}

  // y = A x ; y = A' x
  virtual void times(const V* x, V* y) const = 0;
  SArray<V> times(const SArray<V>& x) const {
    CHECK_EQ(x.size(), cols());
    SArray<V> y(rows());
    y.setZero();
    times(x.data(), y.data());
    return y;
  }
  SArray<V> operator*(const SArray<V>& x) const { return times(x); }
  virtual MatrixPtr<V> trans() const = 0;
  SArray<V> transTimes(const SArray<V>& x) const { return trans()->times(x); }
  // A .* A (element-wise square), used for the diagonal Hessian
  virtual MatrixPtr<V> dotTimes(const MatrixPtr<V>& B) const = 0;

  virtual MatrixPtr<V> colBlock(SizeR range) const = 0;
  virtual MatrixPtr<V> rowBlock(SizeR range) const = 0;
  virtual MatrixPtr<V> alterStorage() const = 0;
  MatrixPtr<V> toRowMajor() { return rowMajor() ? MatrixPtr<V>() : alterStorage(); }
  MatrixPtr<V> toColMajor() { return colMajor() ? MatrixPtr<V>() : alterStorage(); }

  virtual bool writeToBinFile(string file) const = 0;
  virtual string debugString() const {
    std::ostringstream ss;
    ss << info_.ShortDebugString() << " value " << value_;
    return ss.str();
  }

  MatrixInfo& info() { return info_; }
  const MatrixInfo& info() const { return info_; }
  size_t rows() const { return info_.row().end() - info_.row().begin(); }
  size_t cols() const { return info_.col().end() - info_.col().begin(); }
  size_t nnz() const { return info_.nnz(); }
  bool rowMajor() const { return info_.row_major(); }
  bool colMajor() const { return !info_.row_major(); }
  bool sparse() const { return info_.type() != MatrixInfoView::DENSE; }
  bool binary() const { return info_.type() == MatrixInfoView::SPARSE_BINARY; }
  bool empty() const { return rows() == 0 || cols() == 0; }
  size_t innerSize() const { return rowMajor() ? cols() : rows(); }
  size_t outerSize() const { return rowMajor() ? rows() : cols(); }
  SArray<V>& value() { return value_; }
  const SArray<V>& value() const { return value_; }
  void tranposeInfo() {
    info_.set_row_major(!info_.row_major());
    PbRange r(info_.row()), c(info_.col());
    info_.mutable_row().raw()->copyFrom(*c.raw());
    info_.mutable_col().raw()->copyFrom(*r.raw());
  }

 protected:
  MatrixInfo info_;
  SArray<V> value_;
};

// ---- dense ------------------------------------------------------------------------------------
template <typename V>
class DenseMatrix : public Matrix<V> {
 public:
  using Matrix<V>::info_;
  using Matrix<V>::value_;
  using Matrix<V>::rows;
  using Matrix<V>::cols;
  DenseMatrix() {}
  DenseMatrix(size_t rows, size_t cols, bool row_major = true) { resize(rows, cols, rows * cols, row_major); }
  DenseMatrix(const MatrixInfoView& info, const SArray<V>& value) : Matrix<V>(info, value) {}
  void resize(size_t rows, size_t cols, size_t nnz, bool row_major) {
    info_.set_type(MatrixInfoView::DENSE);
    info_.set_row_major(row_major);
    info_.mutable_row().set_begin(0);
    info_.mutable_row().set_end(rows);
    info_.mutable_col().set_begin(0);
    info_.mutable_col().set_end(cols);
    info_.set_nnz(nnz);
    info_.set_sizeof_value(sizeof(V));
    value_.resize(nnz);
    value_.setZero();
  }
  void times(const V* x, V* y) const override {
    const size_t R = rows(), C = cols();
    if (this->rowMajor()) {
      for (size_t i = 0; i < R; ++i) {
        V s = 0;
        for (size_t j = 0; j < C; ++j) s += value_[i * C + j] * x[j];
        y[i] = s;
      }
    } else {
      for (size_t i = 0; i < R; ++i) y[i] = 0;
      for (size_t j = 0; j < C; ++j)
        for (size_t i = 0; i < R; ++i) y[i] += value_[j * R + i] * x[j];
    }
  }
  MatrixPtr<V> trans() const override {
    auto* m = new DenseMatrix<V>(info_, value_);  // same buffer, flipped interpretation
    m->tranposeInfo();
    return MatrixPtr<V>(m);
  }
  MatrixPtr<V> dotTimes(const MatrixPtr<V>& B) const override {
    SArray<V> v(value_.size());
    for (size_t i = 0; i < v.size(); ++i) v[i] = value_[i] * B->value()[i];
    return MatrixPtr<V>(new DenseMatrix<V>(info_, v));
  }
  MatrixPtr<V> colBlock(SizeR range) const override {
    CHECK(this->colMajor());
    auto info = info_;
    info.mutable_col().set_begin(0);
    info.mutable_col().set_end(range.size());
    info.set_nnz(range.size() * rows());
    return MatrixPtr<V>(new DenseMatrix<V>(info, value_.segment(range * rows())));
  }
  MatrixPtr<V> rowBlock(SizeR range) const override {
    CHECK(this->rowMajor());
    auto info = info_;
    info.mutable_row().set_begin(0);
    info.mutable_row().set_end(range.size());
    info.set_nnz(range.size() * cols());
    return MatrixPtr<V>(new DenseMatrix<V>(info, value_.segment(range * cols())));
  }
  MatrixPtr<V> alterStorage() const override {
    const size_t R = rows(), C = cols();
    SArray<V> v(value_.size());
    if (this->rowMajor()) {
      for (size_t i = 0; i < R; ++i) for (size_t j = 0; j < C; ++j) v[j * R + i] = value_[i * C + j];
    } else {
      for (size_t i = 0; i < R; ++i) for (size_t j = 0; j < C; ++j) v[i * C + j] = value_[j * R + i];
    }
    auto info = info_;
    info.set_row_major(!info_.row_major());
    return MatrixPtr<V>(new DenseMatrix<V>(info, v));
  }
  bool writeToBinFile(string name) const override {
    return writeProtoToASCIIFile(info_, name + ".info") && value_.writeToFile(name + ".value");
  }
};

// ---- sparse (Yale format) ----------------------------------------------------------------------
template <typename I, typename V>
class SparseMatrix : public Matrix<V> {
 public:
  using Matrix<V>::info_;
  using Matrix<V>::value_;
  using Matrix<V>::rows;
  using Matrix<V>::cols;
  SparseMatrix() {}
  SparseMatrix(const MatrixInfoView& info, const SArray<size_t>& offset, const SArray<I>& index,
               const SArray<V>& value)
      : Matrix<V>(info, value), offset_(offset), index_(index) {}

  // y = A x, threaded over the outer dimension for row-major and over row ranges otherwise
  void times(const V* x, V* y) const override {
    const size_t R = rows(), C = cols();
    const bool bin = this->binary();
    const size_t* off = offset_.data();
    const I* idx = index_.data();
    const V* val = value_.data();
    int nt = std::max(1, (int)FLAGS_num_threads);
    if (this->rowMajor()) {
      parallelFor(nt, nt, [&](int t) {
        SizeR rg = SizeR(0, R).evenDivide(nt, t);
        for (size_t i = rg.begin(); i < rg.end(); ++i) {
          V s = 0;
          if (bin) for (size_t o = off[i]; o < off[i + 1]; ++o) s += x[idx[o]];
          else for (size_t o = off[i]; o < off[i + 1]; ++o) s += x[idx[o]] * val[o];
          y[i] = s;
        }
      });
    } else {
      // column-major: every thread owns a slice of the output rows and walks all columns
      memset(y, 0, sizeof(V) * R);
      parallelFor(nt, nt, [&](int t) {
        SizeR rg = SizeR(0, R).evenDivide(nt, t);
        for (size_t j = 0; j < C; ++j) {
          const V xj = x[j];
          if (xj == 0) continue;
          for (size_t o = off[j]; o < off[j + 1]; ++o) {
            const size_t i = idx[o];
            if (i >= rg.begin() && i < rg.end()) y[i] += bin ? xj : xj * val[o];
          }
        }
      });
    }
  }
  // the transpose of CSR is CSC over the same buffers
  MatrixPtr<V> trans() const override {
    auto* m = new SparseMatrix<I, V>(info_, offset_, index_, value_);
    m->tranposeInfo();
    return MatrixPtr<V>(m);
  }
  MatrixPtr<V> dotTimes(const MatrixPtr<V>& B) const override {
    auto C = std::static_pointer_cast<const SparseMatrix<I, V>>(B);
    SArray<V> v;
    if (!this->binary()) {
      v.resize(value_.size());
      for (size_t i = 0; i < v.size(); ++i) v[i] = value_[i] * (C->binary() ? V(1) : C->value()[i]);
    }
    return MatrixPtr<V>(new SparseMatrix<I, V>(info_, offset_, index_, v));
  }
  // zero-copy block of outer entries
  MatrixPtr<V> colBlock(SizeR range) const override {
    CHECK(this->colMajor());
    auto info = info_;
    info.mutable_col().set_begin(0);
    info.mutable_col().set_end(range.size());
    return outerBlock(range, info);
  }
  MatrixPtr<V> rowBlock(SizeR range) const override {
    CHECK(this->rowMajor());
    auto info = info_;
    info.mutable_row().set_begin(0);
    info.mutable_row().set_end(range.size());
    return outerBlock(range, info);
  }
  // CSR <-> CSC by counting sort (stable: inner order stays ascending)
  MatrixPtr<V> alterStorage() const override {
    const size_t outer = this->outerSize(), inner = this->innerSize(), n = index_.size();
    SArray<size_t> off(inner + 1);
    off.setZero();
    for (size_t k = offset_[0]; k < offset_[outer]; ++k) ++off[index_[k] + 1];
    for (size_t i = 0; i < inner; ++i) off[i + 1] += off[i];
    SArray<I> idx(offset_[outer] - offset_[0]);
    SArray<V> val(this->binary() ? 0 : idx.size());
    std::vector<size_t> cur(off.begin(), off.begin() + inner);
    for (size_t o = 0; o < outer; ++o) {
      for (size_t k = offset_[o]; k < offset_[o + 1]; ++k) {
        size_t p = cur[index_[k]]++;
        idx[p] = (I)o;
        if (!this->binary()) val[p] = value_[k];
      }
    }
    (void)n;
    auto info = info_;
    info.set_row_major(!info_.row_major());
    info.set_nnz(idx.size());
    return MatrixPtr<V>(new SparseMatrix<I, V>(info, off, idx, val));
  }
  bool writeToBinFile(string name) const override {
    return writeProtoToASCIIFile(info_, name + ".info") && offset_.writeToFile(name + ".offset") &&
           index_.writeToFile(name + ".index") &&
           (this->binary() || value_.writeToFile(name + ".value"));
  }
  string debugString() const override {
    std::ostringstream ss;
    ss << info_.ShortDebugString() << " offset " << offset_ << " index " << index_ << " value " << value_;
    return ss.str();
  }
  SArray<size_t>& offset() { return offset_; }
  const SArray<size_t>& offset() const { return offset_; }
  SArray<I>& index() { return index_; }
  const SArray<I>& index() const { return index_; }

 private:
  MatrixPtr<V> outerBlock(SizeR range, MatrixInfo& info) const {
    CHECK_LE(range.end(), this->outerSize());
    SArray<size_t> off = offset_.segment(SizeR(range.begin(), range.end() + 1));
    info.set_nnz(off.back() - off.front());
    // offsets stay absolute; index/value are passed whole so off[] indexes them directly
    return MatrixPtr<V>(new SparseMatrix<I, V>(info, off, index_, value_));
  }
  SArray<size_t> offset_;
  SArray<I> index_;
};

template <typename I, typename V>
using SparseMatrixPtr = shared_ptr<SparseMatrix<I, V>>;

// load "<name>.info" + arrays written by writeToBinFile
template <typename V>
MatrixPtr<V> readMatrixFromBin(const string& name) {
  MatrixInfo info;
  if (!readFileToProto(name + ".info", &info)) return MatrixPtr<V>();
  SArray<V> value;
  if (info.type() == MatrixInfoView::DENSE) {
    if (!value.readFromFile(name + ".value")) return MatrixPtr<V>();
    return MatrixPtr<V>(new DenseMatrix<V>(info, value));
  }
  SArray<size_t> offset;
  if (!offset.readFromFile(name + ".offset")) return MatrixPtr<V>();
  if (info.type() != MatrixInfoView::SPARSE_BINARY && !value.readFromFile(name + ".value"))
    return MatrixPtr<V>();
  if (info.sizeof_index() == 8) {
    SArray<uint64> index;
    if (!index.readFromFile(name + ".index")) return MatrixPtr<V>();
    return MatrixPtr<V>(new SparseMatrix<uint64, V>(info, offset, index, value));
  }
  SArray<uint32> index;
  if (!index.readFromFile(name + ".index")) return MatrixPtr<V>();
  return MatrixPtr<V>(new SparseMatrix<uint32, V>(info, offset, index, value));
}

}  // namespace PS
