// Sorting / merging helpers and the Localizer (global 64-bit keys -> dense local columns).
// ref src/util/parallel_sort.h, src/util/parallel_ordered_match.h, src/util/localizer.h,
//     src/util/evaluation.h, src/util/auc.h
// Device counterparts: csrc/cuda/localizer.cu (K10-K12) and csrc/cuda/eval.cu (K14).
#pragma once
#include "util/matrix.h"

namespace PS {

// ---- parallel sort: bisect, sort halves on threads, merge ----------------------------------------
template <typename T, class Fn>
void parallelSort(T* data, size_t len, size_t grainsize, const Fn& cmp) {
  if (len <= grainsize) {
    std::sort(data, data + len, cmp);
    return;
  }
  std::thread thr(parallelSort<T, Fn>, data, len / 2, grainsize, cmp);
  parallelSort(data + len / 2, len - len / 2, grainsize, cmp);
  thr.join();
  std::inplace_merge(data, data + len / 2, data + len, cmp);
}
template <typename T, class Fn>
void parallelSort(SArray<T>* arr, int num_threads, const Fn& cmp) {
  CHECK_GT(num_threads, 0);
  size_t grain = std::max(arr->size() / num_threads + 5, (size_t)1024 * 16);
  parallelSort(arr->data(), arr->size(), grain, cmp);
}

// ---- ordered match: dst[k] op= src[k] for keys present in both sorted key arrays -------------------
template <typename V> struct OpAssign { void operator()(const V* s, V* d) const { *d = *s; } };
template <typename V> struct OpPlus { void operator()(const V* s, V* d) const { *d += *s; } };
template <typename V> struct OpOr { void operator()(const V* s, V* d) const { *d |= *s; } };

template <typename K, typename V, class Op>
void orderedMatchRange(const K* sk, const K* sk_end, const V* sv, const K* dk, const K* dk_end,
                       V* dv, int k, const Op& op, size_t grainsize, size_t* matched) {
  size_t sn = sk_end - sk, dn = dk_end - dk;
  if (sn == 0 || dn == 0) return;
  // narrow both sides to the overlapping key span
  const K* s0 = std::lower_bound(sk, sk_end, *dk);
  sv += (s0 - sk) * k;
  sk = s0;
  const K* d0 = std::lower_bound(dk, dk_end, sk < sk_end ? *sk : *(dk_end - 1) + 1);
  dv += (d0 - dk) * k;
  dk = d0;
  sn = sk_end - sk;
  dn = dk_end - dk;
  if (sn == 0 || dn == 0) return;
  if (dn <= grainsize) {
    size_t m = 0;
    while (sk < sk_end && dk < dk_end) {
      if (*sk < *dk) {
        ++sk; sv += k;
      } else if (*dk < *sk) {
        ++dk; dv += k;
      } else {
        for (int i = 0; i < k; ++i) op(sv + i, dv + i);
        ++sk; sv += k; ++dk; dv += k; ++m;
      }
    }
    if (matched) __sync_fetch_and_add(matched, m);
    return;
  }
  // split the destination in the middle, recurse on a thread
  const K* dm = dk + dn / 2;
  const K* sm = std::lower_bound(sk, sk_end, *dm);
  std::thread thr([&] {
    orderedMatchRange(sk, sm, sv, dk, dm, dv, k, op, grainsize, matched);
  });
  orderedMatchRange(sm, sk_end, sv + (sm - sk) * k, dm, dk_end, dv + (dm - dk) * k, k, op,
                    grainsize, matched);
  thr.join();
}

// dst_val must be sized dst_key.size()*k (it is resized and zeroed when empty); returns #matches
template <typename K, typename V, class Op = OpAssign<V>>
size_t parallelOrderedMatch(const SArray<K>& src_key, const SArray<V>& src_val,
                            const SArray<K>& dst_key, SArray<V>* dst_val, int k = 1,
                            const Op& op = Op(), int num_threads = 0) {
  if (num_threads <= 0) num_threads = std::max(1, (int)FLAGS_num_threads);
  CHECK_EQ(src_key.size() * k, src_val.size());
  if (dst_val->size() != dst_key.size() * k) {
    dst_val->resize(dst_key.size() * k);
    dst_val->setZero();
  }
  size_t grain = std::max(dst_key.size() / num_threads + 5, (size_t)1024 * 1024);
  size_t matched = 0;
  orderedMatchRange(src_key.begin(), src_key.end(), src_val.data(), dst_key.begin(), dst_key.end(),
                    dst_val->data(), k, op, grain, &matched);
  return matched;
}

// union of two sorted key sets with values merged by op (values of keys in only one set are copied)
template <typename K, typename V, class Op = OpPlus<V>>
void parallelUnion(const SArray<K>& k1, const SArray<V>& v1, const SArray<K>& k2,
                   const SArray<V>& v2, SArray<K>* ko, SArray<V>* vo, int k = 1,
                   const Op& op = Op()) {
  *ko = k1.setUnion(k2);
  vo->resize(0);
  vo->resize(ko->size() * k);
  vo->setZero();
  parallelOrderedMatch<K, V, OpAssign<V>>(k1, v1, *ko, vo, k);
  parallelOrderedMatch<K, V, Op>(k2, v2, *ko, vo, k, op);
}

// ---- Localizer -----------------------------------------------------------------------------------
template <typename I, typename V = double>
class Localizer {
 public:
  // unique keys of `idx` (sorted) with saturating counts.  The reference casts the count to C
  // and lets uint8 wrap (localizer.h:94,100); a key seen 256 times would look unseen, so we clamp.
  template <typename C>
  void countUniqIndex(const SArray<I>& idx, SArray<I>* uniq_idx, SArray<C>* idx_frq = nullptr) {
    CHECK(uniq_idx);
    const size_t n = idx.size();
    pair_.resize(n);
    for (size_t i = 0; i < n; ++i) { pair_[i].k = idx[i]; pair_[i].i = (uint32)i; }
    parallelSort(&pair_, std::max(1, (int)FLAGS_num_threads),
                 [](const Pair& a, const Pair& b) { return a.k < b.k; });
    uniq_idx->clear();
    if (idx_frq) idx_frq->clear();
    const uint64 cmax = std::numeric_limits<C>::max();
    size_t run = 0;
    for (size_t i = 0; i < n; ++i) {
      ++run;
      if (i + 1 == n || pair_[i + 1].k != pair_[i].k) {
        const I key = pair_[i].k;  // the pairs are packed: never bind a reference to the field
        uniq_idx->pushBack(key);
        if (idx_frq) idx_frq->pushBack((C)std::min<uint64>(run, cmax));
        run = 0;
      }
    }
  }

  // keep only columns in `idx_dict` (sorted), renumber them 0..|dict|-1, drop empty positions;
  // the result is row-major with uint32 local indices (ref localizer.h:126-191)
  MatrixPtr<V> remapIndex(const MatrixInfoView& info, const SArray<size_t>& offset,
                          const SArray<I>& index, const SArray<V>& value,
                          const SArray<I>& idx_dict) const {
    if (index.empty() || idx_dict.empty()) return MatrixPtr<V>();
    CHECK_EQ(offset.back(), index.size());
    CHECK_EQ(index.size(), pair_.size()) << "call countUniqIndex on the same index array first";
    const bool bin = value.empty();
    if (!bin) CHECK_EQ(value.size(), index.size());
    // merge-join the sorted (key,pos) pairs against the dictionary
    std::vector<uint32> remapped(pair_.size(), 0);
    const I* dict = idx_dict.begin();
    const I* dict_end = idx_dict.end();
    size_t matched = 0;
    for (size_t p = 0; p < pair_.size() && dict < dict_end;) {
      if (pair_[p].k < *dict) {
        ++p;
      } else if (*dict < pair_[p].k) {
        ++dict;
      } else {
        remapped[pair_[p].i] = (uint32)(dict - idx_dict.begin()) + 1;
        ++matched;
        ++p;
      }
    }
    SArray<uint32> new_index(matched);
    SArray<size_t> new_offset(offset.size());
    SArray<V> new_value(bin ? 0 : matched);
    new_offset[0] = 0;
    size_t k = 0;
    for (size_t r = 0; r + 1 < offset.size(); ++r) {
      for (size_t o = offset[r]; o < offset[r + 1]; ++o) {
        if (remapped[o] == 0) continue;
        if (!bin) new_value[k] = value[o];
        new_index[k++] = remapped[o] - 1;
      }
      new_offset[r + 1] = k;
    }
    CHECK_EQ(k, matched);
    MatrixInfo inf(info);
    inf.set_sizeof_index(sizeof(uint32));
    inf.set_nnz(matched);
    inf.set_row_major(true);
    inf.mutable_col().set_begin(0);
    inf.mutable_col().set_end(idx_dict.size());
    inf.mutable_row().set_begin(0);
    inf.mutable_row().set_end(offset.size() - 1);
    return MatrixPtr<V>(new SparseMatrix<uint32, V>(inf, new_offset, new_index, new_value));
  }
  void clear() { pair_.clear(); }
  size_t memSize() const { return pair_.memSize(); }

 private:
#pragma pack(push, 4)
  struct Pair {
    I k;
    uint32 i;
  };
#pragma pack(pop)
  SArray<Pair> pair_;
};

// ---- evaluation ---------------------------------------------------------------------------------
template <typename V>
class Evaluation {
 public:
  // exact AUC by sorting on the prediction (ref evaluation.h:22-48)
  static V auc(const SArray<V>& label, const SArray<V>& predict) {
    const size_t n = label.size();
    CHECK_EQ(n, predict.size());
    std::vector<std::pair<V, V>> buf(n);
    for (size_t i = 0; i < n; ++i) buf[i] = std::make_pair(label[i], predict[i]);
    std::sort(buf.begin(), buf.end(),
              [](const std::pair<V, V>& a, const std::pair<V, V>& b) { return a.second < b.second; });
    V area = 0, cum_tp = 0;
    for (size_t i = 0; i < n; ++i) {
      if (buf[i].first > 0) cum_tp += 1;
      else area += cum_tp;
    }
    if (cum_tp == 0 || cum_tp == (V)n) return 1;
    area /= cum_tp * (n - cum_tp);
    return area < 0.5 ? 1 - area : area;
  }
  static V accuracy(const SArray<V>& label, const SArray<V>& predict, V threshold = 0) {
    const size_t n = label.size();
    CHECK_EQ(n, predict.size());
    V correct = 0;
    for (size_t i = 0; i < n; ++i)
      if ((label[i] > 0 && predict[i] > threshold) || (label[i] < 0 && predict[i] <= threshold))
        correct += 1;
    V acc = n ? correct / n : 0;
    return acc > 0.5 ? acc : 1 - acc;
  }
  static V logloss(const SArray<V>& label, const SArray<V>& predict) {
    V s = 0;
    for (size_t i = 0; i < label.size(); ++i) s += log(1 + exp(-label[i] * predict[i]));
    return s;
  }
};

// mergeable AUC over bucketed predictions (ref auc.h:8-103)
class AUC {
 public:
  void setGoodness(int g) { goodness_ = g; }
  template <typename V>
  void compute(const SArray<V>& y, const SArray<V>& predict, AUCDataView data) {
    CHECK_EQ(y.size(), predict.size());
    CHECK_GT(goodness_, 0);
    clear();
    for (size_t i = 0; i < y.size(); ++i) {
      int64 k = (int64)(predict[i] * goodness_);
      if (y[i] > 0) ++tp_count_[k]; else ++fp_count_[k];
    }
    data.Clear();
    for (auto& it : tp_count_) { data.add_tp_key(it.first); data.add_tp_count(it.second); }
    for (auto& it : fp_count_) { data.add_fp_key(it.first); data.add_fp_count(it.second); }
  }
  void merge(const AUCDataView& data) {
    CHECK_EQ(data.tp_key_size(), data.tp_count_size());
    for (int i = 0; i < data.tp_key_size(); ++i) tp_count_[data.tp_key(i)] += data.tp_count(i);
    for (int i = 0; i < data.fp_key_size(); ++i) fp_count_[data.fp_key(i)] += data.fp_count(i);
  }
  // P(score of a positive > score of a negative) + half of the ties, from the bucket histograms
  double evaluate() {
    if (tp_count_.empty() || fp_count_.empty()) return 0.5;
    double pos_total = 0, neg_total = 0, lower = 0, pos_below = 0;
    for (auto& it : tp_count_) pos_total += it.second;
    auto tp_it = tp_count_.begin();
    for (auto& fp_it : fp_count_) {
      while (tp_it != tp_count_.end() && tp_it->first < fp_it.first) {
        pos_below += tp_it->second;
        ++tp_it;
      }
      double pos_equal = (tp_it != tp_count_.end() && tp_it->first == fp_it.first) ? tp_it->second : 0;
      lower += (double)fp_it.second * (pos_below + 0.5 * pos_equal);
      neg_total += fp_it.second;
    }
    double auc = 1 - lower / (pos_total * neg_total);
    return auc < 0.5 ? 1 - auc : auc;
  }
  double accuracy(double threshold = 0) {
    double total = 0, correct = 0;
    for (auto& it : tp_count_) { total += it.second; if (it.first >= threshold * goodness_) correct += it.second; }
    for (auto& it : fp_count_) { total += it.second; if (it.first < threshold * goodness_) correct += it.second; }
    return total ? correct / total : 0;
  }
  void clear() { tp_count_.clear(); fp_count_.clear(); }

 private:
  int goodness_ = 1000;
  std::map<int64, uint64> tp_count_, fp_count_;
};

}  // namespace PS
