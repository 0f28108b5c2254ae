// Data-pipeline kernels: unique/count (K10), dictionary remap + row compaction (K11),
// CSR <-> CSC (K12), shard split (K15), key-set signature (K16), work segments for Darlin.
//
// ref: src/util/localizer.h:69-191 (countUniqIndex / remapIndex), src/util/parallel_sort.h,
//      src/util/sparse_matrix.h:185-241 (alterStorage), src/system/message.h:120-159
//      (sliceKeyOrderedMsg), src/filter/key_caching.h:18,74 (signature).
// The reference sorts (key,pos) pairs on CPU threads; here sorting / scanning / run-length
// encoding use CUB device primitives and the joins are binary-search kernels.
//
// Entry points that need temporary storage follow the CUB convention: call with temp==NULL
// to get the size in *temp_bytes, then call again with a buffer of at least that size.
#include "common.cuh"
#include "../../include/psb200.h"
#include <cub/cub.cuh>

namespace psb {

__global__ void k_lookup_sorted(const uint64_t* __restrict__ dict, int64_t nd,
                                const uint64_t* __restrict__ keys, int64_t n,
                                uint32_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t k = keys[i];
    int64_t lo = 0, hi = nd;
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if (dict[mid] < k) lo = mid + 1; else hi = mid;
    }
    out[i] = (lo < nd && dict[lo] == k) ? (uint32_t)(lo + 1) : 0u;
  }
}

__global__ void k_lower_bound_u64(const uint64_t* __restrict__ sorted, int64_t n,
                                  const uint64_t* __restrict__ probes, int64_t np,
                                  int64_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < np;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t k = probes[i];
    int64_t lo = 0, hi = n;
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if (sorted[mid] < k) lo = mid + 1; else hi = mid;
    }
    out[i] = lo;
  }
}

// lower bound of every column id 0..n_cols in a sorted uint32 column array -> col_ptr
__global__ void k_col_ptr_from_sorted(const uint32_t* __restrict__ sorted_col, int64_t nnz,
                                      int64_t n_cols, int64_t* __restrict__ col_ptr) {
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c <= n_cols;
       c += (int64_t)gridDim.x * blockDim.x) {
    int64_t lo = 0, hi = nnz;
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if ((int64_t)sorted_col[mid] < c) lo = mid + 1; else hi = mid;
    }
    col_ptr[c] = lo;
  }
}

__global__ void k_expand_rows(const int64_t* __restrict__ row_ptr, int64_t n_rows,
                              uint32_t* __restrict__ row_of) {
  // one warp per row: rows are short (tens to hundreds of entries)
  const int lane = threadIdx.x & 31;
  const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = w; r < n_rows; r += nw)
    for (int64_t e = row_ptr[r] + lane; e < row_ptr[r + 1]; e += 32) row_of[e] = (uint32_t)r;
}

__global__ void k_iota_u32(uint32_t* p, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = (uint32_t)i;
}

// gather through the sort permutation; sval = y[row] * val (label folded into the value)
__global__ void k_gather_csc(const uint32_t* __restrict__ perm, const uint32_t* __restrict__ row_of,
                             const float* __restrict__ val, const float* __restrict__ y,
                             int64_t nnz, uint32_t* __restrict__ row_idx,
                             float* __restrict__ val_out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t p = perm[i];
    const uint32_t r = row_of[p];
    row_idx[i] = r;
    if (val_out) val_out[i] = y ? y[r] * val[p] : val[p];
  }
}

__global__ void k_flag_nonzero(const uint32_t* __restrict__ idx1, int64_t n,
                               int64_t* __restrict__ flag) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    flag[i] = idx1[i] != 0;
}

__global__ void k_compact_scatter(const uint32_t* __restrict__ idx1, const float* __restrict__ val,
                                  const int64_t* __restrict__ pos, int64_t n,
                                  uint32_t* __restrict__ col_out, float* __restrict__ val_out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t c = idx1[i];
    if (c) {
      col_out[pos[i]] = c - 1;
      if (val_out) val_out[pos[i]] = val[i];
    }
  }
}

__global__ void k_remap_row_ptr(const int64_t* __restrict__ row_ptr, int64_t n_rows,
                                const int64_t* __restrict__ pos, int64_t nnz_old,
                                int64_t nnz_new, int64_t* __restrict__ out) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r <= n_rows;
       r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = row_ptr[r];
    out[r] = e >= nnz_old ? nnz_new : pos[e];
  }
}

__global__ void k_seg_count(const int64_t* __restrict__ col_ptr, int64_t n_cols, int seg_len,
                            int64_t* __restrict__ cnt) {
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < n_cols;
       c += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = col_ptr[c + 1] - col_ptr[c];
    cnt[c] = n == 0 ? 1 : (n + seg_len - 1) / seg_len;  // empty columns keep one (empty) segment
  }
}

__global__ void k_seg_fill(const int64_t* __restrict__ col_ptr, const int64_t* __restrict__ col_seg,
                           int64_t n_cols, int seg_len, int32_t* __restrict__ seg_col,
                           int64_t* __restrict__ seg_beg, int4* __restrict__ seg_meta) {
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < n_cols;
       c += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s0 = col_seg[c], s1 = col_seg[c + 1];
    for (int64_t s = s0; s < s1; ++s) {
      const int64_t beg = col_ptr[c] + (s - s0) * seg_len;
      seg_col[s] = (int32_t)c;
      seg_beg[s] = beg;
      if (seg_meta) {  // {col, len | first<<16 | single<<17, beg lo, beg hi}
        int64_t len = col_ptr[c + 1] - beg;
        len = len < 0 ? 0 : (len > seg_len ? seg_len : len);
        const int flags = (int)len | ((s == s0) << 16) | ((s1 - s0 == 1) << 17);
        seg_meta[s] = make_int4((int)c, flags, (int)(uint32_t)(beg & 0xffffffffLL), (int)(beg >> 32));
      }
    }
  }
}

__device__ __forceinline__ uint64_t mix64(uint64_t x) {  // splitmix64 finaliser
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
  x ^= x >> 27; x *= 0x94d049bb133111ebULL;
  x ^= x >> 31;
  return x;
}

// order-sensitive, reduction-order-insensitive signature: sum_i mix(key_i + phi*i) mod 2^64
__global__ void k_key_signature(const uint64_t* __restrict__ keys, int64_t n,
                                unsigned long long* out) {
  unsigned long long acc = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    acc += mix64(keys[i] + 0x9e3779b97f4a7c15ULL * (uint64_t)(i + 1));
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}

}  // namespace psb

using namespace psb;

static inline int grid_for(int64_t n, int threads = 256) {
  int64_t g = (n + threads - 1) / threads;
  int64_t cap = (int64_t)sm_count() * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// ---- sort / unique / scan (CUB) -----------------------------------------------------------
PSB_API int psb_sort_pairs_u64_u32(const uint64_t* keys_in, uint64_t* keys_out,
                                   const uint32_t* vals_in, uint32_t* vals_out, int64_t n,
                                   int begin_bit, int end_bit, void* temp, int64_t* temp_bytes,
                                   void* stream) {
  size_t bytes = temp ? (size_t)*temp_bytes : 0;
  PSB_CUDA(cub::DeviceRadixSort::SortPairs(temp, bytes, keys_in, keys_out, vals_in, vals_out, n,
                                           begin_bit, end_bit, (cudaStream_t)stream));
  if (!temp) *temp_bytes = (int64_t)bytes;
  return 0;
}
PSB_API int psb_sort_pairs_u32_u32(const uint32_t* keys_in, uint32_t* keys_out,
                                   const uint32_t* vals_in, uint32_t* vals_out, int64_t n,
                                   int begin_bit, int end_bit, void* temp, int64_t* temp_bytes,
                                   void* stream) {
  size_t bytes = temp ? (size_t)*temp_bytes : 0;
  PSB_CUDA(cub::DeviceRadixSort::SortPairs(temp, bytes, keys_in, keys_out, vals_in, vals_out, n,
                                           begin_bit, end_bit, (cudaStream_t)stream));
  if (!temp) *temp_bytes = (int64_t)bytes;
  return 0;
}
PSB_API int psb_sort_keys_u64(const uint64_t* keys_in, uint64_t* keys_out, int64_t n,
                              int begin_bit, int end_bit, void* temp, int64_t* temp_bytes,
                              void* stream) {
  size_t bytes = temp ? (size_t)*temp_bytes : 0;
  PSB_CUDA(cub::DeviceRadixSort::SortKeys(temp, bytes, keys_in, keys_out, n, begin_bit, end_bit,
                                          (cudaStream_t)stream));
  if (!temp) *temp_bytes = (int64_t)bytes;
  return 0;
}
// run-length encode a sorted key array: uniq[], counts[] (uint32, NOT wrapped to uint8 — the
// reference casts to uint8, localizer.h:94,100; callers saturate instead), *num_runs (device)
PSB_API int psb_rle_u64(const uint64_t* sorted, int64_t n, uint64_t* uniq, uint32_t* counts,
                        int64_t* num_runs_dev, void* temp, int64_t* temp_bytes, void* stream) {
  PSB_REQUIRE(n < (1LL << 31), "psb_rle_u64: n=%lld exceeds 2^31-1", (long long)n);
  size_t bytes = temp ? (size_t)*temp_bytes : 0;
  PSB_CUDA(cub::DeviceRunLengthEncode::Encode(temp, bytes, sorted, uniq, counts, num_runs_dev,
                                              (int)n, (cudaStream_t)stream));
  if (!temp) *temp_bytes = (int64_t)bytes;
  return 0;
}
PSB_API int psb_exclusive_sum_i64(const int64_t* in, int64_t* out, int64_t n, void* temp,
                                  int64_t* temp_bytes, void* stream) {
  size_t bytes = temp ? (size_t)*temp_bytes : 0;
  PSB_CUDA(cub::DeviceScan::ExclusiveSum(temp, bytes, in, out, n, (cudaStream_t)stream));
  if (!temp) *temp_bytes = (int64_t)bytes;
  return 0;
}

// ---- joins ------------------------------------------------------------------------------
// out[i] = 1 + position of keys[i] in the sorted dictionary, 0 if absent (ref localizer.h:146-165)
PSB_API int psb_lookup_sorted_u64(const uint64_t* dict, int64_t n_dict, const uint64_t* keys,
                                  int64_t n, uint32_t* out_idx_plus1, void* stream) {
  PSB_REQUIRE(n_dict < 0xffffffffLL, "psb_lookup_sorted_u64: dictionary too large for uint32");
  if (n == 0) return 0;
  k_lookup_sorted<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(dict, n_dict, keys, n, out_idx_plus1);
  PSB_LAUNCH_CHECK();
  return 0;
}
// K15: lower bounds of `probes` (e.g. server range boundaries, block boundaries) in sorted keys
PSB_API int psb_lower_bound_u64(const uint64_t* sorted, int64_t n, const uint64_t* probes,
                                int64_t n_probes, int64_t* out, void* stream) {
  if (n_probes == 0) return 0;
  k_lower_bound_u64<<<grid_for(n_probes), 256, 0, (cudaStream_t)stream>>>(sorted, n, probes, n_probes, out);
  PSB_LAUNCH_CHECK();
  return 0;
}

// ---- K11: drop unmapped entries, rebuild offsets -------------------------------------------
// idx1: 1+local column or 0 (drop).  scan_tmp: int64[nnz] scratch (flags, then positions).
PSB_API int psb_compact_csr(const int64_t* row_ptr, int64_t n_rows, const uint32_t* idx1,
                            const float* val, int64_t nnz, int64_t* scan_tmp,
                            int64_t* row_ptr_out, uint32_t* col_out, float* val_out,
                            void* temp, int64_t* temp_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (!temp) return psb_exclusive_sum_i64(scan_tmp, scan_tmp, nnz, nullptr, temp_bytes, stream);
  if (nnz > 0) {
    k_flag_nonzero<<<grid_for(nnz), 256, 0, st>>>(idx1, nnz, scan_tmp);
    PSB_LAUNCH_CHECK();
    if (int rc = psb_exclusive_sum_i64(scan_tmp, scan_tmp, nnz, temp, temp_bytes, stream)) return rc;
    k_compact_scatter<<<grid_for(nnz), 256, 0, st>>>(idx1, val, scan_tmp, nnz, col_out, val_out);
    PSB_LAUNCH_CHECK();
  }
  return 0;
}
// second half: new row_ptr once the caller knows nnz_new (= pos[nnz-1] + flag[nnz-1])
PSB_API int psb_compact_row_ptr(const int64_t* row_ptr, int64_t n_rows, const int64_t* pos,
                                int64_t nnz_old, int64_t nnz_new, int64_t* row_ptr_out,
                                void* stream) {
  k_remap_row_ptr<<<grid_for(n_rows + 1), 256, 0, (cudaStream_t)stream>>>(row_ptr, n_rows, pos, nnz_old,
                                                                         nnz_new, row_ptr_out);
  PSB_LAUNCH_CHECK();
  return 0;
}

// ---- K12: CSR -> CSC (stable: rows ascending inside every column) --------------------------
// work: uint32[3*nnz] scratch (row_of, perm_in/out); sorted_col: uint32[nnz] scratch.
// y != NULL folds the label into the values (sval = y_i * x_ij).
PSB_API int psb_csr_to_csc(const int64_t* row_ptr, const uint32_t* col, const float* val,
                           const float* y, int64_t n_rows, int64_t n_cols, int64_t nnz,
                           int64_t* col_ptr, uint32_t* row_idx, float* val_out, uint32_t* work,
                           uint32_t* sorted_col, void* temp, int64_t* temp_bytes, void* stream) {
  PSB_REQUIRE(nnz < 0xffffffffLL, "psb_csr_to_csc: nnz=%lld needs 64-bit permutation", (long long)nnz);
  cudaStream_t st = (cudaStream_t)stream;
  int end_bit = 1;
  while (end_bit < 32 && (1LL << end_bit) < n_cols) ++end_bit;
  uint32_t* row_of = work;
  uint32_t* perm_in = work + nnz;
  uint32_t* perm_out = work + 2 * nnz;
  if (!temp)
    return psb_sort_pairs_u32_u32(col, sorted_col, perm_in, perm_out, nnz, 0, end_bit, nullptr,
                                  temp_bytes, stream);
  if (nnz > 0) {
    k_expand_rows<<<grid_for(n_rows * 32), 256, 0, st>>>(row_ptr, n_rows, row_of);
    PSB_LAUNCH_CHECK();
    k_iota_u32<<<grid_for(nnz), 256, 0, st>>>(perm_in, nnz);
    PSB_LAUNCH_CHECK();
    if (int rc = psb_sort_pairs_u32_u32(col, sorted_col, perm_in, perm_out, nnz, 0, end_bit, temp,
                                        temp_bytes, stream))
      return rc;
    k_gather_csc<<<grid_for(nnz), 256, 0, st>>>(perm_out, row_of, val, y, nnz, row_idx, val_out);
    PSB_LAUNCH_CHECK();
  }
  k_col_ptr_from_sorted<<<grid_for(n_cols + 1), 256, 0, st>>>(sorted_col, nnz, n_cols, col_ptr);
  PSB_LAUNCH_CHECK();
  return 0;
}

// ---- Darlin work segments --------------------------------------------------------------------
// step 1: cnt[c] = segments of column c (write into col_seg[0..n_cols), then exclusive-scan
// n_cols+1 entries in place); step 2: fill seg_col / seg_beg.
PSB_API int psb_segments_count(const int64_t* col_ptr, int64_t n_cols, int32_t seg_len,
                               int64_t* col_seg, void* temp, int64_t* temp_bytes, void* stream) {
  PSB_REQUIRE(seg_len >= 32, "psb_segments_count: seg_len must be >= 32");
  if (!temp) return psb_exclusive_sum_i64(col_seg, col_seg, n_cols + 1, nullptr, temp_bytes, stream);
  PSB_CUDA(cudaMemsetAsync(col_seg + n_cols, 0, sizeof(int64_t), (cudaStream_t)stream));
  if (n_cols > 0) {
    k_seg_count<<<grid_for(n_cols), 256, 0, (cudaStream_t)stream>>>(col_ptr, n_cols, seg_len, col_seg);
    PSB_LAUNCH_CHECK();
  }
  return psb_exclusive_sum_i64(col_seg, col_seg, n_cols + 1, temp, temp_bytes, stream);
}
PSB_API int psb_segments_fill(const int64_t* col_ptr, const int64_t* col_seg, int64_t n_cols,
                              int32_t seg_len, int32_t* seg_col, int64_t* seg_beg,
                              void* seg_meta, void* stream) {
  PSB_REQUIRE(seg_len <= 32768, "psb_segments_fill: seg_len must be <= 32768");
  if (n_cols == 0) return 0;
  k_seg_fill<<<grid_for(n_cols), 256, 0, (cudaStream_t)stream>>>(col_ptr, col_seg, n_cols, seg_len,
                                                                 seg_col, seg_beg, (int4*)seg_meta);
  PSB_LAUNCH_CHECK();
  return 0;
}

// ---- K16 --------------------------------------------------------------------------------------
// 64-bit signature of a key array; *out_dev must be zeroed by the caller (memset on `stream`).
PSB_API int psb_key_signature(const uint64_t* keys, int64_t n, uint64_t* out_dev, void* stream) {
  if (n == 0) return 0;
  k_key_signature<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(
      keys, n, reinterpret_cast<unsigned long long*>(out_dev));
  PSB_LAUNCH_CHECK();
  return 0;
}
