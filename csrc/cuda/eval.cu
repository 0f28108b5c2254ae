// Evaluation and small utility kernels: exact AUC / accuracy / logloss (K14), CountMin sketch
// insert / query (K13), fixed-point value codec (K9), constant / Gaussian fills (K19), bitmap
// popcount (K18), CSR SpMV (K6 stand-alone form, also used by model evaluation K17).
//
// ref: src/util/evaluation.h:22-63, src/util/countmin.h:21-41 + src/util/sketch.h:20-33,
//      src/parameter/frequency_filter.h:26-43, src/filter/fixing_float.h:44-95,
//      src/util/shared_array_inl.h:103-130, src/util/bitmap.h:35-75,
//      src/util/sparse_matrix.h:72-86, src/app/linear_method/model_evaluation.h:40-66
#include "common.cuh"
#include "sketch.cuh"
#include "../../include/psb200.h"
#include <cub/cub.cuh>

namespace psb {

// ---- K14 ----------------------------------------------------------------------------------------
__global__ void k_pos_flags(const float* __restrict__ label_sorted, int64_t n, int64_t* __restrict__ flag) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    flag[i] = label_sorted[i] > 0;
}
// area = sum over negatives of (#positives ranked below); out[0] += area, out[1] += #positives
__global__ void k_auc_area(const float* __restrict__ label_sorted, const int64_t* __restrict__ cum_pos,
                           int64_t n, double* out) {
  double area = 0, pos = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    if (label_sorted[i] > 0) pos += 1;
    else area += (double)cum_pos[i];
  }
  area = warp_sum(area);
  pos = warp_sum(pos);
  if ((threadIdx.x & 31) == 0) {
    if (area != 0) atomicAdd(out, area);
    if (pos != 0) atomicAdd(out + 1, pos);
  }
}
// out[0] += correct, out[1] += logloss
__global__ void k_acc_logloss(const float* __restrict__ label, const float* __restrict__ predict,
                              int64_t n, float threshold, double* out) {
  double correct = 0, ll = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float y = label[i], p = predict[i];
    if ((y > 0 && p > threshold) || (y < 0 && p <= threshold)) correct += 1;
    const double m = -(double)y * p;
    ll += m > 0 ? m + log1p(exp(-m)) : log1p(exp(m));
  }
  correct = warp_sum(correct);
  ll = warp_sum(ll);
  if ((threadIdx.x & 31) == 0) {
    if (correct != 0) atomicAdd(out, correct);
    if (ll != 0) atomicAdd(out + 1, ll);
  }
}

// ---- K13 ----------------------------------------------------------------------------------------
__global__ void k_countmin_insert(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ counts,
                                  int64_t n, uint8_t* table, uint32_t cm_n, int cm_k, uint32_t vmax) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t h = sketch_hash(keys[i]);
    const uint32_t delta = (h >> 17) | (h << 15);
    const uint32_t c = min(counts[i], 255u);
    for (int j = 0; j < cm_k; ++j) {
      sat_add_u8(table, h % cm_n, c, vmax);
      h += delta;
    }
  }
}
__global__ void k_countmin_query(const uint64_t* __restrict__ keys, int64_t n, const uint8_t* __restrict__ table,
                                 uint32_t cm_n, int cm_k, int freq, uint8_t* __restrict__ keep) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t h = sketch_hash(keys[i]);
    const uint32_t delta = (h >> 17) | (h << 15);
    uint32_t est = 255;
    for (int j = 0; j < cm_k; ++j) {
      est = min(est, (uint32_t)table[h % cm_n]);
      h += delta;
    }
    keep[i] = (int)est > freq;
  }
}

// ---- K9 -----------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t splitmix(uint64_t x) {
  x += 0x9e3779b97f4a7c15ULL;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
  return x ^ (x >> 31);
}
template <typename V>
__global__ void k_fixfloat_encode(const V* __restrict__ in, int64_t n, int nbytes, double min_v, double max_v,
                                  uint64_t seed, uint8_t* __restrict__ code) {
  const double ratio = (double)(1ULL << (nbytes * 8)) - 2, bin = max_v - min_v;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const double x = (double)in[i];
    const double proj = x > max_v ? max_v : (x < min_v ? min_v : x);
    const double t = (proj - min_v) / bin * ratio;
    const double u = (double)(splitmix(seed + (uint64_t)i) >> 11) * (1.0 / 9007199254740992.0);
    uint64_t r = (uint64_t)floor(t + u);  // stochastic rounding
    for (int j = 0; j < nbytes; ++j) {
      code[i * nbytes + j] = (uint8_t)(r & 0xff);
      r >>= 8;
    }
  }
}
template <typename V>
__global__ void k_fixfloat_decode(const uint8_t* __restrict__ code, int64_t n, int nbytes, double min_v,
                                  double max_v, V* __restrict__ out) {
  const double ratio = (double)(1ULL << (nbytes * 8)) - 2, bin = max_v - min_v;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t r = 0;
    for (int j = 0; j < nbytes; ++j) r |= (uint64_t)code[i * nbytes + j] << (8 * j);
    out[i] = (V)((double)r / ratio * bin + min_v);
  }
}

// ---- K19 / K18 --------------------------------------------------------------------------------------
__global__ void k_fill_gaussian(double* p, int64_t n, double mean, double std, uint64_t seed) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t a = splitmix(seed + 2 * (uint64_t)i), b = splitmix(seed + 2 * (uint64_t)i + 1);
    const double u1 = ((double)(a >> 11) + 1.0) * (1.0 / 9007199254740993.0);
    const double u2 = (double)(b >> 11) * (1.0 / 9007199254740992.0);
    p[i] = mean + std * sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
  }
}
__global__ void k_fill_const(double* p, int64_t n, double v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}
__global__ void k_bitmap_popcount(const uint32_t* __restrict__ words, int64_t n_words, unsigned long long* out) {
  unsigned long long c = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words;
       i += (int64_t)gridDim.x * blockDim.x)
    c += __popc(words[i]);
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// ---- K6 / K17: y = X w for CSR with 64-bit keys looked up in a sorted dictionary ----------------------
// (model evaluation: the model is a sorted (key, weight) list, data streams by in minibatches)
__global__ void k_spmv_lookup(const int64_t* __restrict__ row_ptr, const uint64_t* __restrict__ keys,
                              const float* __restrict__ val, int64_t n_rows, const uint64_t* __restrict__ dict,
                              const double* __restrict__ w, int64_t n_dict, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = gw; r < n_rows; r += nw) {
    double s = 0;
    for (int64_t e = row_ptr[r] + lane; e < row_ptr[r + 1]; e += 32) {
      const uint64_t k = keys[e];
      int64_t lo = 0, hi = n_dict;
      while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (dict[mid] < k) lo = mid + 1; else hi = mid;
      }
      if (lo < n_dict && dict[lo] == k) s += val ? w[lo] * (double)val[e] : w[lo];
    }
    s = warp_sum(s);
    if (lane == 0) out[r] = (float)s;
  }
}

}  // namespace psb

using namespace psb;

static inline int grid_for(int64_t n) {
  int64_t g = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// K14: exact AUC.  work: float[2n] (sorted predictions, sorted labels) + int64[n] (flags -> cumulative
// positives); out_dev: double[2] zeroed by the caller, receives {area, #positives}; temp per CUB convention.
PSB_API int psb_auc_exact(const float* label, const float* predict, int64_t n, float* work_f,
                          int64_t* work_i, double* out_dev, void* temp, int64_t* temp_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  size_t b1 = 0, b2 = 0;
  PSB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, b1, predict, work_f, label, work_f + n, n, 0, 32, st));
  PSB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, b2, work_i, work_i, n, st));
  const size_t need = b1 > b2 ? b1 : b2;
  if (!temp) {
    *temp_bytes = (int64_t)need;
    return 0;
  }
  PSB_REQUIRE((size_t)*temp_bytes >= need, "psb_auc_exact: temp too small");
  if (n == 0) return 0;
  size_t bytes = (size_t)*temp_bytes;
  PSB_CUDA(cub::DeviceRadixSort::SortPairs(temp, bytes, predict, work_f, label, work_f + n, n, 0, 32, st));
  k_pos_flags<<<grid_for(n), 256, 0, st>>>(work_f + n, n, work_i);
  PSB_LAUNCH_CHECK();
  bytes = (size_t)*temp_bytes;
  PSB_CUDA(cub::DeviceScan::ExclusiveSum(temp, bytes, work_i, work_i, n, st));
  k_auc_area<<<grid_for(n), 256, 0, st>>>(work_f + n, work_i, n, out_dev);
  PSB_LAUNCH_CHECK();
  return 0;
}
// out_dev: double[2] zeroed by the caller, receives {#correct, total logloss}
PSB_API int psb_accuracy_logloss(const float* label, const float* predict, int64_t n, float threshold,
                                 double* out_dev, void* stream) {
  if (n == 0) return 0;
  k_acc_logloss<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(label, predict, n, threshold, out_dev);
  PSB_LAUNCH_CHECK();
  return 0;
}

// K13: CountMin with 8-bit saturating counters (table: cm_n bytes, zeroed by the caller, cm_n % 4 == 0)
PSB_API int psb_countmin_insert(const uint64_t* keys, const uint32_t* counts, int64_t n, uint8_t* table,
                                int32_t cm_n, int32_t cm_k, void* stream) {
  PSB_REQUIRE(cm_n >= 64 && cm_n % 4 == 0 && cm_k >= 1 && cm_k <= 30, "psb_countmin_insert: bad sketch size");
  if (n == 0) return 0;
  k_countmin_insert<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(keys, counts, n, table, (uint32_t)cm_n, cm_k, 254u);
  PSB_LAUNCH_CHECK();
  return 0;
}
PSB_API int psb_countmin_query(const uint64_t* keys, int64_t n, const uint8_t* table, int32_t cm_n, int32_t cm_k,
                               int32_t freq, uint8_t* keep_out, void* stream) {
  if (n == 0) return 0;
  k_countmin_query<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(keys, n, table, (uint32_t)cm_n, cm_k, freq, keep_out);
  PSB_LAUNCH_CHECK();
  return 0;
}

// K9: the FIXING_FLOAT wire codec (nbytes 1..7, little endian, ratio 2^(8 nbytes) - 2)
PSB_API int psb_fixfloat_encode_f32(const float* in, int64_t n, int32_t nbytes, double min_v, double max_v,
                                    uint64_t seed, uint8_t* code, void* stream) {
  PSB_REQUIRE(nbytes >= 1 && nbytes <= 7 && max_v > min_v, "psb_fixfloat_encode: bad config");
  if (n == 0) return 0;
  k_fixfloat_encode<float><<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(in, n, nbytes, min_v, max_v, seed, code);
  PSB_LAUNCH_CHECK();
  return 0;
}
PSB_API int psb_fixfloat_decode_f32(const uint8_t* code, int64_t n, int32_t nbytes, double min_v, double max_v,
                                    float* out, void* stream) {
  PSB_REQUIRE(nbytes >= 1 && nbytes <= 7 && max_v > min_v, "psb_fixfloat_decode: bad config");
  if (n == 0) return 0;
  k_fixfloat_decode<float><<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(code, n, nbytes, min_v, max_v, out);
  PSB_LAUNCH_CHECK();
  return 0;
}
PSB_API int psb_fixfloat_encode_f64(const double* in, int64_t n, int32_t nbytes, double min_v, double max_v,
                                    uint64_t seed, uint8_t* code, void* stream) {
  PSB_REQUIRE(nbytes >= 1 && nbytes <= 7 && max_v > min_v, "psb_fixfloat_encode: bad config");
  if (n == 0) return 0;
  k_fixfloat_encode<double><<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(in, n, nbytes, min_v, max_v, seed, code);
  PSB_LAUNCH_CHECK();
  return 0;
}
PSB_API int psb_fixfloat_decode_f64(const uint8_t* code, int64_t n, int32_t nbytes, double min_v, double max_v,
                                    double* out, void* stream) {
  PSB_REQUIRE(nbytes >= 1 && nbytes <= 7 && max_v > min_v, "psb_fixfloat_decode: bad config");
  if (n == 0) return 0;
  k_fixfloat_decode<double><<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(code, n, nbytes, min_v, max_v, out);
  PSB_LAUNCH_CHECK();
  return 0;
}

// K19: ParameterInitConfig ZERO / CONSTANT / GAUSSIAN on a double array
PSB_API int psb_fill_f64(double* p, int64_t n, int32_t type, double constant, double mean, double std,
                         uint64_t seed, void* stream) {
  if (n == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (type == 3) k_fill_gaussian<<<grid_for(n), 256, 0, st>>>(p, n, mean, std, seed);
  else k_fill_const<<<grid_for(n), 256, 0, st>>>(p, n, type == 2 ? constant : 0.0);
  PSB_LAUNCH_CHECK();
  return 0;
}
// K18: popcount of a bitmap; *out_dev zeroed by the caller
PSB_API int psb_bitmap_popcount(const uint32_t* words, int64_t n_words, uint64_t* out_dev, void* stream) {
  if (n_words == 0) return 0;
  k_bitmap_popcount<<<grid_for(n_words), 256, 0, (cudaStream_t)stream>>>(words, n_words, (unsigned long long*)out_dev);
  PSB_LAUNCH_CHECK();
  return 0;
}
// K6 / K17: margins of CSR rows with 64-bit keys against a sorted (key, weight) model
PSB_API int psb_spmv_lookup(const int64_t* row_ptr, const uint64_t* keys, const float* val, int64_t n_rows,
                            const uint64_t* dict, const double* w, int64_t n_dict, float* out, void* stream) {
  if (n_rows == 0) return 0;
  k_spmv_lookup<<<grid_for(n_rows * 32), 256, 0, (cudaStream_t)stream>>>(row_ptr, keys, val, n_rows, dict, w,
                                                                         n_dict, out);
  PSB_LAUNCH_CHECK();
  return 0;
}
