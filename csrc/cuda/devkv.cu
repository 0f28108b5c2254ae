// Device side of the public key-value containers (KVVector / KVBufferedVector / KVMap with
// their storage in HBM): typed apply-with-operator, gather, key matching and the hash map.
//
// The host runtime (csrc/host/parameter/kv.h, device backend) keeps keys sorted on the host for
// slicing and holds values here.  A pushed value frame arrives as a pointer into the SENDER's
// staging arena (CUDA IPC mapping): psb_dkv_apply reads it over NVLink while it applies the
// operator, so the payload crosses the link exactly once and never touches host memory.
// ref: KVVector::setValue / getValue (src/parameter/kv_vector.h:46-93), KVMap (kv_map.h:4-45),
// compAssOp (shared_parameter.h:173-193), parallelOrderedMatch (parallel_ordered_match.h:5-44)
#include "common.cuh"
#include "../../include/psb200.h"

namespace psb {

constexpr unsigned long long kDmapEmpty = ~0ull;

__device__ __forceinline__ uint64_t dmap_hash(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

template <typename T>
__device__ __forceinline__ T apply_op(T cur, T in, int op) {
  switch (op) {
    case PSB_OP_ASSIGN: return in;
    case PSB_OP_PLUS: return cur + in;
    case PSB_OP_MINUS: return cur - in;
    case PSB_OP_TIMES: return cur * in;
    case PSB_OP_DIVIDE: return cur / in;
    default: return cur;
  }
}
template <typename T>
__device__ __forceinline__ T apply_op_int(T cur, T in, int op) {
  switch (op) {
    case PSB_OP_AND: return cur & in;
    case PSB_OP_OR: return cur | in;
    case PSB_OP_XOR: return cur ^ in;
    default: return apply_op(cur, in, op);
  }
}
__device__ __forceinline__ float apply_any(float c, float i, int op) { return apply_op(c, i, op); }
__device__ __forceinline__ double apply_any(double c, double i, int op) { return apply_op(c, i, op); }
__device__ __forceinline__ int32_t apply_any(int32_t c, int32_t i, int op) { return apply_op_int(c, i, op); }
__device__ __forceinline__ int64_t apply_any(int64_t c, int64_t i, int op) { return apply_op_int(c, i, op); }
__device__ __forceinline__ uint8_t apply_any(uint8_t c, uint8_t i, int op) { return apply_op_int(c, i, op); }

// dst[pos[i]*k + c] = dst[...] (op) src[i*k + c];  pos == NULL: contiguous; pos[i] < 0: skipped
template <typename T>
__global__ void k_dkv_apply(T* __restrict__ dst, const int64_t* __restrict__ pos, const T* __restrict__ src,
                            int64_t n, int k, int op) {
  const int64_t total = n * k;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = t / k;
    const int c = (int)(t - i * k);
    const int64_t p = pos ? pos[i] : i;
    if (p < 0) continue;
    T* d = dst + p * k + c;
    *d = apply_any(*d, src[t], op);  // src may be a peer mapping: the load crosses NVLink here
  }
}

// dst[i*kb .. +kb) = src[pos[i]*kb .. +kb) as bytes (dst may be a peer mapping); pos[i] < 0 -> zeros
__global__ void k_dkv_gather(unsigned char* __restrict__ dst, const unsigned char* __restrict__ src,
                             const int64_t* __restrict__ pos, int64_t n, int kb) {
  const int words = kb / 4;  // kb is a multiple of 4 (checked by the entry point) or 1
  if (kb % 4 == 0) {
    const int64_t total = n * words;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
      const int64_t i = t / words;
      const int c = (int)(t - i * words);
      const int64_t p = pos ? pos[i] : i;
      reinterpret_cast<uint32_t*>(dst)[t] = p < 0 ? 0u : reinterpret_cast<const uint32_t*>(src)[p * words + c];
    }
  } else {
    const int64_t total = n * kb;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
      const int64_t i = t / kb;
      const int64_t p = pos ? pos[i] : i;
      dst[t] = p < 0 ? 0 : src[p * kb + (t - i * kb)];
    }
  }
}

// pos[i] = position of keys[i] in the sorted, duplicate-free my_keys, or -1
__global__ void k_dkv_positions(const uint64_t* __restrict__ my_keys, int64_t n_my,
                                const uint64_t* __restrict__ keys, int64_t n, int64_t* __restrict__ pos) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t key = keys[i];
    int64_t lo = 0, hi = n_my;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (my_keys[mid] < key) lo = mid + 1; else hi = mid;
    }
    pos[i] = (lo < n_my && my_keys[lo] == key) ? lo : -1;
  }
}

// open addressing, linear probing; insert != 0 claims an empty slot for an unseen key
__global__ void k_dmap_slots(unsigned long long* __restrict__ table, int64_t cap,
                             const uint64_t* __restrict__ keys, int64_t n, int insert,
                             int64_t* __restrict__ slot, int* __restrict__ overflow) {
  const uint64_t mask = (uint64_t)cap - 1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long key = keys[i];
    uint64_t h = dmap_hash(key) & mask;
    int64_t found = -1;
    for (uint64_t probe = 0; probe <= mask; ++probe) {
      unsigned long long k = *reinterpret_cast<volatile unsigned long long*>(table + h);
      if (k == kDmapEmpty) {
        if (!insert) break;
        k = atomicCAS(table + h, kDmapEmpty, key);
        if (k == kDmapEmpty) k = key;
      }
      if (k == key) { found = (int64_t)h; break; }
      h = (h + 1) & mask;
    }
    if (found < 0 && insert) *overflow = 1;
    slot[i] = found;
  }
}

__global__ void k_dmap_export(const unsigned long long* __restrict__ table, int64_t cap,
                              unsigned long long* __restrict__ out_keys, int64_t* __restrict__ out_slot,
                              unsigned long long* __restrict__ count) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cap;
       i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long k = table[i];
    if (k == kDmapEmpty) continue;
    const unsigned long long at = atomicAdd(count, 1ull);
    out_keys[at] = k;
    out_slot[at] = i;
  }
}

__global__ void k_fill_words(unsigned long long* p, int64_t n, unsigned long long v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

static inline int grid_of(int64_t n) {
  int64_t g = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace psb

using namespace psb;

PSB_API int psb_device_count(int32_t* out) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) { cudaGetLastError(); n = 0; }
  *out = n;
  return 0;
}
PSB_API int psb_set_device(int32_t dev) {
  PSB_CUDA(cudaSetDevice(dev));
  return 0;
}
PSB_API int psb_stream_create(void** out) {
  cudaStream_t s;
  PSB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *out = (void*)s;
  return 0;
}
PSB_API int psb_stream_destroy(void* s) {
  if (s) PSB_CUDA(cudaStreamDestroy((cudaStream_t)s));
  return 0;
}

PSB_API int psb_dkv_apply(void* dst, const int64_t* pos, const void* src, int64_t n, int32_t k,
                          int32_t dtype, int32_t op, void* stream) {
  PSB_REQUIRE(dst && src && n >= 0 && k >= 1, "psb_dkv_apply: bad argument");
  PSB_REQUIRE(op >= PSB_OP_ASSIGN && op <= PSB_OP_XOR, "psb_dkv_apply: operator %d", op);
  if (n == 0) return 0;
  const bool fl = dtype == PSB_DT_F32 || dtype == PSB_DT_F64;
  PSB_REQUIRE(!(fl && op >= PSB_OP_AND), "psb_dkv_apply: AND / OR / XOR need an integer value type");
  cudaStream_t s = (cudaStream_t)stream;
  const int g = grid_of(n * k);
  switch (dtype) {
    case PSB_DT_F32: k_dkv_apply<float><<<g, 256, 0, s>>>((float*)dst, pos, (const float*)src, n, k, op); break;
    case PSB_DT_F64: k_dkv_apply<double><<<g, 256, 0, s>>>((double*)dst, pos, (const double*)src, n, k, op); break;
    case PSB_DT_I32: k_dkv_apply<int32_t><<<g, 256, 0, s>>>((int32_t*)dst, pos, (const int32_t*)src, n, k, op); break;
    case PSB_DT_I64: k_dkv_apply<int64_t><<<g, 256, 0, s>>>((int64_t*)dst, pos, (const int64_t*)src, n, k, op); break;
    case PSB_DT_U8: k_dkv_apply<uint8_t><<<g, 256, 0, s>>>((uint8_t*)dst, pos, (const uint8_t*)src, n, k, op); break;
    default: PSB_REQUIRE(false, "psb_dkv_apply: value type %d", dtype);
  }
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int psb_dkv_gather(void* dst, const void* src, const int64_t* pos, int64_t n, int32_t entry_bytes,
                           void* stream) {
  PSB_REQUIRE(dst && src && n >= 0 && entry_bytes >= 1, "psb_dkv_gather: bad argument");
  if (n == 0) return 0;
  k_dkv_gather<<<grid_of(n * (entry_bytes % 4 == 0 ? entry_bytes / 4 : entry_bytes)), 256, 0, (cudaStream_t)stream>>>(
      (unsigned char*)dst, (const unsigned char*)src, pos, n, entry_bytes);
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int psb_dkv_positions(const uint64_t* my_keys, int64_t n_my, const uint64_t* keys, int64_t n,
                              int64_t* pos, void* stream) {
  PSB_REQUIRE(pos && n >= 0 && n_my >= 0, "psb_dkv_positions: bad argument");
  if (n == 0) return 0;
  k_dkv_positions<<<grid_of(n), 256, 0, (cudaStream_t)stream>>>(my_keys, n_my, keys, n, pos);
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int psb_dmap_clear(uint64_t* table_keys, int64_t cap, void* stream) {
  PSB_REQUIRE(cap > 0 && (cap & (cap - 1)) == 0, "psb_dmap_clear: capacity must be a power of two");
  k_fill_words<<<grid_of(cap), 256, 0, (cudaStream_t)stream>>>((unsigned long long*)table_keys, cap, kDmapEmpty);
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int psb_dmap_slots(uint64_t* table_keys, int64_t cap, const uint64_t* keys, int64_t n,
                           int32_t insert, int64_t* slot, int32_t* overflow_dev, void* stream) {
  PSB_REQUIRE(cap > 0 && (cap & (cap - 1)) == 0, "psb_dmap_slots: capacity must be a power of two");
  if (n == 0) return 0;
  k_dmap_slots<<<grid_of(n), 256, 0, (cudaStream_t)stream>>>((unsigned long long*)table_keys, cap, keys, n,
                                                           insert, slot, overflow_dev);
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int psb_dmap_export(const uint64_t* table_keys, int64_t cap, uint64_t* out_keys, int64_t* out_slot,
                            uint64_t* count_dev, void* stream) {
  k_dmap_export<<<grid_of(cap), 256, 0, (cudaStream_t)stream>>>(
      (const unsigned long long*)table_keys, cap, (unsigned long long*)out_keys, out_slot,
      (unsigned long long*)count_dev);
  PSB_LAUNCH_CHECK();
  return 0;
}
