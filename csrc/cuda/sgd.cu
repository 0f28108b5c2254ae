// Asynchronous minibatch SGD on sm_100a: sharded hash store, fused pull -> SpMV -> gradient ->
// push kernel, server-side FTRL / AdaGrad / SGD updates, block-scaled int8 push.
//
// ref (CPU loops reproduced): src/app/linear_method/async_sgd.h:101-124 (FTRLEntry), :72-99
// (AdaGrad / SGD entries as intended), :241-289 (computeGradient), src/app/linear_method/loss.h,
// penalty.h:38-43, learning_rate.h:15-22, src/parameter/kv_store.h:37-57 (KVStore),
// src/filter/fixing_float.h (lossy push), src/util/sparse_matrix.h:72-107 (SpMV).
//
// Store (K8): every server shard is an open-addressing table {key u64, w, z, n f32} in its own
// HBM.  Workers resolve a minibatch's unique keys to (owner, slot) once with CAS inserts on the
// owner's key array (peer atomics over NVLink) and cache the result per minibatch — the
// successor of the KEY_CACHING filter: afterwards only slots and values move.
//
// One worker step is ONE cooperative kernel (K6 + K7 + C3 + C4):
//   phase 1  pull: w_local[j] = owner_table.w[slot_j]        (loads from peer HBM)
//   phase 2  warp per row: margin, loss, tau; g_local[col] += -y tau x   (fp32 atomics, local)
//   phase 3  push: (slot, g) of every owner segment stored into the owner's landing region of
//            this worker, optionally as int8 with one fp32 scale per 128 values (K9), then a
//            release.sys flag
// The owner applies the landed gradients worker by worker in fixed order (k_sgd_server), which is
// what the reference's server does with arriving pushes, and makes runs reproducible.
#include "common.cuh"
#include <cuda_fp8.h>
#include "../../include/psb200.h"

namespace psb {

constexpr int kSgdThreads = 256;
constexpr int kSgdWarps = kSgdThreads / 32;
constexpr uint64_t kEmptyKey = ~0ULL;
constexpr int kQuantBlock = 128;
constexpr uint32_t kFilteredColumn = 0xffffffffu;  // col entry of a key the tail filter dropped (sgd_localize.cu)

__device__ __forceinline__ uint64_t hash_key(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

// ---- K8a: resolve keys to slots, inserting unseen keys ---------------------------------------------
struct TablePtrs {
  unsigned long long* keys[PSB_MAX_PEERS];
  int64_t cap[PSB_MAX_PEERS];
};

__global__ void k_kv_lookup_insert(const uint64_t* __restrict__ keys, int64_t n,
                                   const int32_t* __restrict__ owner, const __grid_constant__ TablePtrs T,
                                   int64_t* __restrict__ slot_out, int* __restrict__ overflow) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t key = keys[i];
    const int o = owner[i];
    unsigned long long* tk = T.keys[o];
    const uint64_t mask = (uint64_t)T.cap[o] - 1;
    uint64_t h = hash_key(key) & mask;
    int64_t found = -1;
    for (uint64_t probe = 0; probe <= mask; ++probe) {
      unsigned long long k = *reinterpret_cast<volatile unsigned long long*>(tk + h);
      if (k == kEmptyKey) k = atomicCAS_system(tk + h, (unsigned long long)kEmptyKey, (unsigned long long)key);
      if (k == kEmptyKey || k == key) {
        found = (int64_t)h;
        break;
      }
      h = (h + 1) & mask;
    }
    if (found < 0) *overflow = 1;
    slot_out[i] = found;
  }
}

// ---- bounded waits -----------------------------------------------------------------------------------
// A wait on a peer's flag gives up after ~2^21 polls (seconds) instead of hanging the GPU: it
// records {code, index, wanted, seen} in a host-pinned sink (psb_sgd_set_fault_sink) and lets
// the kernel run on, so every rank drains and the host can raise.  code 1: worker waiting for
// an owner's acknowledgement, code 2: server waiting for a worker's push.
__device__ unsigned long long* g_sgd_fault_sink = nullptr;
__device__ __forceinline__ void spin_or_fault(const uint64_t* flag, uint64_t want, int code, int index) {
  uint32_t polls = 0;
  uint64_t seen;
  while ((seen = ld_acquire_sys(flag)) < want) {
    __nanosleep(20);
    if (++polls > (1u << 21)) {
      unsigned long long* sink = g_sgd_fault_sink;
      if (sink && atomicCAS_system(sink, 0ull, (unsigned long long)code) == 0ull) {
        sink[1] = (unsigned long long)index;
        sink[2] = want;
        sink[3] = seen;
      }
      return;
    }
  }
}

// ---- worker step -------------------------------------------------------------------------------------
typedef psb_sgd_worker SgdWorker;
typedef psb_sgd_server SgdServer;

__device__ __forceinline__ float loss_and_residual(int loss, float y, float m, float* r) {
  // returns the loss of one example and sets r = d loss / d margin
  if (loss == 2) {
    const float ym = y * m;
    const float l = ym > 0 ? log1pf(__expf(-ym)) : (-ym + log1pf(__expf(ym)));
    const float tau = 1.f / (1.f + __expf(ym));
    *r = -y * tau;
    return l;
  } else if (loss == 4) {
    const float h = fmaxf(0.f, 1.f - y * m);
    *r = -2.f * y * h;
    return h * h;
  }
  const float d = m - y;
  *r = d;
  return 0.5f * d * d;
}

__global__ void __launch_bounds__(kSgdThreads, 2)
k_sgd_worker_step(const __grid_constant__ SgdWorker W, uint64_t seq, uint32_t* bar) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t gt = (int64_t)blockIdx.x * kSgdThreads + threadIdx.x, nt = (int64_t)gridDim.x * kSgdThreads;
  const int64_t gw = (int64_t)blockIdx.x * kSgdWarps + wid, nw = (int64_t)gridDim.x * kSgdWarps;
  const uint32_t G = gridDim.x;

  // Bounded delay (ref linear.proto:47 max_delay): minibatch `seq` may start only when every
  // owner has applied ALL workers' pushes up to seq - depth (depth = max_delay + 1).  The
  // acknowledgement also frees ring region seq % depth for this step's push.  Waiting here, before
  // the pull, bounds the staleness of the pulled weights from below; it does not bound it from
  // above: an owner applies a push as soon as it arrives, so a slower worker may already see part
  // of step `seq` (asynchronous SGD, as in the reference - not sequentially consistent).
  if (W.n_workers > 1 || W.depth > 1) {
    if (threadIdx.x < W.n_servers && seq > (uint64_t)W.depth)
      spin_or_fault(W.ack + threadIdx.x, seq - (uint64_t)W.depth, 1, W.rank * 100 + threadIdx.x);
    __syncthreads();
  }

  // ---- phase 1: pull the weights of this minibatch's columns from their owners ----
  const int64_t n_cols = W.n_cols_dev ? *W.n_cols_dev : W.n_cols;
  for (int64_t j = gt; j < n_cols; j += nt) {
    const int64_t s = W.col_slot[j];
    float w = 0.f;
    if (s >= 0) w = *reinterpret_cast<const volatile float*>(W.tab_w[W.col_owner[j]] + s);
    W.w_local[j] = w;
    W.g_local[j] = 0.f;
  }
  grid_arrive(bar);
  grid_wait(bar, G);

  // ---- phase 2: margins, loss, gradient ----
  double loss_acc = 0, correct = 0;
  for (int64_t r = gw; r < W.n_rows; r += nw) {
    const int64_t b = W.row_ptr[r], e = W.row_ptr[r + 1];
    float m = 0.f;
    for (int64_t k = b + lane; k < e; k += 32) {
      const uint32_t c = W.col[k];
      if (c == kFilteredColumn) continue;
      const float wv = __ldcg(W.w_local + c);
      m += W.val ? wv * W.val[k] : wv;
    }
    m = warp_sum(m);
    const float yr = W.y[r];
    float res;
    const float l = loss_and_residual(W.loss, yr, m, &res);
    if (lane == 0) {
      W.xw[r] = m;
      loss_acc += l;
      correct += ((yr > 0 && m > 0) || (yr < 0 && m <= 0)) ? 1.0 : 0.0;
    }
    if (res != 0.f)
      for (int64_t k = b + lane; k < e; k += 32) {
        const uint32_t c = W.col[k];
        if (c != kFilteredColumn) atomicAdd(W.g_local + c, W.val ? res * W.val[k] : res);
      }
  }
  if (lane == 0 && (loss_acc != 0 || correct != 0)) {
    atomicAdd(W.stats, loss_acc);
    atomicAdd(W.stats + 1, correct);
  }
  grid_arrive(bar);
  grid_wait(bar, 2 * G);

  // ---- phase 3: push every owner's segment into region seq % depth of its ring ----
  const int64_t region = (int64_t)(seq % (uint64_t)W.depth);
  const int64_t sblk = W.land_cap / kQuantBlock + 1;
  for (int s = 0; s < W.n_servers; ++s) {
    const int64_t lo = W.owner_begin[s], hi = W.owner_begin[s + 1];
    const int64_t n = min(hi - lo, W.land_cap);  // a longer segment was reported by the localisation
    int32_t* d_slot = W.land_slot[s] + region * W.land_cap;
    if (W.quant == 0) {
      float* d_val = W.land_val[s] + region * W.land_cap;
      for (int64_t i = gt; i < n; i += nt) {
        d_slot[i] = (int32_t)W.col_slot[lo + i];
        d_val[i] = __ldcg(W.g_local + lo + i);
      }
    } else {
      // one warp per block of 128 values: scale = max|g| / 127, 4 values per lane
      int8_t* d_q = W.land_q[s] + region * W.land_cap;
      float* d_scale = W.land_scale[s] + region * sblk;
      const int64_t nblk = (n + kQuantBlock - 1) / kQuantBlock;
      for (int64_t b = gw; b < nblk; b += nw) {
        float v[4];
        float mx = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int64_t i = b * kQuantBlock + k * 32 + lane;
          v[k] = i < n ? __ldcg(W.g_local + lo + i) : 0.f;
          mx = fmaxf(mx, fabsf(v[k]));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, o));
        // int8: 127 levels per sign; fp8 e4m3: the block maximum maps to 448, the largest finite
        // value, and small entries keep 3 mantissa bits instead of vanishing below one level
        const float top = W.quant == 2 ? 448.f : 127.f;
        const float scale = mx > 0.f ? mx / top : 1.f;
        if (lane == 0) d_scale[b] = scale;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int64_t i = b * kQuantBlock + k * 32 + lane;
          if (i < n) {
            d_slot[i] = (int32_t)W.col_slot[lo + i];
            if (W.quant == 2)
              d_q[i] = (int8_t)__nv_cvt_float_to_fp8(v[k] / scale, __NV_SATFINITE, __NV_E4M3);
            else
              d_q[i] = (int8_t)__float2int_rn(v[k] / scale);
          }
        }
      }
    }
  }
  grid_arrive(bar);
  if (blockIdx.x != 0) return;
  grid_wait(bar, 3 * G);
  if (threadIdx.x < W.n_servers) {
    const int s = threadIdx.x;
    W.land_cnt[s][region] = min(W.owner_begin[s + 1] - W.owner_begin[s], W.land_cap);
    __threadfence_system();
    st_release_sys(W.land_flag[s], seq);
  }
  if (threadIdx.x == 0) {
    atomicAdd(W.stats + 2, (double)W.n_rows);
    *bar = 0;
  }
}

// ---- server update -----------------------------------------------------------------------------------
__device__ __forceinline__ float prox_enet(float z, float eta, float l1, float l2) {
  const float leta = l1 * eta;
  if (z <= leta && z >= -leta) return 0.f;
  return z > 0 ? (z - leta) / (1.f + l2 * eta) : (z + leta) / (1.f + l2 * eta);
}

// apply push `seq` of worker `wkr` to the shard (threads gt, gt + nt, ...); returns the statistics deltas
__device__ __forceinline__ void sgd_apply_worker(const SgdServer& S, int wkr, uint64_t seq, int64_t gt, int64_t nt,
                                                 double& d_nnz, double& d_w2, double& d_d2) {
  const int64_t region = (int64_t)(seq % (uint64_t)S.depth);
  const int64_t n = __ldcg(S.land_cnt + (int64_t)wkr * S.depth + region);
  const int64_t base = ((int64_t)wkr * S.depth + region) * S.land_cap;
  const int64_t sbase = ((int64_t)wkr * S.depth + region) * (S.land_cap / kQuantBlock + 1);
  for (int64_t i = gt; i < n; i += nt) {
    const int32_t slot = __ldcg(S.land_slot + base + i);
    if (slot < 0) continue;
    float g;
    if (S.quant == 0) {
      g = __ldcg(S.land_val + base + i);
    } else {
      const int8_t code = __ldcg(S.land_q + base + i);
      const float unit = S.quant == 2
          ? __half2float(__half(__nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)code, __NV_E4M3)))
          : (float)code;
      g = unit * __ldcg(S.land_scale + sbase + i / kQuantBlock);
    }
    const float w_old = S.w[slot];
    float w_new;
    if (S.algo == 0) {  // FTRL-proximal
      const float sn = S.n[slot];
      const float sn_new = sqrtf(sn * sn + g * g);
      const float sigma = (sn_new - sn) / S.alpha;
      const float z = S.z[slot] + g - sigma * w_old;
      S.z[slot] = z;
      S.n[slot] = sn_new;
      const float eta = S.lr_constant ? S.alpha : S.alpha / (sn_new + S.beta);
      w_new = prox_enet(-z * eta, eta, S.l1, S.l2);
    } else if (S.algo == 1) {  // AdaGrad
      const float ss = S.z[slot] + g * g;
      S.z[slot] = ss;
      const float eta = S.lr_constant ? S.alpha : S.alpha / (sqrtf(ss) + S.beta);
      w_new = prox_enet(w_old - eta * g, eta, S.l1, S.l2);
    } else {  // SGD
      const float eta = S.lr_constant ? S.alpha : S.alpha / (S.sgd_t + S.beta);
      w_new = prox_enet(w_old - eta * g, eta, S.l1, S.l2);
    }
    S.w[slot] = w_new;
    if (w_new == 0.f && w_old != 0.f) d_nnz -= 1;
    else if (w_new != 0.f && w_old == 0.f) d_nnz += 1;
    d_w2 += (double)w_new * w_new;
    d_d2 += (double)(w_new - w_old) * (w_new - w_old);
  }
}

__device__ __forceinline__ void sgd_flush_stats(const SgdServer& S, double d_nnz, double d_w2, double d_d2,
                                                double (*sh)[kSgdWarps]) {
  d_nnz = warp_sum(d_nnz); d_w2 = warp_sum(d_w2); d_d2 = warp_sum(d_d2);
  const int wid = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { sh[0][wid] = d_nnz; sh[1][wid] = d_w2; sh[2][wid] = d_d2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0, c = 0;
    for (int i = 0; i < kSgdWarps; ++i) { a += sh[0][i]; b += sh[1][i]; c += sh[2][i]; }
    if (a != 0) atomicAdd(S.stats, a);
    if (b != 0) atomicAdd(S.stats + 1, b);
    if (c != 0) atomicAdd(S.stats + 2, c);
  }
}

// one launch per worker (rounds must not interleave: two workers may push the same key)
__global__ void __launch_bounds__(kSgdThreads)
k_sgd_server(const __grid_constant__ SgdServer S, int wkr, uint64_t seq) {
  __shared__ double sh[3][kSgdWarps];
  const int64_t gt = (int64_t)blockIdx.x * kSgdThreads + threadIdx.x, nt = (int64_t)gridDim.x * kSgdThreads;
  double d_nnz = 0, d_w2 = 0, d_d2 = 0;
  if (threadIdx.x == 0) spin_or_fault(S.land_flag + wkr, seq, 2, S.rank * 100 + wkr);
  __syncthreads();
  sgd_apply_worker(S, wkr, seq, gt, nt, d_nnz, d_w2, d_d2);
  sgd_flush_stats(S, d_nnz, d_w2, d_d2, sh);
}

// The same in ONE cooperative launch: workers in rank order, a grid barrier between two workers
// (their pushes may contain the same key), acknowledgements at the end.  2 launches per minibatch
// step instead of 2 + n_workers.
__global__ void __launch_bounds__(kSgdThreads)
k_sgd_server_all(const __grid_constant__ SgdServer S, uint64_t seq, uint32_t* bar) {
  __shared__ double sh[3][kSgdWarps];
  const int64_t gt = (int64_t)blockIdx.x * kSgdThreads + threadIdx.x, nt = (int64_t)gridDim.x * kSgdThreads;
  const uint32_t G = gridDim.x;
  double d_nnz = 0, d_w2 = 0, d_d2 = 0;
  for (int wkr = 0; wkr < S.n_workers; ++wkr) {
    if (threadIdx.x == 0) spin_or_fault(S.land_flag + wkr, seq, 2, S.rank * 100 + wkr);
    __syncthreads();
    sgd_apply_worker(S, wkr, seq, gt, nt, d_nnz, d_w2, d_d2);
    grid_arrive(bar);
    if (wkr + 1 < S.n_workers) grid_wait(bar, (uint32_t)(wkr + 1) * G);
  }
  sgd_flush_stats(S, d_nnz, d_w2, d_d2, sh);
  if (blockIdx.x != 0) return;  // everybody has arrived at the last barrier; only block 0 waits
  grid_wait(bar, (uint32_t)S.n_workers * G);
  if (threadIdx.x < S.n_workers) {
    __threadfence_system();
    st_release_sys(S.ack_out[threadIdx.x], seq);
  }
  if (threadIdx.x == 0) *bar = 0;
}

// tell every worker that its push `seq` has been applied: its ring slot may be reused
__global__ void k_sgd_ack(const __grid_constant__ SgdServer S, uint64_t seq) {
  if (threadIdx.x < S.n_workers) {
    __threadfence_system();
    st_release_sys(S.ack_out[threadIdx.x], seq);
  }
}

__global__ void k_fill_u64(unsigned long long* p, int64_t n, unsigned long long v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}

// dump the non-empty entries of a table (for saving the model)
__global__ void k_kv_export(const unsigned long long* keys, const float* w, int64_t cap,
                            unsigned long long* out_keys, float* out_w, unsigned long long* count) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cap;
       i += (int64_t)gridDim.x * blockDim.x) {
    if (keys[i] != kEmptyKey && w[i] != 0.f) {
      unsigned long long p = atomicAdd(count, 1ULL);
      out_keys[p] = keys[i];
      out_w[p] = w[i];
    }
  }
}

}  // namespace psb

using namespace psb;

static inline int grid_for(int64_t n) {
  int64_t g = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

PSB_API int psb_kv_table_clear(uint64_t* keys, int64_t cap, void* stream) {
  PSB_REQUIRE(cap > 0 && (cap & (cap - 1)) == 0, "psb_kv_table_clear: capacity must be a power of two");
  k_fill_u64<<<grid_for(cap), 256, 0, (cudaStream_t)stream>>>((unsigned long long*)keys, cap, kEmptyKey);
  PSB_LAUNCH_CHECK();
  return 0;
}

// K8a: slots of `keys` in their owners' tables (peer CAS inserts). table_keys_host[s] / cap_host[s]:
// this process' mapping of server s' key array and its capacity. *overflow_dev is set to 1 when a
// table is full.
PSB_API int psb_kv_lookup_insert(const uint64_t* keys, int64_t n, const int32_t* owner,
                                 uint64_t* const* table_keys_host, const int64_t* cap_host,
                                 int32_t n_servers, int64_t* slot_out, int32_t* overflow_dev,
                                 void* stream) {
  PSB_REQUIRE(n_servers >= 1 && n_servers <= PSB_MAX_PEERS, "psb_kv_lookup_insert: n_servers=%d", n_servers);
  if (n == 0) return 0;
  TablePtrs T;
  for (int s = 0; s < n_servers; ++s) {
    PSB_REQUIRE(cap_host[s] > 0 && (cap_host[s] & (cap_host[s] - 1)) == 0, "table capacity must be a power of two");
    T.keys[s] = (unsigned long long*)table_keys_host[s];
    T.cap[s] = cap_host[s];
  }
  k_kv_lookup_insert<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(keys, n, owner, T, slot_out, overflow_dev);
  PSB_LAUNCH_CHECK();
  return 0;
}

// `host_pinned`: 4 zeroed uint64 in pinned (mapped) host memory, or NULL to disable
PSB_API int psb_sgd_set_fault_sink(uint64_t* host_pinned) {
  unsigned long long* p = reinterpret_cast<unsigned long long*>(host_pinned);
  PSB_CUDA(cudaMemcpyToSymbol(g_sgd_fault_sink, &p, sizeof(p)));
  return 0;
}

PSB_API int psb_sgd_max_grid(void) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_sgd_worker_step, kSgdThreads, 0) != cudaSuccess ||
      per_sm < 1)
    return 0;
  return per_sm * sm_count();
}

// One worker minibatch step (pull, margins / loss / gradient, push) in one cooperative launch.
// `bar`: one zeroed uint32 owned by the caller.
PSB_API int psb_sgd_worker_step(const psb_sgd_worker* wk, uint64_t seq, int32_t grid, uint32_t* bar,
                                void* stream) {
  PSB_REQUIRE(wk && bar, "psb_sgd_worker_step: NULL argument");
  PSB_REQUIRE(grid >= 1 && grid <= psb_sgd_max_grid(), "psb_sgd_worker_step: grid=%d not co-resident", grid);
  PSB_REQUIRE(wk->n_servers <= PSB_MAX_PEERS, "n_servers > PSB_MAX_PEERS");
  PSB_REQUIRE(wk->depth >= 1, "psb_sgd_worker_step: depth must be >= 1");
  void* args[] = {(void*)wk, (void*)&seq, (void*)&bar};
  PSB_CUDA(cudaLaunchCooperativeKernel((const void*)k_sgd_worker_step, dim3(grid), dim3(kSgdThreads), args, 0,
                                       (cudaStream_t)stream));
  return 0;
}

// K8: wait on the device for push `seq` of every worker, apply them in rank order (one launch
// each so that two workers' updates of one key never interleave), then acknowledge.
PSB_API int psb_sgd_server_update(const psb_sgd_server* sv, uint64_t seq, int32_t grid, void* stream) {
  PSB_REQUIRE(sv && grid >= 1, "psb_sgd_server_update: bad argument");
  PSB_REQUIRE(sv->n_workers >= 1 && sv->n_workers <= PSB_MAX_PEERS && sv->depth >= 1, "psb_sgd_server_update: bad struct");
  for (int w = 0; w < sv->n_workers; ++w) {
    k_sgd_server<<<grid, kSgdThreads, 0, (cudaStream_t)stream>>>(*sv, w, seq);
    PSB_LAUNCH_CHECK();
  }
  k_sgd_ack<<<1, 32, 0, (cudaStream_t)stream>>>(*sv, seq);
  PSB_LAUNCH_CHECK();
  return 0;
}

// The same as psb_sgd_server_update in one cooperative launch (k_sgd_server_all).  `bar`: one zeroed
// uint32 owned by the caller; `grid` is clamped to what is co-resident.
PSB_API int psb_sgd_server_update_fused(const psb_sgd_server* sv, uint64_t seq, int32_t grid, uint32_t* bar,
                                        void* stream) {
  PSB_REQUIRE(sv && bar && grid >= 1, "psb_sgd_server_update_fused: bad argument");
  PSB_REQUIRE(sv->n_workers >= 1 && sv->n_workers <= PSB_MAX_PEERS && sv->depth >= 1,
              "psb_sgd_server_update_fused: bad struct");
  int per_sm = 0;
  PSB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_sgd_server_all, kSgdThreads, 0));
  PSB_REQUIRE(per_sm >= 1, "psb_sgd_server_update_fused: kernel does not fit on an SM");
  grid = min(grid, per_sm * sm_count());
  void* args[] = {(void*)sv, (void*)&seq, (void*)&bar};
  PSB_CUDA(cudaLaunchCooperativeKernel((const void*)k_sgd_server_all, dim3(grid), dim3(kSgdThreads), args, 0,
                                       (cudaStream_t)stream));
  return 0;
}

// non-zero weights of a shard: *count_dev must be zeroed by the caller; out arrays sized cap
PSB_API int psb_kv_export_nonzero(const uint64_t* keys, const float* w, int64_t cap, uint64_t* out_keys,
                                  float* out_w, uint64_t* count_dev, void* stream) {
  k_kv_export<<<grid_for(cap), 256, 0, (cudaStream_t)stream>>>(
      (const unsigned long long*)keys, w, cap, (unsigned long long*)out_keys, out_w,
      (unsigned long long*)count_dev);
  PSB_LAUNCH_CHECK();
  return 0;
}
