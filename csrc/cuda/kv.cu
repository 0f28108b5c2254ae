// Generic sharded KV push / pull over peer memory (C1-C4 in SURVEY 2.9 as plain data movement).
//
// ref: SharedParameter::push / pull + KVVector::setValue / getValue (src/parameter/
// shared_parameter.h:17-29,97-162, kv_vector.h:46-93) and the slicing of a sorted key list by
// server range (src/system/message.h:120-159).  A handle (owner[], slot[]) is resolved once per
// key set (K15 + dictionary lookup or hash insert); afterwards
//   push  = every value is added into the owner's array with a reduction that travels over
//           NVLink (red.global.add, 128-bit wide where 4 consecutive slots are contiguous)
//   pull  = every value is loaded from the owner's array in peer memory
// No staging buffer, no host involvement, no NCCL.
#include "common.cuh"
#include "../../include/psb200.h"

namespace psb {

struct KvPeers {
  float* val[PSB_MAX_PEERS];
};

// groups of 4 consecutive elements with the same owner and consecutive, 16-byte aligned slots
// move as one 128-bit transaction
__global__ void k_kv_push_add_f32(const float* __restrict__ vals, const int32_t* __restrict__ owner,
                                  const int64_t* __restrict__ slot, int64_t n,
                                  const __grid_constant__ KvPeers P) {
  const int64_t n4 = n / 4;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n4 + (n & 3);
       q += (int64_t)gridDim.x * blockDim.x) {
    if (q < n4) {
      const int64_t i = q * 4;
      const int o = owner[i];
      const int64_t s0 = slot[i];
      const float4 v = *reinterpret_cast<const float4*>(vals + i);
      const bool run = owner[i + 3] == o && slot[i + 3] == s0 + 3 && slot[i + 1] == s0 + 1 &&
                       slot[i + 2] == s0 + 2 && owner[i + 1] == o && owner[i + 2] == o && (s0 & 3) == 0;
      if (run) {
        atomicAdd(reinterpret_cast<float4*>(P.val[o] + s0), v);
      } else {
        atomicAdd(P.val[owner[i]] + slot[i], v.x);
        atomicAdd(P.val[owner[i + 1]] + slot[i + 1], v.y);
        atomicAdd(P.val[owner[i + 2]] + slot[i + 2], v.z);
        atomicAdd(P.val[owner[i + 3]] + slot[i + 3], v.w);
      }
    } else {
      const int64_t i = n4 * 4 + (q - n4);
      atomicAdd(P.val[owner[i]] + slot[i], vals[i]);
    }
  }
}

__global__ void k_kv_pull_f32(float* __restrict__ out, const int32_t* __restrict__ owner,
                              const int64_t* __restrict__ slot, int64_t n,
                              const __grid_constant__ KvPeers P) {
  const int64_t n4 = n / 4;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n4 + (n & 3);
       q += (int64_t)gridDim.x * blockDim.x) {
    if (q < n4) {
      const int64_t i = q * 4;
      const int o = owner[i];
      const int64_t s0 = slot[i];
      const bool run = owner[i + 3] == o && slot[i + 3] == s0 + 3 && slot[i + 1] == s0 + 1 &&
                       slot[i + 2] == s0 + 2 && owner[i + 1] == o && owner[i + 2] == o && (s0 & 3) == 0;
      float4 v;
      if (run) {
        asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(P.val[o] + s0) : "memory");
      } else {
        v.x = *reinterpret_cast<const volatile float*>(P.val[owner[i]] + slot[i]);
        v.y = *reinterpret_cast<const volatile float*>(P.val[owner[i + 1]] + slot[i + 1]);
        v.z = *reinterpret_cast<const volatile float*>(P.val[owner[i + 2]] + slot[i + 2]);
        v.w = *reinterpret_cast<const volatile float*>(P.val[owner[i + 3]] + slot[i + 3]);
      }
      *reinterpret_cast<float4*>(out + i) = v;
    } else {
      const int64_t i = n4 * 4 + (q - n4);
      out[i] = *reinterpret_cast<const volatile float*>(P.val[owner[i]] + slot[i]);
    }
  }
}

// owner and slot of sorted keys against a global sorted dictionary that is split into
// contiguous shards: shard_begin[s] = first dictionary position of server s (K15 result)
__global__ void k_kv_route_dense(const uint32_t* __restrict__ idx1, int64_t n,
                                 const int64_t* __restrict__ shard_begin, int n_servers,
                                 int32_t* __restrict__ owner, int64_t* __restrict__ slot) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pos = (int64_t)idx1[i] - 1;  // lookup_sorted result; -1 = unknown key
    int o = 0;
    while (o + 1 < n_servers && pos >= shard_begin[o + 1]) ++o;
    owner[i] = pos < 0 ? 0 : o;
    slot[i] = pos < 0 ? -1 : pos - shard_begin[o];
  }
}

}  // namespace psb

using namespace psb;

static inline int grid_for(int64_t n) {
  int64_t g = (n + 255) / 256, cap = (int64_t)sm_count() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// vals[i] is added to server owner[i]'s value array at slot[i]; peers_host[s] = mapping of server s'
// float array.  vals must be 16-byte aligned.
PSB_API int psb_kv_push_add_f32(const float* vals, const int32_t* owner, const int64_t* slot, int64_t n,
                                float* const* peers_host, int32_t n_servers, void* stream) {
  PSB_REQUIRE(n_servers >= 1 && n_servers <= PSB_MAX_PEERS, "psb_kv_push_add_f32: n_servers=%d", n_servers);
  if (n == 0) return 0;
  KvPeers P;
  for (int s = 0; s < n_servers; ++s) P.val[s] = peers_host[s];
  k_kv_push_add_f32<<<grid_for(n / 4 + 4), 256, 0, (cudaStream_t)stream>>>(vals, owner, slot, n, P);
  PSB_LAUNCH_CHECK();
  return 0;
}
PSB_API int psb_kv_pull_f32(float* out, const int32_t* owner, const int64_t* slot, int64_t n,
                            float* const* peers_host, int32_t n_servers, void* stream) {
  PSB_REQUIRE(n_servers >= 1 && n_servers <= PSB_MAX_PEERS, "psb_kv_pull_f32: n_servers=%d", n_servers);
  if (n == 0) return 0;
  KvPeers P;
  for (int s = 0; s < n_servers; ++s) P.val[s] = peers_host[s];
  k_kv_pull_f32<<<grid_for(n / 4 + 4), 256, 0, (cudaStream_t)stream>>>(out, owner, slot, n, P);
  PSB_LAUNCH_CHECK();
  return 0;
}
PSB_API int psb_kv_route_dense(const uint32_t* idx1, int64_t n, const int64_t* shard_begin_dev,
                               int32_t n_servers, int32_t* owner, int64_t* slot, void* stream) {
  if (n == 0) return 0;
  k_kv_route_dense<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(idx1, n, shard_begin_dev, n_servers, owner, slot);
  PSB_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Bulk path: bandwidth-bound pushes / pulls.  Scattered 4-byte reductions over NVLink pay a
// packet per element; here only *contiguous value segments* cross the link (128-bit stores into
// the owner's landing region), and the owner applies them locally with the slot list it cached
// when the handle was registered (the KEY_CACHING idea: keys/slots travel once).
//   bulk_send        worker: segment of every owner -> that owner's landing region, then flag
//   bulk_apply_add   owner : wait flags, val[slot_w[i]] += landing_w[i], worker by worker
//   bulk_gather_send owner : landing of worker w <- val[slot_w[i]] (peer stores), then flag
// ---------------------------------------------------------------------------------------------
namespace psb {

struct BulkSend {
  const float* src;                     // values sorted by owner
  int64_t seg_begin[PSB_MAX_PEERS + 1];
  float* dst[PSB_MAX_PEERS];            // on owner s: landing region of this worker
  uint64_t* flag[PSB_MAX_PEERS];        // on owner s: &flag[rank]
  uint32_t* done_ctr;
  int32_t n_servers;
};

__global__ void __launch_bounds__(256) k_kv_bulk_send(const __grid_constant__ BulkSend B, uint64_t seq) {
  const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
  for (int s = 0; s < B.n_servers; ++s) {
    const int64_t lo = B.seg_begin[s], n = B.seg_begin[s + 1] - lo;
    const float* src = B.src + lo;
    float* dst = B.dst[s];
    // the landing region is 16-byte aligned: 128-bit peer stores, (possibly unaligned) local loads
    const int64_t n4 = n / 4;
    for (int64_t i = gt; i < n4; i += nt) {
      float4 v;
      v.x = src[4 * i]; v.y = src[4 * i + 1]; v.z = src[4 * i + 2]; v.w = src[4 * i + 3];
      *reinterpret_cast<float4*>(dst + 4 * i) = v;
    }
    for (int64_t i = 4 * n4 + gt; i < n; i += nt) dst[i] = src[i];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(B.done_ctr, 1u) == gridDim.x - 1) {
      *B.done_ctr = 0;
      __threadfence_system();
      for (int s = 0; s < B.n_servers; ++s) st_release_sys(B.flag[s], seq);
    }
  }
}

__global__ void __launch_bounds__(256)
k_kv_bulk_apply_add(float* __restrict__ val, const float* __restrict__ landing,
                    const int64_t* __restrict__ slots, int64_t n, const uint64_t* flag, uint64_t seq) {
  if (threadIdx.x == 0) spin_until_ge(flag, seq);
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    val[slots[i]] += __ldcg(landing + i);  // slots are unique inside one worker's segment
}

struct BulkGather {
  const float* val;
  const int64_t* slots[PSB_MAX_PEERS];  // slot list of worker w (cached at register time)
  int64_t count[PSB_MAX_PEERS];
  float* dst[PSB_MAX_PEERS];            // on worker w: where this owner's segment goes
  uint64_t* flag[PSB_MAX_PEERS];        // on worker w: &flag[rank]
  uint32_t* done_ctr;
  int32_t n_workers;
};

__global__ void __launch_bounds__(256) k_kv_bulk_gather_send(const __grid_constant__ BulkGather B, uint64_t seq) {
  const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
  for (int w = 0; w < B.n_workers; ++w) {
    const int64_t* sl = B.slots[w];
    float* dst = B.dst[w];
    for (int64_t i = gt; i < B.count[w]; i += nt) dst[i] = B.val[sl[i]];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(B.done_ctr, 1u) == gridDim.x - 1) {
      *B.done_ctr = 0;
      __threadfence_system();
      for (int w = 0; w < B.n_workers; ++w) st_release_sys(B.flag[w], seq);
    }
  }
}

__global__ void k_kv_wait_flags(const uint64_t* flags, int n, uint64_t seq) {
  if (threadIdx.x < n) spin_until_ge(flags + threadIdx.x, seq);
}

}  // namespace psb

// ---------------------------------------------------------------------------------------------
// Bulk path, second generation: receiver-driven push.
//
// A worker leaves its values (sorted by owner) in a staging buffer of its OWN arena; after a rank
// barrier every owner runs ONE cooperative kernel that walks the workers in a staggered order
// (rank, rank+1, ...: no two owners read the same worker at the same time, so no egress link is
// oversubscribed), loads that worker's segment straight from peer memory with 128-bit loads and
// adds it into its shard at the cached 32-bit slots.  The payload crosses NVLink once and is
// never written to or re-read from a landing region: per key the owner moves 4 B over the link,
// 4 B of slot and one 8-byte read-modify-write - against 20 B with a landing region.  A grid
// barrier separates two workers (their segments may hit the same slot), which also makes the
// sum order fixed: run-to-run deterministic.
// ---------------------------------------------------------------------------------------------
namespace psb {

struct BulkPullApply {
  float* val;
  const float* src[PSB_MAX_PEERS];     // worker w's staging, already offset to its segment for me
  const int32_t* slots[PSB_MAX_PEERS]; // slots of that segment in my shard
  int64_t count[PSB_MAX_PEERS];
  int32_t n_workers, rank;
};

__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_peer_f1(const float* p) {
  float v;
  asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(256) k_kv_bulk_pull_apply(const __grid_constant__ BulkPullApply B, uint32_t* bar) {
  const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
  for (int k = 0; k < B.n_workers; ++k) {
    const int w = (B.rank + k) % B.n_workers;
    const float* src = B.src[w];
    const int32_t* sl = B.slots[w];
    const int64_t n = B.count[w];
    // scalar head up to the first 16-byte boundary of the peer segment, then 4 values per load
    int64_t head = ((16 - ((uintptr_t)src & 15)) & 15) / 4;
    if (head > n) head = n;
    for (int64_t i = gt; i < head; i += nt) B.val[sl[i]] += ld_peer_f1(src + i);
    const int64_t n4 = (n - head) / 4;
    // two independent 128-bit peer loads per thread and trip: NVLink latency (~2 us) times the
    // link rate needs a few MB in flight
    int64_t q = gt;
    for (; q + nt < n4; q += 2 * nt) {
      const int64_t i = head + 4 * q, j = head + 4 * (q + nt);
      const float4 v = ld_peer_f4(src + i);
      const float4 u = ld_peer_f4(src + j);
      const int32_t s0 = sl[i], s1 = sl[i + 1], s2 = sl[i + 2], s3 = sl[i + 3];
      const int32_t t0 = sl[j], t1 = sl[j + 1], t2 = sl[j + 2], t3 = sl[j + 3];
      B.val[s0] += v.x; B.val[s1] += v.y; B.val[s2] += v.z; B.val[s3] += v.w;  // slots of one worker are distinct
      B.val[t0] += u.x; B.val[t1] += u.y; B.val[t2] += u.z; B.val[t3] += u.w;
    }
    for (; q < n4; q += nt) {
      const int64_t i = head + 4 * q;
      const float4 v = ld_peer_f4(src + i);
      const int32_t s0 = sl[i], s1 = sl[i + 1], s2 = sl[i + 2], s3 = sl[i + 3];
      B.val[s0] += v.x; B.val[s1] += v.y; B.val[s2] += v.z; B.val[s3] += v.w;
    }
    for (int64_t i = head + 4 * n4 + gt; i < n; i += nt) B.val[sl[i]] += ld_peer_f1(src + i);
    grid_arrive(bar);
    if (k + 1 < B.n_workers) grid_wait(bar, (uint32_t)(k + 1) * gridDim.x);
  }
  if (blockIdx.x == 0) {
    grid_wait(bar, (uint32_t)B.n_workers * gridDim.x);
    if (threadIdx.x == 0) *bar = 0;
  }
}

// owner -> workers: gather my values at worker w's slots and store them contiguously into w's
// landing buffer (128-bit peer stores), workers in staggered order, then one flag per worker
struct BulkGather32 {
  const float* val;
  const int32_t* slots[PSB_MAX_PEERS];
  int64_t count[PSB_MAX_PEERS];
  float* dst[PSB_MAX_PEERS];            // on worker w: where this owner's segment goes (16-byte aligned)
  uint64_t* flag[PSB_MAX_PEERS];        // on worker w: &flag[rank]
  uint32_t* done_ctr;
  int32_t n_workers, rank;
};

__global__ void __launch_bounds__(256) k_kv_bulk_gather_send32(const __grid_constant__ BulkGather32 B, uint64_t seq) {
  const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
  for (int k = 0; k < B.n_workers; ++k) {
    const int w = (B.rank + k) % B.n_workers;
    const int32_t* sl = B.slots[w];
    float* dst = B.dst[w];
    const int64_t n = B.count[w], n4 = n / 4;
    for (int64_t q = gt; q < n4; q += nt) {
      const int4 s4 = *reinterpret_cast<const int4*>(sl + 4 * q);
      float4 v;
      v.x = B.val[s4.x]; v.y = B.val[s4.y]; v.z = B.val[s4.z]; v.w = B.val[s4.w];
      *reinterpret_cast<float4*>(dst + 4 * q) = v;
    }
    for (int64_t i = 4 * n4 + gt; i < n; i += nt) dst[i] = B.val[sl[i]];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(B.done_ctr, 1u) == gridDim.x - 1) {
      *B.done_ctr = 0;
      __threadfence_system();
      for (int w = 0; w < B.n_workers; ++w) st_release_sys(B.flag[w], seq);
    }
  }
}

}  // namespace psb

// ONE cooperative launch: every worker's segment is read from peer memory and added into `val`.
// src_host[w]: this process' mapping of worker w's staging, offset to its segment for this owner;
// slots_host[w]: device array of int32 slots (16-byte aligned); bar: zeroed uint32.  The caller
// guarantees (rank barrier) that every worker's staging is complete before the launch.
PSB_API int psb_kv_bulk_pull_apply(float* val, const float* const* src_host, const int32_t* const* slots_host,
                                   const int64_t* count_host, int32_t n_workers, int32_t rank, uint32_t* bar,
                                   void* stream) {
  PSB_REQUIRE(n_workers >= 1 && n_workers <= PSB_MAX_PEERS, "psb_kv_bulk_pull_apply: n_workers=%d", n_workers);
  BulkPullApply B;
  B.val = val;
  int64_t most = 0;
  for (int w = 0; w < n_workers; ++w) {
    B.src[w] = src_host[w]; B.slots[w] = slots_host[w]; B.count[w] = count_host[w];
    most = count_host[w] > most ? count_host[w] : most;
  }
  B.n_workers = n_workers;
  B.rank = rank;
  int per_sm = 0;
  PSB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_kv_bulk_pull_apply, 256, 0));
  int grid = per_sm * sm_count();
  const int64_t want = (most / 4 + 255) / 256 + 1;
  if (want < grid) grid = (int)want;
  void* args[] = {(void*)&B, (void*)&bar};
  PSB_CUDA(cudaLaunchCooperativeKernel((const void*)k_kv_bulk_pull_apply, dim3(grid), dim3(256), args, 0,
                                       (cudaStream_t)stream));
  return 0;
}

PSB_API int psb_kv_bulk_gather_send32(const float* val, const int32_t* const* slots_host, const int64_t* count_host,
                                      float* const* dst_host, uint64_t* const* flag_host, int32_t n_workers,
                                      int32_t rank, uint64_t seq, uint32_t* done_ctr, void* stream) {
  PSB_REQUIRE(n_workers >= 1 && n_workers <= PSB_MAX_PEERS, "psb_kv_bulk_gather_send32: n_workers=%d", n_workers);
  BulkGather32 B;
  B.val = val;
  int64_t total = 0;
  for (int w = 0; w < n_workers; ++w) {
    B.slots[w] = slots_host[w]; B.count[w] = count_host[w]; B.dst[w] = dst_host[w]; B.flag[w] = flag_host[w];
    total += count_host[w];
  }
  B.done_ctr = done_ctr;
  B.n_workers = n_workers;
  B.rank = rank;
  k_kv_bulk_gather_send32<<<grid_for(total / 4 + 1), 256, 0, (cudaStream_t)stream>>>(B, seq);
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int psb_kv_bulk_send(const float* src, const int64_t* seg_begin_host, float* const* dst_host,
                             uint64_t* const* flag_host, int32_t n_servers, uint64_t seq,
                             uint32_t* done_ctr, void* stream) {
  PSB_REQUIRE(n_servers >= 1 && n_servers <= PSB_MAX_PEERS, "psb_kv_bulk_send: n_servers=%d", n_servers);
  BulkSend B;
  B.src = src;
  for (int s = 0; s <= n_servers; ++s) B.seg_begin[s] = seg_begin_host[s];
  for (int s = 0; s < n_servers; ++s) { B.dst[s] = dst_host[s]; B.flag[s] = flag_host[s]; }
  B.done_ctr = done_ctr;
  B.n_servers = n_servers;
  const int64_t n = seg_begin_host[n_servers];
  k_kv_bulk_send<<<grid_for(n / 4 + 4), 256, 0, (cudaStream_t)stream>>>(B, seq);
  PSB_LAUNCH_CHECK();
  return 0;
}
// one launch per worker, in rank order
PSB_API int psb_kv_bulk_apply_add(float* val, const float* landing, const int64_t* slots, int64_t n,
                                  const uint64_t* flag, uint64_t seq, void* stream) {
  k_kv_bulk_apply_add<<<grid_for(n + 1), 256, 0, (cudaStream_t)stream>>>(val, landing, slots, n, flag, seq);
  PSB_LAUNCH_CHECK();
  return 0;
}
PSB_API int psb_kv_bulk_gather_send(const float* val, const int64_t* const* slots_host, const int64_t* count_host,
                                    float* const* dst_host, uint64_t* const* flag_host, int32_t n_workers,
                                    uint64_t seq, uint32_t* done_ctr, void* stream) {
  PSB_REQUIRE(n_workers >= 1 && n_workers <= PSB_MAX_PEERS, "psb_kv_bulk_gather_send: n_workers=%d", n_workers);
  BulkGather B;
  B.val = val;
  int64_t total = 0;
  for (int w = 0; w < n_workers; ++w) {
    B.slots[w] = slots_host[w]; B.count[w] = count_host[w]; B.dst[w] = dst_host[w]; B.flag[w] = flag_host[w];
    total += count_host[w];
  }
  B.done_ctr = done_ctr;
  B.n_workers = n_workers;
  k_kv_bulk_gather_send<<<grid_for(total + 1), 256, 0, (cudaStream_t)stream>>>(B, seq);
  PSB_LAUNCH_CHECK();
  return 0;
}
PSB_API int psb_kv_wait_flags(const uint64_t* flags, int32_t n, uint64_t seq, void* stream) {
  k_kv_wait_flags<<<1, 32, 0, (cudaStream_t)stream>>>(flags, n, seq);
  PSB_LAUNCH_CHECK();
  return 0;
}
