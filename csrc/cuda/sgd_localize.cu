// Minibatch localisation for asynchronous SGD without a single host synchronisation.
//
// What MinibatchReader::read does per minibatch on the CPU (ref src/learner/sgd.h:131-150:
// countUniqIndex -> CountMin insert + query -> remapIndex) plus the slicing of the key list by
// server range (src/system/message.h:120-159) and the key -> table-slot resolution, as ONE
// enqueue: every size that depends on the data (unique keys, kept keys, owner segments) stays in
// device memory and the later kernels read it from there, so the host never waits for a count.
// The step kernel (k_sgd_worker_step) takes the number of columns from the same place.
//
//   keys[nnz] --radix sort--> sorted --run-length--> uniq, cnt, *n_uniq
//     --CountMin insert, then query--> flags --select--> cols[*n_cols]         (tail filter)
//     --binary search per non-zero--> col[nnz] (0xffffffff: filtered key, skipped by the step)
//     --lower bounds at the shard boundaries--> owner_begin[W+1], col_owner[*n_cols]
//     --open-addressing lookup / CAS insert in the owners' tables (peer memory)--> col_slot
#include <cstdlib>
#include <ctime>

#include <cub/cub.cuh>

#include "common.cuh"
#include "sketch.cuh"
#include "../../include/psb200.h"

namespace psb {

constexpr unsigned long long kTableEmpty = ~0ULL;

__device__ __forceinline__ uint64_t table_hash(uint64_t x) {  // the table hash of sgd.cu
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

__global__ void k_loc_sketch_insert(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ counts,
                                    const int64_t* __restrict__ n_dev, uint8_t* table, uint32_t cm_n, int cm_k) {
  const int64_t n = *n_dev;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t h = sketch_hash(keys[i]);
    const uint32_t delta = (h >> 17) | (h << 15);
    const uint32_t c = min(counts[i], 255u);
    for (int j = 0; j < cm_k; ++j) {
      sat_add_u8(table, h % cm_n, c, 254u);
      h += delta;
    }
  }
}
// flag[i] = key i is kept; entries at and beyond *n get 0 so that a select over the capacity works
__global__ void k_loc_sketch_query(const uint64_t* __restrict__ keys, const int64_t* __restrict__ n_dev, int64_t cap,
                                   const uint8_t* __restrict__ table, uint32_t cm_n, int cm_k, int freq,
                                   uint8_t* __restrict__ flag) {
  const int64_t n = *n_dev;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += (int64_t)gridDim.x * blockDim.x) {
    uint8_t keep = 0;
    if (i < n) {
      if (!table) {
        keep = 1;
      } else {
        uint32_t h = sketch_hash(keys[i]);
        const uint32_t delta = (h >> 17) | (h << 15);
        uint32_t est = 255;
        for (int j = 0; j < cm_k; ++j) {
          est = min(est, (uint32_t)table[h % cm_n]);
          h += delta;
        }
        keep = (int)est > freq;
      }
    }
    flag[i] = keep;
  }
}

// local column of every non-zero (position of its key among the kept keys)
__global__ void k_loc_columns(const uint64_t* __restrict__ cols, const int64_t* __restrict__ n_cols_dev,
                              const uint64_t* __restrict__ keys, int64_t nnz, uint32_t* __restrict__ col) {
  const int64_t nd = *n_cols_dev;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t k = keys[i];
    int64_t lo = 0, hi = nd;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (cols[mid] < k) lo = mid + 1; else hi = mid;
    }
    col[i] = (lo < nd && cols[lo] == k) ? (uint32_t)lo : 0xffffffffu;
  }
}

// owner_begin[s] = first column whose key lies in server s' range; one thread per boundary
__global__ void k_loc_owner_begin(const uint64_t* __restrict__ cols, const int64_t* __restrict__ n_cols_dev,
                                  const uint64_t* __restrict__ bounds, int n_servers, int64_t land_cap,
                                  int64_t* __restrict__ owner_begin, int* __restrict__ overflow) {
  const int64_t n = *n_cols_dev;
  const int s = threadIdx.x;
  if (s > n_servers) return;
  int64_t at = 0;
  if (s == n_servers) {
    at = n;
  } else if (s > 0) {
    const uint64_t b = bounds[s - 1];
    int64_t lo = 0, hi = n;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (cols[mid] < b) lo = mid + 1; else hi = mid;
    }
    at = lo;
  }
  owner_begin[s] = at;
  __syncthreads();
  if (s < n_servers && owner_begin[s + 1] - owner_begin[s] > land_cap) *overflow = 2;  // ring region too small
}

struct LocTables {
  unsigned long long* keys[PSB_MAX_PEERS];
  int64_t cap[PSB_MAX_PEERS];
};

// owner of every column and its slot in the owner's table (inserted when new)
__global__ void k_loc_slots(const uint64_t* __restrict__ cols, const int64_t* __restrict__ n_cols_dev,
                            const int64_t* __restrict__ owner_begin, int n_servers,
                            const __grid_constant__ LocTables T, int32_t* __restrict__ col_owner,
                            int64_t* __restrict__ col_slot, int* __restrict__ overflow) {
  const int64_t n = *n_cols_dev;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int o = 0;
    while (o + 1 < n_servers && i >= owner_begin[o + 1]) ++o;
    col_owner[i] = o;
    const unsigned long long key = cols[i];
    unsigned long long* tk = T.keys[o];
    const uint64_t mask = (uint64_t)T.cap[o] - 1;
    uint64_t h = table_hash(key) & mask;
    int64_t found = -1;
    for (uint64_t probe = 0; probe <= mask; ++probe) {
      unsigned long long k = *reinterpret_cast<volatile unsigned long long*>(tk + h);
      if (k == kTableEmpty) k = atomicCAS_system(tk + h, kTableEmpty, key);
      if (k == kTableEmpty || k == key) { found = (int64_t)h; break; }
      h = (h + 1) & mask;
    }
    if (found < 0) *overflow = 1;
    col_slot[i] = found;
  }
}

// PSB_LOC_TIMING=1: synchronise after every stage and accumulate its wall time (diagnosis only: the
// synchronisations serialise the copy stream against the host)
struct LocTiming {
  bool on = false;
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long calls = 0;
  LocTiming() { const char* e = getenv("PSB_LOC_TIMING"); on = e && *e && *e != '0'; }
};
static LocTiming g_loc_timing;
static inline double now_s() {
  timespec t; clock_gettime(CLOCK_MONOTONIC, &t);
  return t.tv_sec + 1e-9 * t.tv_nsec;
}
#define LOC_LAP(k)                                         \
  if (g_loc_timing.on) {                                   \
    cudaStreamSynchronize(st);                             \
    const double t_now = now_s();                          \
    g_loc_timing.acc[k] += t_now - t_lap;                  \
    t_lap = t_now;                                         \
  }

static inline int grid_of(int64_t n) {
  int64_t g = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace psb

using namespace psb;

// Bytes of temporary storage psb_sgd_localize needs for minibatches of at most `max_nnz` non-zeros.
PSB_API int64_t psb_sgd_localize_temp_bytes(int64_t max_nnz) {
  size_t a = 0, b = 0, c = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, a, (const uint64_t*)nullptr, (uint64_t*)nullptr, max_nnz);
  cub::DeviceRunLengthEncode::Encode(nullptr, b, (const uint64_t*)nullptr, (uint64_t*)nullptr, (uint32_t*)nullptr,
                                     (int64_t*)nullptr, (int)max_nnz);
  cub::DeviceSelect::Flagged(nullptr, c, (const uint64_t*)nullptr, (const uint8_t*)nullptr, (uint64_t*)nullptr,
                             (int64_t*)nullptr, (int)max_nnz);
  size_t m = a > b ? a : b;
  m = m > c ? m : c;
  return (int64_t)m + 256;
}

// The localisation in two stages, so that a caller may overlap stage A of one minibatch with stage B of
// the one before (stage B touches the sketch, whose updates must stay in minibatch order).
//
// Stage A: keys[nnz] -> sorted[nnz] -> uniq / cnt (capacity nnz), counts_dev[0] = number of unique keys.
PSB_API int psb_sgd_localize_sort(const uint64_t* keys, int64_t nnz, uint64_t* sorted, uint64_t* uniq, uint32_t* cnt,
                                  int64_t* counts_dev, void* temp, int64_t temp_bytes, void* stream) {
  PSB_REQUIRE(nnz >= 0 && nnz < (1LL << 31), "psb_sgd_localize: nnz=%lld", (long long)nnz);
  cudaStream_t st = (cudaStream_t)stream;
  if (nnz == 0) {
    PSB_CUDA(cudaMemsetAsync(counts_dev, 0, 2 * sizeof(int64_t), st));
    return 0;
  }
  double t_lap = 0;
  if (g_loc_timing.on) { cudaStreamSynchronize(st); t_lap = now_s(); ++g_loc_timing.calls; }
  size_t bytes = (size_t)temp_bytes;
  PSB_CUDA(cub::DeviceRadixSort::SortKeys(temp, bytes, keys, sorted, nnz, 0, 64, st));
  LOC_LAP(0)
  bytes = (size_t)temp_bytes;
  PSB_CUDA(cub::DeviceRunLengthEncode::Encode(temp, bytes, sorted, uniq, cnt, counts_dev, (int)nnz, st));
  LOC_LAP(1)
  return 0;
}

// Stage B: tail filter (sketch == NULL: none) -> cols (kept keys), counts_dev[1] = their number -> col (local
// column per non-zero, 0xffffffff = filtered) -> owner_begin[W+1], col_owner -> col_slot.
// bounds_dev[W-1]: first key of servers 1..W-1.  table_keys_host[s] / cap_host[s]: this process' mapping
// of server s' key table.  *overflow_dev: 1 = a table is full, 2 = an owner segment exceeds land_cap.
PSB_API int psb_sgd_localize_finish(const uint64_t* keys, int64_t nnz, const uint64_t* uniq, const uint32_t* cnt,
                                    uint8_t* flag, uint64_t* cols, uint32_t* col, int64_t* counts_dev,
                                    uint8_t* sketch, int32_t cm_n, int32_t cm_k, int32_t freq,
                                    const uint64_t* bounds_dev, int32_t n_servers, int64_t land_cap,
                                    int64_t* owner_begin, uint64_t* const* table_keys_host, const int64_t* cap_host,
                                    int32_t* col_owner, int64_t* col_slot, int32_t* overflow_dev, void* temp,
                                    int64_t temp_bytes, void* stream) {
  PSB_REQUIRE(n_servers >= 1 && n_servers <= PSB_MAX_PEERS, "psb_sgd_localize: n_servers=%d", n_servers);
  cudaStream_t st = (cudaStream_t)stream;
  if (nnz == 0) {
    PSB_CUDA(cudaMemsetAsync(owner_begin, 0, (n_servers + 1) * sizeof(int64_t), st));
    return 0;
  }
  double t_lap = 0;
  if (g_loc_timing.on) { cudaStreamSynchronize(st); t_lap = now_s(); }
  if (sketch) {
    PSB_REQUIRE(cm_n >= 64 && cm_n % 4 == 0 && cm_k >= 1 && cm_k <= 30, "psb_sgd_localize: bad sketch size");
    k_loc_sketch_insert<<<grid_of(nnz), 256, 0, st>>>(uniq, cnt, counts_dev, sketch, (uint32_t)cm_n, cm_k);
  }
  k_loc_sketch_query<<<grid_of(nnz), 256, 0, st>>>(uniq, counts_dev, nnz, sketch, (uint32_t)cm_n, cm_k, freq, flag);
  LOC_LAP(2)
  size_t bytes = (size_t)temp_bytes;
  PSB_CUDA(cub::DeviceSelect::Flagged(temp, bytes, uniq, flag, cols, counts_dev + 1, (int)nnz, st));
  LOC_LAP(3)
  k_loc_columns<<<grid_of(nnz), 256, 0, st>>>(cols, counts_dev + 1, keys, nnz, col);
  LOC_LAP(4)
  k_loc_owner_begin<<<1, 32, 0, st>>>(cols, counts_dev + 1, bounds_dev, n_servers, land_cap, owner_begin, overflow_dev);
  LOC_LAP(5)
  LocTables T;
  for (int s = 0; s < n_servers; ++s) {
    PSB_REQUIRE(cap_host[s] > 0 && (cap_host[s] & (cap_host[s] - 1)) == 0, "table capacity must be a power of two");
    T.keys[s] = (unsigned long long*)table_keys_host[s];
    T.cap[s] = cap_host[s];
  }
  k_loc_slots<<<grid_of(nnz), 256, 0, st>>>(cols, counts_dev + 1, owner_begin, n_servers, T, col_owner, col_slot,
                                           overflow_dev);
  PSB_LAUNCH_CHECK();
  LOC_LAP(6)
  return 0;
}

// Both stages as one enqueue, no host synchronisation.  Device buffers of capacity nnz: sorted, uniq, cnt,
// flag, cols, col, col_owner, col_slot; counts_dev[2] <- {unique keys, kept keys = columns}.
PSB_API int psb_sgd_localize(const uint64_t* keys, int64_t nnz, uint64_t* sorted, uint64_t* uniq, uint32_t* cnt,
                             uint8_t* flag, uint64_t* cols, uint32_t* col, int64_t* counts_dev,
                             uint8_t* sketch, int32_t cm_n, int32_t cm_k, int32_t freq,
                             const uint64_t* bounds_dev, int32_t n_servers, int64_t land_cap, int64_t* owner_begin,
                             uint64_t* const* table_keys_host, const int64_t* cap_host, int32_t* col_owner,
                             int64_t* col_slot, int32_t* overflow_dev, void* temp, int64_t temp_bytes,
                             void* stream) {
  int rc = psb_sgd_localize_sort(keys, nnz, sorted, uniq, cnt, counts_dev, temp, temp_bytes, stream);
  if (rc) return rc;
  return psb_sgd_localize_finish(keys, nnz, uniq, cnt, flag, cols, col, counts_dev, sketch, cm_n, cm_k, freq,
                                 bounds_dev, n_servers, land_cap, owner_begin, table_keys_host, cap_host, col_owner,
                                 col_slot, overflow_dev, temp, temp_bytes, stream);
}

// PSB_LOC_TIMING=1: mean milliseconds per call of {sort, run-length, sketch insert + query, select,
// columns, owner boundaries, table slots}; returns the number of calls (0: timing is off)
PSB_API int64_t psb_sgd_localize_timing(double* out7) {
  if (!g_loc_timing.on || !g_loc_timing.calls) return 0;
  for (int k = 0; k < 7; ++k) out7[k] = 1e3 * g_loc_timing.acc[k] / (double)g_loc_timing.calls;
  return g_loc_timing.calls;
}
