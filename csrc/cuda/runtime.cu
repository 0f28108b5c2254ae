// Error plumbing, device memory, CUDA-IPC symmetric buffers and the cross-rank device barrier.
//
// This is the B200 replacement of the reference's Van for *data*: instead of serialising
// key/value frames through ZeroMQ sockets (ref src/system/van.cc:117-223) every rank maps the
// other ranks' shard / landing / flag buffers once (CUDA IPC over NVLink) and kernels then
// read and write them directly.  Control messages stay on the host runtime.
#include "common.cuh"
#include "../../include/psb200.h"
#include <stdarg.h>
#include <string.h>

namespace psb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what, const char* file, int line) {
  if (e == cudaSuccess) return 0;
  set_error("CUDA error %d (%s) at %s:%d in %s", (int)e, cudaGetErrorString(e), file, line, what);
  return (int)e ? (int)e : 1;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!cached[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// Cross-rank barrier over peer-mapped flag words: rank r writes `epoch` into slot r of every
// peer's flag array, then waits until all slots of its own array reached `epoch`.
struct PeerPtrs {
  uint64_t* p[PSB_MAX_PEERS];
};
__global__ void k_rank_barrier_v(const __grid_constant__ PeerPtrs peers, int rank, int world,
                                 uint64_t epoch, const uint64_t* epoch_dev) {
  const int t = threadIdx.x;
  if (epoch_dev) epoch = *epoch_dev;
  if (t < world) {
    __threadfence_system();
    st_release_sys(peers.p[t] + rank, epoch);
    spin_until_ge(peers.p[rank] + t, epoch);
  }
}

}  // namespace psb

using namespace psb;

PSB_API const char* psb_last_error(void) { return g_err; }
PSB_API int psb_abi_version(void) { return PSB_ABI_VERSION; }
PSB_API int psb_device_sm_count(void) { return sm_count(); }

// ---- device memory (these DO allocate; used for buffers that must be IPC-exportable) --------
PSB_API int psb_malloc(void** out, int64_t bytes) {
  PSB_REQUIRE(out && bytes >= 0, "psb_malloc: bad argument");
  *out = nullptr;
  if (bytes == 0) return 0;
  PSB_CUDA(cudaMalloc(out, (size_t)bytes));
  return 0;
}
PSB_API int psb_free(void* p) {
  if (p) PSB_CUDA(cudaFree(p));
  return 0;
}
PSB_API int psb_memset_async(void* p, int value, int64_t bytes, void* stream) {
  PSB_CUDA(cudaMemsetAsync(p, value, (size_t)bytes, (cudaStream_t)stream));
  return 0;
}
PSB_API int psb_memcpy_async(void* dst, const void* src, int64_t bytes, void* stream) {
  PSB_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return 0;
}
PSB_API int psb_stream_sync(void* stream) {
  PSB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return 0;
}
PSB_API int psb_host_alloc_pinned(void** out, int64_t bytes) {
  PSB_CUDA(cudaHostAlloc(out, (size_t)bytes, cudaHostAllocDefault));
  return 0;
}
PSB_API int psb_host_free_pinned(void* p) {
  if (p) PSB_CUDA(cudaFreeHost(p));
  return 0;
}

// ---- CUDA IPC ---------------------------------------------------------------------------
PSB_API int psb_ipc_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }
PSB_API int psb_ipc_get_handle(void* dev_ptr, void* handle_out_host) {
  cudaIpcMemHandle_t h;
  PSB_CUDA(cudaIpcGetMemHandle(&h, dev_ptr));
  memcpy(handle_out_host, &h, sizeof(h));
  return 0;
}
PSB_API int psb_ipc_open_handle(const void* handle_host, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle_host, sizeof(h));
  PSB_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
PSB_API int psb_ipc_close_handle(void* mapped) {
  if (mapped) PSB_CUDA(cudaIpcCloseMemHandle(mapped));
  return 0;
}
// same-process multi-device fallback (tests that drive two GPUs from one process)
PSB_API int psb_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  PSB_CUDA(e);
  return 0;
}

// ---- cross-rank device barrier ------------------------------------------------------------
// peers_host[r] = this process' mapping of rank r's flag array (uint64[world]); epoch must
// increase by one per call.  No NCCL involved.
PSB_API int psb_rank_barrier(uint64_t* const* peers_host, int rank, int world, uint64_t epoch,
                             void* stream) {
  PSB_REQUIRE(world >= 1 && world <= PSB_MAX_PEERS, "psb_rank_barrier: world=%d", world);
  PeerPtrs pp;
  for (int i = 0; i < world; ++i) pp.p[i] = peers_host[i];
  k_rank_barrier_v<<<1, 32, 0, (cudaStream_t)stream>>>(pp, rank, world, epoch, nullptr);
  PSB_LAUNCH_CHECK();
  return 0;
}
// same, with the epoch read from device memory at run time (replayable inside a CUDA graph)
PSB_API int psb_rank_barrier_dev(uint64_t* const* peers_host, int rank, int world,
                                 const uint64_t* epoch_dev, void* stream) {
  PSB_REQUIRE(world >= 1 && world <= PSB_MAX_PEERS, "psb_rank_barrier_dev: world=%d", world);
  PeerPtrs pp;
  for (int i = 0; i < world; ++i) pp.p[i] = peers_host[i];
  k_rank_barrier_v<<<1, 32, 0, (cudaStream_t)stream>>>(pp, rank, world, 0, epoch_dev);
  PSB_LAUNCH_CHECK();
  return 0;
}

// ---- micro-benchmarks (used by profiles/ and tests; not on any product path) ------------------
namespace psb {
__global__ void __launch_bounds__(256, 2) k_bench_grid_barrier(uint32_t* bar, int iters, int mode) {
  const uint32_t G = gridDim.x;
  for (int i = 0; i < iters; ++i) {
    if (mode == 0) {  // split arrive/wait, relaxed polling (what k_darlin_pass uses)
      grid_arrive(bar);
      grid_wait(bar, (uint32_t)(i + 1) * G);
    } else {          // acquire polling
      grid_arrive(bar);
      if (threadIdx.x == 0)
        while (ld_acquire_gpu(bar) < (uint32_t)(i + 1) * G) {}
      __syncthreads();
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    // all CTAs passed the last wait only after everybody arrived; safe to clear for the next launch
  }
}
__global__ void k_clear_u32(uint32_t* p) { *p = 0; }
}  // namespace psb

// time per grid barrier: launch cooperative kernel with `iters` barriers; caller times it.
PSB_API int psb_bench_grid_barrier(uint32_t* bar, int32_t iters, int32_t grid, int32_t mode,
                                   void* stream) {
  psb::k_clear_u32<<<1, 1, 0, (cudaStream_t)stream>>>(bar);
  void* args[] = {(void*)&bar, (void*)&iters, (void*)&mode};
  PSB_CUDA(cudaLaunchCooperativeKernel((const void*)psb::k_bench_grid_barrier, dim3(grid), dim3(256),
                                       args, 0, (cudaStream_t)stream));
  return 0;
}
