// The streaming minibatch loop of asynchronous SGD in native code.
//
// The reference's worker loop is C++ (src/learner/sgd.h:120-150 reader thread + the app's
// processMinibatch, src/app/linear_method/async_sgd.h:217-290); here the loop body is a fixed
// sequence of enqueues, so the runner issues it without the interpreter in between:
//
//   copy stream k % L:  [wait: the step that last read this ring set]  H2D row_ptr / keys / vals / y
//                 -> stage A (sort, run-length)  -> [wait `ready` of minibatch k - 1: the sketch is updated
//                 in minibatch order]  -> stage B (tail filter, columns, owners, slots)  -> event `ready`
//   step stream:  [wait `ready`]  worker step (pull, gradient, push)  ->  server update  -> event `done`
//
// L copy streams ("lanes") with their own scratch let upload and stage A of a minibatch overlap stage B
// of those before: all of them are chains of small latency-bound operations.
//
// Nothing on the host waits for the device; sizes that depend on the data never leave it.
#include <cmath>
#include <cstdlib>
#include <new>

#include "common.cuh"
#include "../../include/psb200.h"

namespace psb {

struct SgdRunner {
  psb_sgd_runner_config c;
  cudaEvent_t ready[PSB_SGD_MAX_RING];
  cudaEvent_t done[PSB_SGD_MAX_RING];
  bool used[PSB_SGD_MAX_RING];
  uint64_t next = 0;  // minibatches submitted so far (ring position)
  int prev_slot = -1; // ring set of the minibatch submitted last
  int n_events = 0;
  // PSB_SGD_TIMING=1: device time stamps of the stages of the first kTimedSteps steps
  // (copy stream: start, after H2D, after localise; step stream: start, after worker, after server)
  static constexpr int kTimedSteps = 4096;
  cudaEvent_t* stamp = nullptr;  // [kTimedSteps][6]
  int timed = 0;
};

}  // namespace psb

using namespace psb;

PSB_API void psb_sgd_runner_destroy(void* runner) {
  SgdRunner* r = static_cast<SgdRunner*>(runner);
  if (!r) return;
  for (int i = 0; i < r->n_events; ++i) {
    cudaEventDestroy(r->ready[i]);
    cudaEventDestroy(r->done[i]);
  }
  if (r->stamp) {
    for (int i = 0; i < SgdRunner::kTimedSteps * 6; ++i) cudaEventDestroy(r->stamp[i]);
    delete[] r->stamp;
  }
  delete r;
}

PSB_API void* psb_sgd_runner_create(const psb_sgd_runner_config* cfg) {
  if (!cfg || cfg->n_sets < 1 || cfg->n_sets > PSB_SGD_MAX_RING || cfg->max_rows < 1 || cfg->max_nnz < 1 ||
      !cfg->worker_bar || !cfg->overflow_dev || cfg->n_lanes < 1 || cfg->n_lanes > PSB_SGD_MAX_LANES ||
      cfg->n_sets <= cfg->n_lanes) {
    set_error("psb_sgd_runner_create: bad configuration");
    return nullptr;
  }
  for (int l = 0; l < cfg->n_lanes; ++l)
    if (!cfg->scratch[l].temp || !cfg->scratch[l].sorted || !cfg->copy_stream[l]) {
      set_error("psb_sgd_runner_create: lane %d has no scratch or no stream", l);
      return nullptr;
    }
  SgdRunner* r = new (std::nothrow) SgdRunner();
  if (!r) return nullptr;
  r->c = *cfg;
  for (int i = 0; i < cfg->n_sets; ++i) {
    r->used[i] = false;
    if (cudaEventCreateWithFlags(&r->ready[i], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&r->done[i], cudaEventDisableTiming) != cudaSuccess) {
      set_error("psb_sgd_runner_create: cudaEventCreate failed");
      psb_sgd_runner_destroy(r);
      return nullptr;
    }
    r->n_events = i + 1;
  }
  const char* t = getenv("PSB_SGD_TIMING");
  if (t && *t && *t != '0') {
    r->stamp = new cudaEvent_t[SgdRunner::kTimedSteps * 6];
    for (int i = 0; i < SgdRunner::kTimedSteps * 6; ++i) cudaEventCreate(&r->stamp[i]);
  }
  return r;
}

// Mean milliseconds per step of {H2D, localise, worker kernel, server kernel, copy-stream stage end to
// end, step-stream stage end to end, whole step period} over the timed steps (skipping the first
// tenth); synchronises the device.  Returns the number of steps averaged (0: timing is off).
PSB_API int psb_sgd_runner_timing(void* runner, double* out7) {
  SgdRunner* r = static_cast<SgdRunner*>(runner);
  if (!r || !r->stamp || r->timed < 20) return 0;
  cudaDeviceSynchronize();
  const int lo = r->timed / 10, hi = r->timed;
  double acc[7] = {0, 0, 0, 0, 0, 0, 0};
  for (int i = lo; i < hi; ++i) {
    cudaEvent_t* e = r->stamp + (size_t)i * 6;
    float ms;
    cudaEventElapsedTime(&ms, e[0], e[1]); acc[0] += ms;
    cudaEventElapsedTime(&ms, e[1], e[2]); acc[1] += ms;
    cudaEventElapsedTime(&ms, e[3], e[4]); acc[2] += ms;
    cudaEventElapsedTime(&ms, e[4], e[5]); acc[3] += ms;
    cudaEventElapsedTime(&ms, e[0], e[2]); acc[4] += ms;
    cudaEventElapsedTime(&ms, e[3], e[5]); acc[5] += ms;
  }
  float span;
  cudaEventElapsedTime(&span, r->stamp[(size_t)lo * 6 + 5], r->stamp[(size_t)(hi - 1) * 6 + 5]);
  for (int k = 0; k < 6; ++k) out7[k] = acc[k] / (hi - lo);
  out7[6] = span / (hi - lo - 1);
  return hi - lo;
}

PSB_API int psb_sgd_runner_submit(void* runner, const psb_sgd_host_minibatch* mb, int32_t n, uint64_t seq_first,
                                  void* step_stream) {
  SgdRunner* r = static_cast<SgdRunner*>(runner);
  PSB_REQUIRE(r && (mb || n == 0) && n >= 0 && seq_first >= 1, "psb_sgd_runner_submit: bad argument");
  const psb_sgd_runner_config& c = r->c;
  cudaStream_t step = (cudaStream_t)step_stream;
  cudaStream_t copy = nullptr;
  const int W = c.worker.n_servers;
  for (int i = 0; i < n; ++i) {
    const psb_sgd_host_minibatch& h = mb[i];
    PSB_REQUIRE(h.n_rows >= 0 && h.n_rows <= c.max_rows && h.nnz >= 0 && h.nnz <= c.max_nnz,
                "psb_sgd_runner_submit: minibatch of %lld rows / %lld non-zeros exceeds the ring (%lld / %lld)",
                (long long)h.n_rows, (long long)h.nnz, (long long)c.max_rows, (long long)c.max_nnz);
    const int slot = (int)(r->next % (uint64_t)c.n_sets);
    const psb_sgd_ring_set& s = c.set[slot];
    const uint64_t seq = seq_first + (uint64_t)i;
    const int lane = (int)(r->next % (uint64_t)c.n_lanes);
    copy = (cudaStream_t)c.copy_stream[lane];
    const psb_sgd_scratch& sc = c.scratch[lane];

    // ---- copy stream: upload and localise ----
    if (r->used[slot]) PSB_CUDA(cudaStreamWaitEvent(copy, r->done[slot], 0));
    cudaEvent_t* stamp = (r->stamp && r->timed < SgdRunner::kTimedSteps) ? r->stamp + (size_t)r->timed++ * 6 : nullptr;
    if (stamp) cudaEventRecord(stamp[0], copy);
    if (h.n_rows > 0) {
      PSB_CUDA(cudaMemcpyAsync(s.row_ptr, h.row_ptr, (size_t)(h.n_rows + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, copy));
      PSB_CUDA(cudaMemcpyAsync(s.y, h.y, (size_t)h.n_rows * sizeof(float), cudaMemcpyHostToDevice, copy));
    }
    if (h.nnz > 0) {
      PSB_CUDA(cudaMemcpyAsync(s.keys, h.keys, (size_t)h.nnz * sizeof(uint64_t), cudaMemcpyHostToDevice, copy));
      if (h.vals) {
        PSB_REQUIRE(s.vals, "psb_sgd_runner_submit: minibatch has values, the ring was built for binary features");
        PSB_CUDA(cudaMemcpyAsync(s.vals, h.vals, (size_t)h.nnz * sizeof(float), cudaMemcpyHostToDevice, copy));
      }
    }
    if (stamp) cudaEventRecord(stamp[1], copy);
    int rc = psb_sgd_localize_sort(s.keys, h.nnz, sc.sorted, sc.uniq, sc.cnt, s.counts, sc.temp, c.temp_bytes, copy);
    if (rc) return rc;
    if (r->prev_slot >= 0) PSB_CUDA(cudaStreamWaitEvent(copy, r->ready[r->prev_slot], 0));
    rc = psb_sgd_localize_finish(s.keys, h.nnz, sc.uniq, sc.cnt, sc.flag, s.cols, s.col, s.counts, c.sketch, c.cm_n,
                                 c.cm_k, c.freq, c.bounds_dev, W, c.worker.land_cap, s.owner_begin, c.table_keys,
                                 c.table_cap, s.col_owner, s.col_slot, c.overflow_dev, sc.temp, c.temp_bytes, copy);
    if (rc) return rc;
    if (stamp) cudaEventRecord(stamp[2], copy);
    PSB_CUDA(cudaEventRecord(r->ready[slot], copy));

    // ---- step stream: worker, then this rank's server ----
    PSB_CUDA(cudaStreamWaitEvent(step, r->ready[slot], 0));
    if (stamp) cudaEventRecord(stamp[3], step);
    psb_sgd_worker wk = c.worker;
    wk.row_ptr = s.row_ptr; wk.col = s.col; wk.val = h.vals ? s.vals : nullptr; wk.y = s.y;
    wk.n_rows = h.n_rows; wk.n_cols = h.nnz; wk.n_cols_dev = s.counts + 1;
    wk.col_owner = s.col_owner; wk.col_slot = s.col_slot; wk.owner_begin = s.owner_begin;
    int64_t work = h.n_rows > h.nnz / 8 ? h.n_rows : h.nnz / 8;
    int grid = (int)((work + 7) / 8);
    grid = grid < 1 ? 1 : (grid > c.grid_worker ? c.grid_worker : grid);
    rc = psb_sgd_worker_step(&wk, seq, grid, c.worker_bar, step);
    if (rc) return rc;
    if (stamp) cudaEventRecord(stamp[4], step);
    psb_sgd_server sv = c.server;
    sv.sgd_t = sqrtf((float)seq);
    rc = c.fused_server ? psb_sgd_server_update_fused(&sv, seq, c.grid_server, c.server_bar, step)
                        : psb_sgd_server_update(&sv, seq, c.grid_server, step);
    if (rc) return rc;
    if (stamp) cudaEventRecord(stamp[5], step);
    PSB_CUDA(cudaEventRecord(r->done[slot], step));
    r->used[slot] = true;
    r->prev_slot = slot;
    ++r->next;
  }
  if (c.overflow_host && copy)  // `copy` ran stage B of the last minibatch, which waited for all earlier ones
    PSB_CUDA(cudaMemcpyAsync(c.overflow_host, c.overflow_dev, sizeof(int32_t), cudaMemcpyDeviceToHost, copy));
  return 0;
}
