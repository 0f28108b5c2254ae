// Darlin (L1-regularised logistic regression by block coordinate descent) on sm_100a.
//
// One process per GPU hosts a worker (rows of X, CSC over its local columns) and a server
// shard (a key range of w).  The three phases of a feature-block step are
//
//   grad_push      K1 + C1  worker: g_j,u_j per column -> stores into the owner shard's
//                           per-worker landing slot over NVLink peer memory, then a
//                           release.sys flag                      (ref darlin.h:344-427)
//   server_update  K2 + K3  server: acquire all workers' flags, reduce the slots in fixed
//                           worker order, prox-Newton + trust region + KKT filter, publish
//                           w / dw, release version flags         (ref darlin.h:175-246)
//   pull_update    K4 + C2  worker: acquire the owners' version flags, load w / dw of its
//                           columns from the owner shard in peer memory, update the
//                           margins z_i += dw_j * y_i x_ij        (ref darlin.h:317-342,429-502)
//
// Differences from the reference that are deliberate:
//   * the worker keeps the margin z_i = y_i<x_i,w> (= log dual_i) instead of dual_i, so the
//     column scatter is an fp64 atomicAdd instead of a multiply that would race between
//     columns of one block (the reference avoids the race by row-partitioning threads);
//   * values are pre-multiplied by the label (sval = y_i x_ij), removing the y gather;
//   * nothing is sent for inactive columns (the server's active set is identical, so the
//     NaN the reference tunnels through G/U is redundant); w itself still carries the NaN
//     mark so the worker learns about newly filtered coordinates exactly as in the reference;
//   * the server publishes dw = w_new - w_old next to w, so all segments of a long column
//     can be processed by independent warps without reading a half-updated cached w.
#include "common.cuh"
#include "../../include/psb200.h"
#include <mutex>

namespace psb {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxBps = 32;  // max feature blocks merged into one step (tau+1)

struct StepPlan {
  int64_t first[kMaxBps];   // first segment (worker) / first key (server) of block q
  int64_t pref[kMaxBps + 1];
  uint32_t mask;
};

__device__ __forceinline__ int locate(const StepPlan& p, int nb, int64_t v) {
  int q = 0;
  while (q + 1 < nb && v >= p.pref[q + 1]) ++q;
  return q;
}

__device__ __forceinline__ void plan_worker(StepPlan& p, const psb_darlin_worker& W,
                                            const int32_t* order, int step, int bps) {
  int64_t pref = 0;
  uint32_t mask = 0;
  for (int q = 0; q < bps; ++q) {
    int b = order[(int64_t)step * bps + q];
    int64_t lo = 0, n = 0;
    if (b >= 0) {
      int64_t cl = W.blk_col[2 * b], ch = W.blk_col[2 * b + 1];
      lo = W.col_seg[cl];
      n = W.col_seg[ch] - lo;
      mask |= W.blk_srv_mask[b];
    }
    p.first[q] = lo;
    p.pref[q] = pref;
    pref += n;
  }
  p.pref[bps] = pref;
  p.mask = mask;
}

// -------------------------------------------------------------------------------------------
// K1 + push
// -------------------------------------------------------------------------------------------
__device__ __forceinline__ void push_column(const psb_darlin_worker& W, int64_t j, double g,
                                            double u) {
  int s = W.col_owner[j];
  int64_t at = (int64_t)W.rank * W.land_cap[s] + W.col_off[j];
  W.land_g[s][at] = g;  // peer store when s != rank
  W.land_u[s][at] = u;
}

template <bool BINARY>
__device__ __forceinline__ void grad_phase(const psb_darlin_worker& W, const StepPlan& plan, int bps,
                                           int64_t gw, int64_t nw, int lane) {
  const int64_t total = plan.pref[bps];
  for (int64_t v = gw; v < total; v += nw) {
    const int q = locate(plan, bps, v);
    const int64_t s = plan.first[q] + (v - plan.pref[q]);
    const int64_t j = W.seg_col[s];
    if (!bit_test_cg(W.active, j)) continue;
    const int64_t cs = W.col_seg[j];
    const int nseg = (int)(W.col_seg[j + 1] - cs);
    const int64_t beg = W.seg_beg[s];
    const int64_t end = min(beg + (int64_t)W.seg_len, W.col_ptr[j + 1]);
    double d = __ldcg(W.delta + j);
    if (BINARY) d = exp(d);
    double g = 0, u = 0;
#pragma unroll 2
    for (int64_t e = beg + lane; e < end; e += 32) {
      const uint32_t i = __ldg(W.row_idx + e);
      const double zi = __ldcg(W.z + i);
      // tau = 1/(1+dual) with dual = exp(z)
      const double tau = 1.0 / (1.0 + exp(zi));
      if (BINARY) {
        const double yv = (double)__ldg(W.y + i);
        g -= yv * tau;
        u += fmin(tau * (1.0 - tau) * d, 0.25);
      } else {
        const double sv = (double)__ldg(W.sval + e);
        g -= tau * sv;
        u += fmin(tau * (1.0 - tau) * exp(fabs(sv) * d), 0.25) * sv * sv;
      }
    }
    g = warp_sum(g);
    u = warp_sum(u);
    if (nseg == 1) {
      if (lane == 0) push_column(W, j, g, u);
      continue;
    }
    // long column: park the partial, the last arriving segment reduces in segment order
    int old = 0;
    if (lane == 0) {
      __stcg(W.part + 2 * v, g);
      __stcg(W.part + 2 * v + 1, u);
      __threadfence();
      old = atomicAdd(W.col_cnt + j, 1);
    }
    old = __shfl_sync(FULL, old, 0);
    if (old == nseg - 1) {
      __threadfence();
      const int64_t v0 = v - (s - cs);
      double sg = 0, su = 0;
      for (int k = lane; k < nseg; k += 32) {
        sg += __ldcg(W.part + 2 * (v0 + k));
        su += __ldcg(W.part + 2 * (v0 + k) + 1);
      }
      sg = warp_sum(sg);
      su = warp_sum(su);
      if (lane == 0) {
        push_column(W, j, sg, su);
        W.col_cnt[j] = 0;
      }
    }
  }
}

template <bool BINARY>
__global__ void __launch_bounds__(kThreads)
k_grad_push(const __grid_constant__ psb_darlin_worker W, const psb_darlin_ctrl* __restrict__ ctrl,
            const int32_t* __restrict__ order, int step, int bps) {
  __shared__ StepPlan plan;
  if (threadIdx.x == 0) plan_worker(plan, W, order, step, bps);
  __syncthreads();
  grad_phase<BINARY>(W, plan, bps, (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5),
                     (int64_t)gridDim.x * kWarps, threadIdx.x & 31);
  // all of this CTA's peer stores are ordered before its arrival; the last CTA raises flags
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    unsigned old = atomicAdd(W.done_ctr, 1u);
    if (old == gridDim.x - 1) {
      *W.done_ctr = 0;
      __threadfence_system();
      const uint64_t seq = ctrl->seq_base + (uint64_t)step + 1;
      for (int sv = 0; sv < W.n_servers; ++sv)
        if (plan.mask >> sv & 1u) st_release_sys(W.push_flag[sv] + W.rank, seq);
    }
  }
}

// -------------------------------------------------------------------------------------------
// K2 + K3: server
// -------------------------------------------------------------------------------------------
struct ServerHyper {
  double lambda, eta, delta_max, thr;
};

// thread-strided over the step's coordinates; returns this thread's max violation
__device__ __forceinline__ double server_phase(const psb_darlin_server& S, const ServerHyper& h,
                                               const StepPlan& plan, int bps, int64_t tid,
                                               int64_t nthreads) {
  const int64_t total = plan.pref[bps];
  double vio_max = 0;
  for (int64_t v = tid; v < total; v += nthreads) {
    const int q = locate(plan, bps, v);
    const int64_t k = plan.first[q] + (v - plan.pref[q]);
    if (!bit_test_cg(S.active, k)) continue;
    double w = __ldcg(S.w + k);
    // revived by reset_kkt_filter but still NaN: arithmetic on NaN keeps it NaN in the
    // reference, the bit stays set there too, so only skip
    if (is_marked(w)) continue;
    double g = 0, u = 0;
    for (int r = 0; r < S.n_workers; ++r) {  // fixed order => run-to-run deterministic
      const int64_t at = (int64_t)r * S.land_cap + k;
      g += __ldcg(S.land_g + at);  // slots of workers that lack the column stay 0 forever
      u += __ldcg(S.land_u + at);
    }
    u = u / h.eta + 1e-10;
    const double g_pos = g + h.lambda, g_neg = g - h.lambda;
    double d = -w, vio = 0;
    if (w == 0) {
      if (g_pos < 0) {
        vio = -g_pos;
      } else if (g_neg > 0) {
        vio = g_neg;
      } else if (g_pos > h.thr && g_neg < -h.thr) {
        bit_clear_atomic(S.active, k);
        S.w[k] = nan_mark();
        S.dw[k] = 0;
        continue;
      }
    }
    vio_max = fmax(vio_max, vio);
    if (g_pos <= u * w) {
      d = -g_pos / u;
    } else if (g_neg >= u * w) {
      d = -g_neg / u;
    }
    const double dl = __ldcg(S.delta + k);
    d = fmin(dl, fmax(-dl, d));
    S.delta[k] = fmin(h.delta_max, 2 * fabs(d) + .1);
    const double w_new = w + d;
    S.w[k] = w_new;
    S.dw[k] = w_new - w;  // what the reference worker computes as new_w - cur_w
  }
  return vio_max;
}

__device__ __forceinline__ void plan_server(StepPlan& p, const psb_darlin_server& S,
                                            const int32_t* order, int step, int bps) {
  int64_t pref = 0;
  for (int q = 0; q < bps; ++q) {
    int b = order[(int64_t)step * bps + q];
    int64_t lo = 0, n = 0;
    if (b >= 0) {
      lo = S.blk_rng[2 * b];
      n = S.blk_rng[2 * b + 1] - lo;
    }
    p.first[q] = lo;
    p.pref[q] = pref;
    pref += n;
  }
  p.pref[bps] = pref;
}

__global__ void __launch_bounds__(kThreads)
k_server_update(const __grid_constant__ psb_darlin_server S, const psb_darlin_ctrl* __restrict__ ctrl,
                const int32_t* __restrict__ order, int step, int bps) {
  __shared__ StepPlan plan;
  __shared__ double s_vio[kWarps];
  const uint64_t seq = ctrl->seq_base + (uint64_t)step + 1;
  if (threadIdx.x == 0) plan_server(plan, S, order, step, bps);
  __syncthreads();
  const int64_t total = plan.pref[bps];
  if (total == 0) return;  // "none of my business" (ref darlin.h:188-189): nobody waits on us
  if (threadIdx.x < S.n_workers) spin_until_ge(S.push_flag + threadIdx.x, seq);
  __syncthreads();
  const ServerHyper h{ctrl->lambda, ctrl->eta, ctrl->delta_max, ctrl->kkt_threshold};
  double vio_max = server_phase(S, h, plan, bps, (int64_t)blockIdx.x * kThreads + threadIdx.x,
                                (int64_t)gridDim.x * kThreads);
  vio_max = warp_max(vio_max);
  if ((threadIdx.x & 31) == 0) s_vio[threadIdx.x >> 5] = vio_max;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < kWarps; ++i) vio_max = fmax(vio_max, s_vio[i]);
    if (vio_max > 0) atomic_max_nonneg(S.violation, vio_max);
    __threadfence_system();
    unsigned old = atomicAdd(S.done_ctr, 1u);
    if (old == gridDim.x - 1) {
      *S.done_ctr = 0;
      __threadfence_system();
      for (int r = 0; r < S.n_workers; ++r) st_release_sys(S.ver_flag[r], seq);
    }
  }
}

// -------------------------------------------------------------------------------------------
// K4 + pull
// -------------------------------------------------------------------------------------------
template <bool BINARY>
__device__ __forceinline__ void pull_phase(const psb_darlin_worker& W, const StepPlan& plan, int bps,
                                           double delta_max, int64_t gw, int64_t nw, int lane) {
  const int64_t total = plan.pref[bps];
  for (int64_t v = gw; v < total; v += nw) {
    const int q = locate(plan, bps, v);
    const int64_t s = plan.first[q] + (v - plan.pref[q]);
    const int64_t j = W.seg_col[s];
    double w_new = 0, d = 0;
    if (lane == 0) {
      const int o = W.col_owner[j];
      const int64_t off = W.col_off[j];
      w_new = ld_volatile_f64(W.srv_w[o] + off);  // peer load when o != rank
      d = ld_volatile_f64(W.srv_dw[o] + off);
    }
    w_new = __shfl_sync(FULL, w_new, 0);
    d = __shfl_sync(FULL, d, 0);
    const bool first = (s == W.col_seg[j]);
    if (is_marked(w_new)) {  // the server filtered this coordinate (ref darlin.h:439-444)
      if (first && lane == 0) {
        bit_clear_atomic(W.active, j);
        W.w[j] = 0;
      }
      continue;
    }
    if (!bit_test_cg(W.active, j)) continue;
    if (first && lane == 0) {
      W.w[j] = w_new;
      W.delta[j] = fmin(delta_max, 2 * fabs(d) + .1);
    }
    if (d == 0) continue;
    const int64_t beg = W.seg_beg[s];
    const int64_t end = min(beg + (int64_t)W.seg_len, W.col_ptr[j + 1]);
#pragma unroll 2
    for (int64_t e = beg + lane; e < end; e += 32) {
      const uint32_t i = __ldg(W.row_idx + e);
      const double sv = BINARY ? (double)__ldg(W.y + i) : (double)__ldg(W.sval + e);
      atomicAdd(W.z + i, d * sv);  // dual_i *= exp(y_i d x_ij)  <=>  z_i += d * sval
    }
  }
}

template <bool BINARY>
__global__ void __launch_bounds__(kThreads)
k_pull_update(const __grid_constant__ psb_darlin_worker W, const psb_darlin_ctrl* __restrict__ ctrl,
              const int32_t* __restrict__ order, int step, int bps) {
  __shared__ StepPlan plan;
  if (threadIdx.x == 0) plan_worker(plan, W, order, step, bps);
  __syncthreads();
  const int64_t total = plan.pref[bps];
  if (total == 0) return;
  const uint64_t seq = ctrl->seq_base + (uint64_t)step + 1;
  if (threadIdx.x < W.n_servers && (plan.mask >> threadIdx.x & 1u))
    spin_until_ge(W.ver_flag + threadIdx.x, seq);
  __syncthreads();
  pull_phase<BINARY>(W, plan, bps, ctrl->delta_max, (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5),
                     (int64_t)gridDim.x * kWarps, threadIdx.x & 31);
}

// -------------------------------------------------------------------------------------------
// pass begin, K5 evaluate
// -------------------------------------------------------------------------------------------
__global__ void k_pass_begin(uint32_t* wk_active, int64_t wk_cols, uint32_t* sv_active,
                             int64_t sv_keys, double* sv_w, double* violation,
                             const psb_darlin_ctrl* __restrict__ ctrl) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if (t == 0 && violation) *violation = 0;
  if (!ctrl->reset_kkt_filter) return;
  // fill(true) but keep the padding bits of the last word clear so popcounts stay exact
  for (int64_t i = t; i < (wk_cols + 31) / 32; i += stride) {
    int64_t rem = wk_cols - i * 32;
    wk_active[i] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
  }
  for (int64_t i = t; i < (sv_keys + 31) / 32; i += stride) {
    int64_t rem = sv_keys - i * 32;
    sv_active[i] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
  }
  if (ctrl->revive_on_reset && sv_w)
    for (int64_t i = t; i < sv_keys; i += stride)
      if (is_marked(sv_w[i])) sv_w[i] = 0;
}

__global__ void __launch_bounds__(kThreads)
k_evaluate(const double* __restrict__ z, int64_t n_rows, const double* __restrict__ w,
           const uint32_t* __restrict__ active, int64_t n_keys, const double* violation,
           const psb_darlin_ctrl* __restrict__ ctrl, psb_darlin_progress* out_local,
           psb_darlin_progress* const* out_peer, int n_peers, uint64_t seq, double* scratch) {
  __shared__ double sh[4][kWarps];
  const int64_t t = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  double loss = 0, l1 = 0;
  unsigned long long nnz = 0, nact = 0;
  for (int64_t i = t; i < n_rows; i += stride) {
    const double zi = z[i];
    // log(1 + 1/dual) = log(1+exp(-z)), evaluated without overflow
    loss += zi > 0 ? log1p(exp(-zi)) : (-zi + log1p(exp(zi)));
  }
  for (int64_t i = t; i < n_keys; i += stride) {
    const double wi = w[i];
    if (!is_marked(wi) && wi != 0) {
      ++nnz;
      l1 += fabs(wi);
    }
  }
  for (int64_t i = t; i < (n_keys + 31) / 32; i += stride) nact += __popc(active[i]);
  loss = warp_sum(loss);
  l1 = warp_sum(l1);
  double dn = warp_sum((double)nnz), da = warp_sum((double)nact);
  const int wid = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) {
    sh[0][wid] = loss; sh[1][wid] = l1; sh[2][wid] = dn; sh[3][wid] = da;
  }
  __syncthreads();
  uint32_t* ctr = reinterpret_cast<uint32_t*>(scratch + 4 * (int64_t)gridDim.x);
  __shared__ bool last;
  if (threadIdx.x == 0) {
    for (int c = 0; c < 4; ++c) {
      double a = 0;
      for (int i = 0; i < kWarps; ++i) a += sh[c][i];
      __stcg(scratch + 4 * (int64_t)blockIdx.x + c, a);
    }
    __threadfence();
    last = atomicAdd(ctr, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  // fixed-order final reduction by one warp-strided pass of the last CTA
  double acc[4] = {0, 0, 0, 0};
  for (int b = threadIdx.x; b < gridDim.x; b += kThreads)
    for (int c = 0; c < 4; ++c) acc[c] += __ldcg(scratch + 4 * (int64_t)b + c);
  for (int c = 0; c < 4; ++c) acc[c] = warp_sum(acc[c]);
  if ((threadIdx.x & 31) == 0)
    for (int c = 0; c < 4; ++c) sh[c][wid] = acc[c];
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int c = 0; c < 4; ++c) {
      double a = 0;
      for (int i = 0; i < kWarps; ++i) a += sh[c][i];
      acc[c] = a;
    }
    *ctr = 0;
    psb_darlin_progress p;
    p.loss = acc[0];
    p.penalty = acc[1] * ctrl->lambda;
    p.nnz_w = (uint64_t)(acc[2] + 0.5);
    p.nnz_active = (uint64_t)(acc[3] + 0.5);
    p.violation = violation ? *violation : 0;
    const uint64_t stamp = seq ? seq : ctrl->seq_base + (uint64_t)ctrl->n_steps;
    p.seq = 0;
    for (int r = -1; r < n_peers; ++r) {
      psb_darlin_progress* o = r < 0 ? out_local : out_peer[r];
      if (!o) continue;
      o->loss = p.loss; o->penalty = p.penalty; o->violation = p.violation;
      o->nnz_w = p.nnz_w; o->nnz_active = p.nnz_active;
    }
    __threadfence_system();
    for (int r = -1; r < n_peers; ++r) {
      psb_darlin_progress* o = r < 0 ? out_local : out_peer[r];
      if (o) st_release_sys(&o->seq, stamp);
    }
  }
}


// -------------------------------------------------------------------------------------------
// Persistent whole-pass kernel (cooperative launch, every CTA resident)
//
// One launch runs pass_begin, all steps and the progress evaluation.  The three phases of a
// step are separated by split grid barriers (arrive / wait on one L2 counter) instead of
// kernel boundaries; the next step's plan is fetched between arrive and wait.  Cross-GPU
// ordering is unchanged: after barrier A one thread raises the push flags on the remote
// owners, the server CTAs acquire the remote workers' flags before reducing, after barrier B
// one thread raises the version flags, and the pull phase acquires the remote owners' flags.
// With one rank no system-scope operation is executed at all.
//
// blk_tab[b] = {first segment, #segments, first shard key, #shard keys, server mask, 0}
// -------------------------------------------------------------------------------------------
struct FusedPlan {
  StepPlan wk;   // segments
  StepPlan sv;   // shard coordinates
};

__device__ __forceinline__ void fused_plan(FusedPlan& P, const int64_t* __restrict__ blk_tab,
                                           const int32_t* __restrict__ order, int step, int bps) {
  // executed by warp 0: lane q fetches block q of the step, then a warp scan builds prefixes
  const int lane = threadIdx.x & 31;
  int64_t seg_lo = 0, nseg = 0, key_lo = 0, nkey = 0;
  uint32_t mask = 0;
  if (lane < bps) {
    const int b = __ldcg(order + (int64_t)step * bps + lane);
    if (b >= 0) {
      const int64_t* t = blk_tab + 6 * (int64_t)b;
      seg_lo = __ldg(t); nseg = __ldg(t + 1); key_lo = __ldg(t + 2); nkey = __ldg(t + 3);
      mask = (uint32_t)__ldg(t + 4);
    }
  }
  int64_t ps = nseg, pk = nkey;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int64_t a = __shfl_up_sync(FULL, ps, o), c = __shfl_up_sync(FULL, pk, o);
    if (lane >= o) { ps += a; pk += c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mask |= __shfl_xor_sync(FULL, mask, o);
  if (lane < bps) {
    P.wk.first[lane] = seg_lo; P.wk.pref[lane] = ps - nseg;
    P.sv.first[lane] = key_lo; P.sv.pref[lane] = pk - nkey;
  }
  if (lane == bps - 1) { P.wk.pref[bps] = ps; P.sv.pref[bps] = pk; }
  if (lane == 0) { P.wk.mask = mask; P.sv.mask = mask; }
}

// ---- packed-metadata phases used by the persistent kernel -------------------------------------
struct SegMeta {
  int64_t j, beg;
  int len;
  bool first, single;
};
__device__ __forceinline__ SegMeta decode_meta(const int4 m) {
  SegMeta r;
  r.j = m.x;
  r.len = m.y & 0xffff;
  r.first = (m.y >> 16) & 1;
  r.single = (m.y >> 17) & 1;
  r.beg = (int64_t)(uint32_t)m.z | ((int64_t)m.w << 32);
  return r;
}
__device__ __forceinline__ SegMeta load_meta(const void* seg_meta, int64_t s) {
  return decode_meta(__ldg(reinterpret_cast<const int4*>(seg_meta) + s));
}

__device__ __forceinline__ uint32_t atom_add_acqrel_gpu(uint32_t* p, uint32_t v) {
  uint32_t old;
  asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

constexpr int kUnroll = 4;  // 4 x 32 lanes = one 128-entry segment per batch

// ---- fused multi-rank pass: one NVLink flight per step, replicated coordinate update ----------
// Every rank keeps a replica of (w, delta, active bit) of every column of the global dictionary.
// A worker's push of column j is one tagged 32-byte entry {g, u} stored into the landing table of
// every rank that needs it (the ranks that hold nonzeros of j and the owner of j): lane q of the
// pushing warp stores to rank q, so the N peer stores leave together.  In phase 3 every warp
// that works on a segment of column j polls, in its OWN memory, the entries of the workers
// that hold j (lane r polls worker r), sums them in worker order and applies the proximal update
// itself.  All ranks run the same instructions on the same bits, so the replicas stay identical
// (ref darlin.h:206-246 is evaluated N times instead of once plus a broadcast); the owner's copy
// is the authoritative shard that is evaluated and saved.  Compared with owner-computes this
// removes the owner's serial shard phase and the second NVLink flight (owner -> holders) from
// every step; it costs N-1 more 32-byte stores per column and step.
//
// An entry is {g.lo, tag, g.hi, tag | u.lo, tag, u.hi, tag}: every 8-byte half carries the low
// word of the step's sequence number, so a reader that finds the four tags equal to the step
// it is in holds a complete, fresh (g, u) - no flag, no fence, no reset.  Only 8-byte
// single-copy atomicity of peer stores is assumed.  Entries are double-buffered by step parity
// and a writer starts step s only after every peer finished step s-2 (W.done_local, stored by
// the peers after their last barrier of a step), so an unread entry is never overwritten.
constexpr uint32_t kSpinLimit = 1u << 26;  // ~10 s of polling: trap instead of hanging the GPU

__device__ __forceinline__ void entry_store(void* entry, double a, double b, uint32_t tag) {
  const unsigned long long x = (unsigned long long)__double_as_longlong(a);
  const unsigned long long y = (unsigned long long)__double_as_longlong(b);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %2};" ::"l"(entry), "r"((uint32_t)x),
               "r"(tag), "r"((uint32_t)(x >> 32)) : "memory");
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %2};" ::"l"((char*)entry + 16),
               "r"((uint32_t)y), "r"(tag), "r"((uint32_t)(y >> 32)) : "memory");
}
__device__ __forceinline__ void entry_poll(const void* entry, uint32_t tag, double& a, double& b) {
  uint32_t a0, a1, a2, a3, b0, b1, b2, b3, spins = 0;
  for (;;) {
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3) : "l"(entry) : "memory");
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3) : "l"((const char*)entry + 16) : "memory");
    if (a1 == tag && a3 == tag && b1 == tag && b3 == tag) break;
    if (++spins > kSpinLimit) __trap();
  }
  a = __longlong_as_double((long long)((unsigned long long)a0 | ((unsigned long long)a2 << 32)));
  b = __longlong_as_double((long long)((unsigned long long)b0 | ((unsigned long long)b2 << 32)));
}
// The lossy form of the push (the FIXING_FLOAT idea, ref src/filter/fixing_float.h:44-95, applied
// where the bytes are: the gradient push): (g, u) rounded to fp32 travel as ONE 16-byte store
// {g32, tag, u32, tag}.  Every rank - the sender included - reads the rounded bits back from an
// entry, so the replicas stay identical.  Inactive columns are simply not pushed, which is
// what the reference's NaN marks in G / U say with more bytes.
__device__ __forceinline__ void entry_store_f32(void* entry, float a, float b, uint32_t tag) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %2};" ::"l"(entry), "r"(__float_as_uint(a)),
               "r"(tag), "r"(__float_as_uint(b)) : "memory");
}
__device__ __forceinline__ void entry_poll_f32(const void* entry, uint32_t tag, double& a, double& b) {
  uint32_t a0, a1, a2, a3, spins = 0;
  for (;;) {
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3) : "l"(entry) : "memory");
    if (a1 == tag && a3 == tag) break;
    if (++spins > kSpinLimit) __trap();
  }
  a = (double)__uint_as_float(a0);
  b = (double)__uint_as_float(a2);
}
__device__ __forceinline__ int64_t entry_index(const psb_darlin_worker& W, uint32_t tag, int from,
                                               int64_t j) {
  return (((int64_t)(tag & 1u) * W.n_workers + from) * W.n_cols + j) * 32;
}

__device__ __forceinline__ void spin_flag_ge(const uint64_t* flag, uint64_t want) {
  uint32_t spins = 0;
  while (ld_acquire_sys(flag) < want) {
    __nanosleep(20);
    if (++spins > kSpinLimit) __trap();
  }
}

// Everything about one segment that does not change during a step, held in registers.  The
// persistent kernel stages the first segment of every warp for step s+1 while it waits on the
// last barrier of step s, and keeps it from phase 1 to phase 3 of the step: what remains on the
// critical path after a barrier is one level of loads of mutable state (margins, trust region,
// active bit, weights), the arithmetic and the stores.
struct SegRegs {
  int64_t v;            // step-local segment number this was staged for (-1: nothing staged)
  int64_t j, beg, off;  // column, first entry, coordinate offset inside the owner's shard
  int len, nseg, sc, owner;  // entries; segments of the column; my position among them; owner rank
  bool first, single;
  uint32_t has;         // ranks that hold nonzeros of the column (multi-rank pass)
  uint32_t idx[kUnroll];
  float sv[kUnroll];
};
// the coordinate's state as phase 1 saw it, carried to phase 3 (multi-rank pass): phase 3 of the
// column's first segment overwrites the state while the other segments still need the old one
struct Carry {
  double w, dl;
  bool act;
};

template <bool BINARY, bool SINGLE>
__device__ __forceinline__ void stage_segment(SegRegs& R, const psb_darlin_worker& W,
                                              const StepPlan& plan, int bps, int64_t v, int lane,
                                              const int4 fetched = make_int4(0, -1, 0, 0)) {
  R.v = v;
  const int q = locate(plan, bps, v);
  const int64_t s = plan.first[q] + (v - plan.pref[q]);
  // the descriptor may have been fetched one barrier earlier (prefetch_step): one round trip less
  const SegMeta m = fetched.y >= 0 ? decode_meta(fetched) : load_meta(W.seg_meta, s);
  R.j = m.j; R.beg = m.beg; R.len = m.len; R.first = m.first; R.single = m.single;
#pragma unroll
  for (int k = 0; k < kUnroll; ++k) {
    const int e = k * 32 + lane;
    R.idx[k] = e < m.len ? __ldg(W.row_idx + m.beg + e) : 0xffffffffu;
  }
  R.owner = __ldg(W.col_owner + m.j);
  R.off = __ldg(W.col_off + m.j);
  R.has = SINGLE ? 1u : __ldg(W.col_has + m.j);
  R.nseg = 1; R.sc = 0;
  if (!SINGLE && !m.single) {  // only the multi-rank reduction needs the position among the column's segments
    const int64_t cs = __ldg(W.col_seg + m.j);
    R.nseg = (int)(__ldg(W.col_seg + m.j + 1) - cs);
    R.sc = (int)(s - cs);
  }
#pragma unroll
  for (int k = 0; k < kUnroll; ++k) {
    const bool ok = R.idx[k] != 0xffffffffu;
    if (BINARY) R.sv[k] = ok ? __ldg(W.y + R.idx[k]) : 0.f;
    else R.sv[k] = ok ? __ldg(W.sval + m.beg + k * 32 + lane) : 0.f;
  }
}

// lane q stores the column's (g, u) into rank q's landing table when rank q holds the column
// or owns it; every lane holds the same g, u
__device__ __forceinline__ void push_entries(const psb_darlin_worker& W, const SegRegs& R, double g,
                                             double u, uint32_t tag, int lane) {
  const uint32_t targets = R.has | (1u << R.owner);
  if (lane < W.n_workers && (targets >> lane & 1u)) {
    char* at = W.land2[lane] + entry_index(W, tag, W.rank, R.j);  // peer memory when lane != rank
    if (W.push_fp32) entry_store_f32(at, (float)g, (float)u, tag);
    else entry_store(at, g, u, tag);
  }
}

// ---- batched fp64 exp / reciprocal ----------------------------------------------------------
// The gradient needs, per nonzero, tau = 1/(1+e^z) and e^{|x| delta} in double precision.  The
// library exp() and the IEEE divide are ~30 dependent fp64 operations each and nvcc does not
// interleave neighbouring calls (measured on B200: 245 cycles per sigmoid at any ILP, DFMA
// latency 8.3 cycles, benchmarks/micro/fp64_rate.cu), which made this arithmetic the longest
// item of a step.  Below the same functions are written stage by stage over the kUnroll
// entries of a batch, so the chains of a batch overlap: table-driven exp (e^x = 2^(k/64) *
// p(r), |r| <= ln2/128, degree-5 polynomial, error < 1.5 ulp) and a Newton reciprocal seeded
// by rcp.approx (error < 1 ulp).
constexpr int kExpTab = 64;
__device__ __forceinline__ void fill_exp_table(double* tab) {  // tab[j] = 2^(j/64), shared memory
  for (int j = threadIdx.x; j < kExpTab; j += blockDim.x) tab[j] = exp2((double)j / kExpTab);
}
template <int N>
__device__ __forceinline__ void exp_batch(double (&x)[N], const double* __restrict__ tab) {
  const double kMagic = 6755399441055744.0;  // 1.5 * 2^52: the integer lands in the low word
  double kd[N], r[N], p[N];
  int ki[N];
#pragma unroll
  for (int i = 0; i < N; ++i) x[i] = fmin(fmax(x[i], -708.0), 708.0);
#pragma unroll
  for (int i = 0; i < N; ++i) kd[i] = fma(x[i], 92.33248261689366 /* 64/ln2 */, kMagic);
#pragma unroll
  for (int i = 0; i < N; ++i) { ki[i] = __double2loint(kd[i]); kd[i] -= kMagic; }
#pragma unroll
  for (int i = 0; i < N; ++i) r[i] = fma(kd[i], -0.01083042469326756 /* ln2/64, low 21 bits clear */, x[i]);
#pragma unroll
  for (int i = 0; i < N; ++i) r[i] = fma(kd[i], -2.9815858269852933e-12 /* ln2/64 - hi */, r[i]);
#pragma unroll
  for (int i = 0; i < N; ++i) p[i] = fma(r[i], 1.0 / 120, 1.0 / 24);
#pragma unroll
  for (int i = 0; i < N; ++i) p[i] = fma(p[i], r[i], 1.0 / 6);
#pragma unroll
  for (int i = 0; i < N; ++i) p[i] = fma(p[i], r[i], 0.5);
#pragma unroll
  for (int i = 0; i < N; ++i) p[i] = fma(p[i], r[i], 1.0);
#pragma unroll
  for (int i = 0; i < N; ++i) p[i] = fma(p[i], r[i], 1.0);
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const double t = tab[ki[i] & (kExpTab - 1)] * p[i];
    // scale by 2^(k >> 6): |x| <= 708 keeps the result a normal number
    x[i] = __hiloint2double(__double2hiint(t) + ((ki[i] >> 6) << 20), __double2loint(t));
  }
}
template <int N>
__device__ __forceinline__ void rcp_batch(double (&a)[N]) {  // a >= 1
  double r[N], e[N];
#pragma unroll
  for (int i = 0; i < N; ++i) asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r[i]) : "d"(a[i]));
#pragma unroll
  for (int it = 0; it < 3; ++it) {
#pragma unroll
    for (int i = 0; i < N; ++i) e[i] = fma(-a[i], r[i], 1.0);
#pragma unroll
    for (int i = 0; i < N; ++i) r[i] = fma(r[i], e[i], r[i]);
  }
#pragma unroll
  for (int i = 0; i < N; ++i) a[i] = r[i];
}

// Phase 1 of one segment: partial (g, u) of the column over this segment's rows.  The
// coordinate's state as it is now is returned in `C` for phase 3, where the update happens.
//   one rank      the partials are accumulated with fp64 reductions into W.acc (two fire-and-forget
//                 REDs per segment); phase 3 reads the totals after the grid barrier.  No dependent
//                 chain at all behind the arithmetic.
//   several ranks the column total has to leave for the peers BEFORE the barrier, so the segments
//                 of a column are reduced in fixed order by the last one to arrive (L2 partials +
//                 acq_rel counter) and that warp pushes the entries.
template <bool BINARY, bool SINGLE>
__device__ __forceinline__ double grad_segment(const psb_darlin_worker& W, const psb_darlin_server& S,
                                               const ServerHyper& h, const SegRegs& R, int lane,
                                               const double* __restrict__ exp_tab, uint32_t tag,
                                               Carry& C) {
  // ONE latency level: active word, trust region, replica weight and the margins of the first
  // batch are all requested before anything is known about the column
  const uint32_t aw = __ldcg(W.active + (R.j >> 5));
  double d = __ldcg(W.delta + R.j);
  C.w = __ldcg(W.w + R.j);
  double zz0[kUnroll];
#pragma unroll
  for (int k = 0; k < kUnroll; ++k) zz0[k] = R.idx[k] != 0xffffffffu ? __ldcg(W.z + R.idx[k]) : 0.0;
  C.dl = d;
  // a weight that was KKT-filtered stays marked after a filter reset: nothing
  // is pushed or pulled for it, its bits stay as they are
  C.act = (aw >> (R.j & 31) & 1u) && !is_marked(C.w);
  if (!C.act || R.len == 0) return 0;  // a rank without nonzeros in the column pushes nothing
  if (BINARY) d = exp(d);
  double g = 0, u = 0;
  uint32_t idx[kUnroll];
  float sv[kUnroll];
#pragma unroll
  for (int k = 0; k < kUnroll; ++k) { idx[k] = R.idx[k]; sv[k] = R.sv[k]; }
  for (int base = 0; base < R.len; base += 32 * kUnroll) {
    if (base > 0) {
#pragma unroll
      for (int k = 0; k < kUnroll; ++k) {
        const int e = base + k * 32 + lane;
        const bool ok = e < R.len;
        idx[k] = ok ? __ldg(W.row_idx + R.beg + e) : 0xffffffffu;
        sv[k] = (ok && !BINARY) ? __ldg(W.sval + R.beg + e) : 0.f;
      }
      if (BINARY) {
#pragma unroll
        for (int k = 0; k < kUnroll; ++k) sv[k] = idx[k] != 0xffffffffu ? __ldg(W.y + idx[k]) : 0.f;
      }
    }
    double zz[kUnroll];
#pragma unroll
    for (int k = 0; k < kUnroll; ++k)
      zz[k] = base == 0 ? zz0[k] : (idx[k] != 0xffffffffu ? __ldcg(W.z + idx[k]) : 0.0);
    // branch-free over the batch so the four exp/divide chains interleave (fp64 latency bound);
    // a padding entry has x == 0 and contributes exactly nothing
    double tau[kUnroll], x[kUnroll], cv[kUnroll];
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) x[k] = (double)sv[k];
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) { tau[k] = zz[k]; cv[k] = BINARY ? d : fabs(x[k]) * d; }
    exp_batch(tau, exp_tab);
    if (!BINARY) exp_batch(cv, exp_tab);
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) tau[k] += 1.0;
    rcp_batch(tau);
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      g -= tau[k] * x[k];
      u += fmin(tau[k] * (1.0 - tau[k]) * cv[k], 0.25) * x[k] * x[k];
    }
  }
  g = warp_sum(g);
  u = warp_sum(u);
  if (SINGLE) {
    if (lane == 0) {
      double* acc = W.acc + 2 * ((int64_t)(tag & 1u) * W.n_cols + R.j);
      atomicAdd(acc, g);      // compiled to RED: nobody waits for the result
      atomicAdd(acc + 1, u);
    }
    return 0;
  }
  if (R.single) {
    push_entries(W, R, g, u, tag, lane);
    return 0;
  }
  int old = 0;
  if (lane == 0) {
    __stcg(W.part + 2 * R.v, g);
    __stcg(W.part + 2 * R.v + 1, u);
    // release: the partial is visible before the count; acquire: the last arriver sees all
    old = (int)atom_add_acqrel_gpu(reinterpret_cast<uint32_t*>(W.col_cnt + R.j), 1u);
  }
  old = __shfl_sync(FULL, old, 0);
  if (old == R.nseg - 1) {
    const int64_t v0 = R.v - R.sc;
    double sg = 0, su = 0;
    for (int k = lane; k < R.nseg; k += 32) {  // fixed segment order: bit-reproducible
      sg += __ldcg(W.part + 2 * (v0 + k));
      su += __ldcg(W.part + 2 * (v0 + k) + 1);
    }
    sg = warp_sum(sg);
    su = warp_sum(su);
    push_entries(W, R, sg, su, tag, lane);
    if (lane == 0) W.col_cnt[R.j] = 0;
  }
  return 0;
}

// Segments beyond a warp's first one of a step (steps with more segments than resident warps)
// cannot keep their phase-1 state in registers; it is parked in W.stash[v] = {w, delta | act}.
__device__ __forceinline__ void stash_put(const psb_darlin_worker& W, int64_t v, const Carry& C, int lane) {
  if (lane == 0) {
    __stcg(W.stash + 2 * v, C.act ? C.w : nan_mark());
    __stcg(W.stash + 2 * v + 1, C.dl);
  }
}
__device__ __forceinline__ void stash_get(const psb_darlin_worker& W, int64_t v, Carry& C) {
  C.w = __ldcg(W.stash + 2 * v);
  C.dl = __ldcg(W.stash + 2 * v + 1);
  C.act = !is_marked(C.w);
}

template <bool BINARY, bool SINGLE>
__device__ __forceinline__ double grad_phase_packed(const psb_darlin_worker& W, const psb_darlin_server& S,
                                                    const ServerHyper& h, const StepPlan& plan,
                                                    int bps, int64_t gw, int64_t nw, int lane,
                                                    const SegRegs& pre, const double* __restrict__ exp_tab,
                                                    uint32_t tag, Carry& C0) {
  const int64_t total = plan.pref[bps];
  double vio_max = 0;
  C0.act = false;
  if (gw < total) {
    if (pre.v == gw) {
      vio_max = grad_segment<BINARY, SINGLE>(W, S, h, pre, lane, exp_tab, tag, C0);
    } else {
      SegRegs R;
      stage_segment<BINARY, SINGLE>(R, W, plan, bps, gw, lane);
      vio_max = grad_segment<BINARY, SINGLE>(W, S, h, R, lane, exp_tab, tag, C0);
    }
  }
  for (int64_t v = gw + nw; v < total; v += nw) {
    SegRegs R;
    Carry C;
    stage_segment<BINARY, SINGLE>(R, W, plan, bps, v, lane);
    vio_max = fmax(vio_max, grad_segment<BINARY, SINGLE>(W, S, h, R, lane, exp_tab, tag, C));
    stash_put(W, v, C, lane);
  }
  return vio_max;
}

template <bool BINARY>
__device__ __forceinline__ void scatter_margins(const psb_darlin_worker& W, const SegRegs& R, double d,
                                                int lane) {
  uint32_t idx[kUnroll];
  float sv[kUnroll];
#pragma unroll
  for (int k = 0; k < kUnroll; ++k) { idx[k] = R.idx[k]; sv[k] = R.sv[k]; }
  for (int base = 0; base < R.len; base += 32 * kUnroll) {
    if (base > 0) {
#pragma unroll
      for (int k = 0; k < kUnroll; ++k) {
        const int e = base + k * 32 + lane;
        const bool ok = e < R.len;
        idx[k] = ok ? __ldg(W.row_idx + R.beg + e) : 0xffffffffu;
        sv[k] = (ok && !BINARY) ? __ldg(W.sval + R.beg + e) : 0.f;
      }
      if (BINARY) {
#pragma unroll
        for (int k = 0; k < kUnroll; ++k) sv[k] = idx[k] != 0xffffffffu ? __ldg(W.y + idx[k]) : 0.f;
      }
    }
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      if (idx[k] == 0xffffffffu) continue;
      atomicAdd(W.z + idx[k], d * (double)sv[k]);  // dual_i *= exp(y_i d x_ij)  <=>  z_i += d * sval
    }
  }
}

// Phase 3 of one segment: get the column's total (g, u) - one rank: from the accumulator; several
// ranks: the holders' entries in this rank's landing table, summed in worker order - apply the
// update (ref darlin.h:216-245 on the replica, 436-449), and scatter dw into the margins.  Every
// segment of a column computes the same update; the first one writes the replica and, on the
// owner, the shard.
template <bool BINARY, bool SINGLE>
__device__ __forceinline__ double pull_segment_replica(const psb_darlin_worker& W, const psb_darlin_server& S,
                                                       const ServerHyper& h, const SegRegs& R,
                                                       const Carry& C, int lane, uint32_t tag) {
  const bool mine = SINGLE || R.owner == W.rank;
  if (!C.act || (R.len == 0 && !mine)) return 0;
  double g = 0, u = 0;
  if (SINGLE) {
    if (lane == 0) {
      const double* acc = W.acc + 2 * ((int64_t)(tag & 1u) * W.n_cols + R.j);
      g = __ldcg(acc);
      u = __ldcg(acc + 1);
    }
    g = __shfl_sync(FULL, g, 0);
    u = __shfl_sync(FULL, u, 0);
  } else {
    double gv = 0, uv = 0;
    if (lane < W.n_workers && (R.has >> lane & 1u)) {
      const char* at = W.land2[W.rank] + entry_index(W, tag, lane, R.j);
      if (W.push_fp32) entry_poll_f32(at, tag, gv, uv);
      else entry_poll(at, tag, gv, uv);
    }
    for (int i = 0; i < W.n_workers; ++i) {  // worker order: bit-reproducible, same on every rank
      g += __shfl_sync(FULL, gv, i);
      u += __shfl_sync(FULL, uv, i);
    }
  }
  const double w = C.w, dl = C.dl;
  u = u / h.eta + 1e-10;
  const double g_pos = g + h.lambda, g_neg = g - h.lambda;
  double d = -w, vio = 0;
  if (w == 0) {
    if (g_pos < 0) {
      vio = -g_pos;
    } else if (g_neg > 0) {
      vio = g_neg;
    } else if (g_pos > h.thr && g_neg < -h.thr) {  // KKT filter: the coordinate leaves the active set
      if (R.first && lane == 0) {
        bit_clear_atomic(W.active, R.j);
        W.w[R.j] = nan_mark();
        if (mine) {
          bit_clear_atomic(S.active, R.off);
          S.w[R.off] = nan_mark();
          S.dw[R.off] = 0;
        }
      }
      return 0;
    }
  }
  if (g_pos <= u * w) {
    d = -g_pos / u;
  } else if (g_neg >= u * w) {
    d = -g_neg / u;
  }
  d = fmin(dl, fmax(-dl, d));
  const double w_new = w + d;
  const double dw = w_new - w;
  if (R.first && lane == 0) {
    const double dl_new = fmin(h.delta_max, 2 * fabs(d) + .1);
    W.w[R.j] = w_new;
    W.delta[R.j] = dl_new;
    if (mine) {
      S.w[R.off] = w_new;
      S.dw[R.off] = dw;
      S.delta[R.off] = dl_new;
    }
  }
  if (dw != 0) scatter_margins<BINARY>(W, R, dw, lane);
  return (mine && R.first) ? vio : 0;
}

template <bool BINARY, bool SINGLE>
__device__ __forceinline__ double pull_phase_packed(const psb_darlin_worker& W, const psb_darlin_server& S,
                                                    const ServerHyper& h, const StepPlan& plan, int bps,
                                                    int64_t gw, int64_t nw, int lane, const SegRegs& pre,
                                                    uint32_t tag, const Carry& C0) {
  const int64_t total = plan.pref[bps];
  double vio = 0;
  if (gw < total) {
    if (pre.v == gw) {
      vio = pull_segment_replica<BINARY, SINGLE>(W, S, h, pre, C0, lane, tag);
    } else {
      SegRegs R;
      stage_segment<BINARY, SINGLE>(R, W, plan, bps, gw, lane);
      vio = pull_segment_replica<BINARY, SINGLE>(W, S, h, R, C0, lane, tag);
    }
  }
  for (int64_t v = gw + nw; v < total; v += nw) {
    SegRegs R;
    Carry C;
    stage_segment<BINARY, SINGLE>(R, W, plan, bps, v, lane);
    stash_get(W, v, C);
    vio = fmax(vio, pull_segment_replica<BINARY, SINGLE>(W, S, h, R, C, lane, tag));
  }
  return vio;
}

// One rank: the accumulators of the step's columns go back to zero once every reader is through
// (after the step's last barrier) - the other parity is in use by then, so nothing races.
__device__ __forceinline__ void clear_accumulators(const psb_darlin_worker& W, const StepPlan& plan, int bps,
                                                   int64_t gw, int64_t nw, int lane, int64_t first_j,
                                                   bool first_known, uint32_t tag) {
  const int64_t total = plan.pref[bps];
  double* acc = W.acc + 2 * (int64_t)(tag & 1u) * W.n_cols;
  for (int64_t v = gw; v < total; v += nw) {
    int64_t j = -1;
    if (v == gw && first_known) {
      j = first_j;
    } else {
      const int q = locate(plan, bps, v);
      const SegMeta m = load_meta(W.seg_meta, plan.first[q] + (v - plan.pref[q]));
      if (m.first) j = m.j;
    }
    if (j >= 0 && lane == 0) {
      __stcg(acc + 2 * j, 0.0);
      __stcg(acc + 2 * j + 1, 0.0);
    }
  }
}

__device__ __forceinline__ void prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
// warm L2 with the CSC ranges this CTA's warps will read in the next step
template <bool BINARY>
__device__ __forceinline__ int4 prefetch_step(const psb_darlin_worker& W, const StepPlan& plan, int bps,
                                              int64_t gw, int64_t nw, int lane) {
  const int64_t total = plan.pref[bps];
  int4 first_meta = make_int4(0, -1, 0, 0);  // y < 0: nothing fetched
  for (int64_t v = gw; v < total; v += nw) {
    const int q = locate(plan, bps, v);
    const int64_t s = plan.first[q] + (v - plan.pref[q]);
    const int4 raw = __ldg(reinterpret_cast<const int4*>(W.seg_meta) + s);
    if (v == gw) first_meta = raw;  // kept for the staging that follows one barrier later
    const SegMeta m = decode_meta(raw);
    const int lines = (m.len * 4 + 127) / 128 + 1;
    if (lane < lines) prefetch_l2(reinterpret_cast<const char*>(W.row_idx + m.beg) + lane * 128);
    if (!BINARY && lane >= 16 && lane - 16 < lines)
      prefetch_l2(reinterpret_cast<const char*>(W.sval + m.beg) + (lane - 16) * 128);
  }
  return first_meta;
}


// The thread that signals and polls the grid barriers of a step: lane 0 of the last warp, which
// owns a segment only in steps with more than (kWarps-1) * grid segments.  The other warps use
// the barrier's flight time for the next steps' plan, L2 prefetch and register staging.
constexpr int kSpinner = (kWarps - 1) * 32;
__device__ __forceinline__ void grid_arrive_by(uint32_t* ctr, int leader) {
  __syncthreads();
  if (threadIdx.x == leader) red_release_gpu_add(ctr, 1u);
}
__device__ __forceinline__ void grid_wait_by(const uint32_t* ctr, uint32_t target, int leader) {
  if (threadIdx.x == leader) spin_gpu_ge(ctr, target);
  __syncthreads();
}
// The last barrier of a multi-rank step, plus the flow control of the landing tables: the next
// step (sequence `next`) may overwrite entries of step next-2, so every peer must have finished
// that step.  The flags live in this rank's memory and are loaded before the barrier spin, so
// in the common case (peers at most one step behind) they cost no extra round trip.
__device__ __forceinline__ void grid_wait_flow(const uint32_t* ctr, uint32_t target,
                                               const uint64_t* done_local, int n, uint64_t next) {
  if ((int)threadIdx.x >= kSpinner) {
    const int lane = threadIdx.x - kSpinner;
    const bool peer = lane < n;
    uint64_t seen = peer ? ld_relaxed_sys(done_local + lane) : 0;
    if (lane == 0) spin_gpu_ge(ctr, target);
    uint32_t spins = 0;
    while (peer && seen + 2 < next) {
      seen = ld_relaxed_sys(done_local + lane);
      if (++spins > kSpinLimit) __trap();
    }
  }
  __syncthreads();
}

// bar[0] phase-1 arrivals, bar[2] phase-3 arrivals, bar[3] generic barrier (begin / evaluate).
// All zero between launches.
//
// `G` CTAs numbered `cta` = 0..G-1 form one rank.  In a real run that is the whole grid; the
// emulation kernel below runs several ranks side by side in one cooperative launch.
template <bool BINARY, bool SINGLE>
__device__ __forceinline__ void darlin_pass_body(
    const psb_darlin_worker& W, const psb_darlin_server& S, const psb_darlin_ctrl* __restrict__ ctrl,
    const int32_t* __restrict__ order, const int64_t* __restrict__ blk_tab, int n_steps, int bps,
    uint32_t* bar, psb_darlin_progress* out_local, psb_darlin_progress* const* out_peer, int n_peers,
    double* scratch, unsigned long long* prof, const uint32_t G, const uint32_t cta,
    FusedPlan* plans, double (*sh)[kWarps], double* exp_tab) {
  fill_exp_table(exp_tab);
  // optional per-phase time accounting (thread 0 of every CTA, summed into prof[0..4])
  unsigned long long t_mark = 0, t_acc[5] = {0, 0, 0, 0, 0};
#define PSB_PROF(slot)                                         \
  if (prof && threadIdx.x == 0) {                              \
    unsigned long long now_;                                   \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now_));   \
    t_acc[slot] += now_ - t_mark;                              \
    t_mark = now_;                                             \
  }
  if (prof && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_mark));
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  // segment v of a step goes to CTA v % G (warp v / G): a step with fewer segments than warps
  // is spread over every SM instead of filling the first CTAs
  const int64_t gw = (ctrl->flags & 4) ? (int64_t)cta * kWarps + wid : (int64_t)wid * G + cta;
  const int64_t nw = (int64_t)G * kWarps;
  const int64_t gt = (int64_t)cta * kThreads + threadIdx.x, nt = (int64_t)G * kThreads;
  SegRegs pre;
  pre.v = -1;
  uint32_t nbar = 0;
  const ServerHyper h{ctrl->lambda, ctrl->eta, ctrl->delta_max, ctrl->kkt_threshold};
  const uint64_t seq_base = ctrl->seq_base;
  const bool do_prefetch = !(ctrl->flags & 1);

  // ---- pass begin (ref darlin.h:176-182, 297-300) ----
  if (gt == 0) *S.violation = 0;
  if (ctrl->reset_kkt_filter) {
    for (int64_t i = gt; i < (W.n_cols + 31) / 32; i += nt) {
      const int64_t rem = W.n_cols - i * 32;
      W.active[i] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
    }
    for (int64_t i = gt; i < (S.n_keys + 31) / 32; i += nt) {
      const int64_t rem = S.n_keys - i * 32;
      S.active[i] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
    }
    if (ctrl->revive_on_reset) {
      for (int64_t i = gt; i < S.n_keys; i += nt)
        if (is_marked(__ldcg(S.w + i))) S.w[i] = 0;
      for (int64_t i = gt; i < W.n_cols; i += nt)
        if (is_marked(__ldcg(W.w + i))) W.w[i] = 0;
    }
  }
  if (wid == 0 && n_steps > 0) fused_plan(plans[0], blk_tab, order, 0, bps);
  if (wid == 1 && n_steps > 1) fused_plan(plans[1], blk_tab, order, 1, bps);
  grid_arrive(bar + 3);
  if (n_steps > 0 && gw < plans[0].wk.pref[bps]) stage_segment<BINARY, SINGLE>(pre, W, plans[0].wk, bps, gw, lane);
  grid_wait(bar + 3, ++nbar * G);

  for (int step = 0; step < n_steps; ++step) {
    const FusedPlan& P = plans[step % 3];
    const FusedPlan& Pn = plans[(step + 1) % 3];
    const uint64_t seq = seq_base + (uint64_t)step + 1;
    const bool has_next = step + 1 < n_steps;
    Carry C0;
    // ---- phase 1: gradient, then accumulate (one rank) or push (several) ----
    grad_phase_packed<BINARY, SINGLE>(W, S, h, P.wk, bps, gw, nw, lane, pre, exp_tab, (uint32_t)seq, C0);
    if (prof && threadIdx.x == 0) {  // per-step straggler: max phase-1 time and its CTA
      unsigned long long now_;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now_));
      atomicMax(prof + 8 + 2 * step, ((now_ - t_mark) << 20) | cta);
    }
    PSB_PROF(0)
    // barrier A (local): every margin read of phase 1 precedes every margin update of phase 3.
    // Several ranks: the pushes are in flight meanwhile.
    grid_arrive_by(bar, kSpinner);
    if (wid == 1 && step + 2 < n_steps) fused_plan(plans[(step + 2) % 3], blk_tab, order, step + 2, bps);
    int4 next_meta = make_int4(0, -1, 0, 0);
    if (has_next && do_prefetch && wid != kWarps - 1) next_meta = prefetch_step<BINARY>(W, Pn.wk, bps, gw, nw, lane);
    PSB_PROF(1)
    grid_wait_by(bar, (uint32_t)(step + 1) * G, kSpinner);
    PSB_PROF(2)
    // ---- phase 3: (several ranks: collect the entries, update the replica) margins ----
    {
      double vio = pull_phase_packed<BINARY, SINGLE>(W, S, h, P.wk, bps, gw, nw, lane, pre, (uint32_t)seq, C0);
      vio = warp_max(vio);
      if (lane == 0 && vio > 0) atomic_max_nonneg(S.violation, vio);
    }
    if (prof && threadIdx.x == 0) {
      unsigned long long now_;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now_));
      atomicMax(prof + 9 + 2 * step, ((now_ - t_mark) << 20) | cta);
    }
    PSB_PROF(3)
    grid_arrive_by(bar + 2, kSpinner);
    const bool first_known = pre.v == gw;
    const int64_t first_j = (first_known && pre.first) ? pre.j : -1;
    pre.v = -1;
    if (has_next && wid != kWarps - 1 && gw < Pn.wk.pref[bps])
      stage_segment<BINARY, SINGLE>(pre, W, Pn.wk, bps, gw, lane, next_meta);
    if (SINGLE) {
      grid_wait_by(bar + 2, (uint32_t)(step + 1) * G, kSpinner);
      clear_accumulators(W, P.wk, bps, gw, nw, lane, first_j, first_known, (uint32_t)seq);
    } else {
      grid_wait_flow(bar + 2, (uint32_t)(step + 1) * G, W.done_local, W.n_workers, seq + 1);
      // this rank has consumed every entry of the step: tell the peers (plain stores, the reads
      // they guard have returned)
      if (cta == 0 && (int)threadIdx.x >= kSpinner && (int)threadIdx.x - kSpinner < W.n_workers)
        *reinterpret_cast<volatile uint64_t*>(W.done_peer[threadIdx.x - kSpinner]) = seq;
    }
    PSB_PROF(4)
  }
  if (prof && threadIdx.x == 0)
    for (int i = 0; i < 5; ++i) atomicAdd(prof + i, t_acc[i]);
#undef PSB_PROF

  // ---- evaluate (ref darlin.h:248-265, 504-511) ----
  double loss = 0, l1 = 0, dn = 0, da = 0;
  for (int64_t i = gt; i < W.n_rows; i += nt) {
    const double zi = __ldcg(W.z + i);
    loss += zi > 0 ? log1p(exp(-zi)) : (-zi + log1p(exp(zi)));
  }
  for (int64_t i = gt; i < S.n_keys; i += nt) {
    const double wi = __ldcg(S.w + i);
    if (!is_marked(wi) && wi != 0) { dn += 1; l1 += fabs(wi); }
  }
  for (int64_t i = gt; i < (S.n_keys + 31) / 32; i += nt) da += __popc(__ldcg(S.active + i));
  loss = warp_sum(loss); l1 = warp_sum(l1); dn = warp_sum(dn); da = warp_sum(da);
  if (lane == 0) { sh[0][wid] = loss; sh[1][wid] = l1; sh[2][wid] = dn; sh[3][wid] = da; }
  __syncthreads();
  if (threadIdx.x < 4) {
    double a = 0;
    for (int i = 0; i < kWarps; ++i) a += sh[threadIdx.x][i];
    __stcg(scratch + 4 * (int64_t)cta + threadIdx.x, a);
  }
  grid_arrive(bar + 3);
  if (cta != 0) return;
  grid_wait(bar + 3, ++nbar * G);
  double acc[4] = {0, 0, 0, 0};
  for (int b = threadIdx.x; b < (int)G; b += kThreads)
    for (int c = 0; c < 4; ++c) acc[c] += __ldcg(scratch + 4 * (int64_t)b + c);
  for (int c = 0; c < 4; ++c) acc[c] = warp_sum(acc[c]);
  if (lane == 0)
    for (int c = 0; c < 4; ++c) sh[c][wid] = acc[c];
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int c = 0; c < 4; ++c) {
      double a = 0;
      for (int i = 0; i < kWarps; ++i) a += sh[c][i];
      acc[c] = a;
    }
    // every CTA has arrived at the last barrier; nobody reads the counters again
    bar[0] = 0; bar[1] = 0; bar[2] = 0; bar[3] = 0;
    const double vio = __ldcg(S.violation);
    const uint64_t stamp = seq_base + (uint64_t)n_steps;
    for (int r = -1; r < n_peers; ++r) {
      psb_darlin_progress* o = r < 0 ? out_local : out_peer[r];
      if (!o) continue;
      o->loss = acc[0]; o->penalty = acc[1] * h.lambda; o->violation = vio;
      o->nnz_w = (uint64_t)(acc[2] + 0.5); o->nnz_active = (uint64_t)(acc[3] + 0.5);
    }
    if (n_peers > 0) __threadfence_system();
    for (int r = -1; r < n_peers; ++r) {
      psb_darlin_progress* o = r < 0 ? out_local : out_peer[r];
      if (o) st_release_sys(&o->seq, stamp);
    }
  }
}

template <bool BINARY, bool SINGLE>
__global__ void __launch_bounds__(kThreads, 2)
k_darlin_pass(const __grid_constant__ psb_darlin_worker W, const __grid_constant__ psb_darlin_server S,
              const psb_darlin_ctrl* __restrict__ ctrl, const int32_t* __restrict__ order,
              const int64_t* __restrict__ blk_tab, int n_steps, int bps, uint32_t* bar,
              psb_darlin_progress* out_local, psb_darlin_progress* const* out_peer, int n_peers,
              double* scratch, unsigned long long* prof) {
  __shared__ FusedPlan plans[3];  // steps s, s+1, s+2: built two steps ahead, off the critical path
  __shared__ double sh[4][kWarps];
  __shared__ double exp_tab[kExpTab];
  darlin_pass_body<BINARY, SINGLE>(W, S, ctrl, order, blk_tab, n_steps, bps, bar, out_local, out_peer,
                                   n_peers, scratch, prof, gridDim.x, blockIdx.x, plans, sh, exp_tab);
}

// Several ranks of the multi-rank pass in ONE cooperative launch on one GPU: CTAs
// [r*G, (r+1)*G) are rank r.  Kernels of different ranks that wait on each other must not be
// separate launches on one GPU (nothing guarantees that they run together); here all ranks'
// CTAs are co-resident by construction, so the single-GPU tests exercise exactly the code of the
// multi-GPU run: tagged entries, replicated update, flow-control flags.  The "peers" are plain
// pointers into the same device memory.
template <bool BINARY>
__global__ void __launch_bounds__(kThreads, 2)
k_darlin_pass_emulated(const psb_darlin_rank_args* __restrict__ ranks, int G, int n_steps, int bps,
                       int n_peers) {
  __shared__ FusedPlan plans[3];
  __shared__ double sh[4][kWarps];
  __shared__ double exp_tab[kExpTab];
  __shared__ psb_darlin_rank_args A;
  const int r = blockIdx.x / G;
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ranks + r);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&A);
    for (int i = threadIdx.x; i < (int)(sizeof(A) / 4); i += kThreads) dst[i] = src[i];
  }
  __syncthreads();
  darlin_pass_body<BINARY, false>(A.wk, A.sv, A.ctrl, A.order, A.blk_tab, n_steps, bps, A.bar, A.out_local,
                                  A.out_peer, n_peers, A.scratch, nullptr, (uint32_t)G,
                                  (uint32_t)(blockIdx.x % G), plans, sh, exp_tab);
}

}  // namespace psb

using namespace psb;

PSB_API int psb_darlin_pass_begin(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                                  const psb_darlin_ctrl* ctrl, void* stream) {
  PSB_REQUIRE(ctrl, "psb_darlin_pass_begin: ctrl is NULL");
  int64_t n = 1;
  if (wk) n = max(n, (wk->n_cols + 31) / 32);
  if (sv) n = max(n, sv->n_keys);
  int grid = min(div_up(n, 256), 4 * sm_count());
  k_pass_begin<<<grid, 256, 0, (cudaStream_t)stream>>>(
      wk ? wk->active : nullptr, wk ? wk->n_cols : 0, sv ? sv->active : nullptr,
      sv ? sv->n_keys : 0, sv ? sv->w : nullptr, sv ? sv->violation : nullptr, ctrl);
  PSB_LAUNCH_CHECK();
  return 0;
}

static int check_step_args(const char* who, int bps, int grid) {
  PSB_REQUIRE(bps >= 1 && bps <= kMaxBps, "%s: blocks_per_step=%d outside [1,%d]", who, bps, kMaxBps);
  PSB_REQUIRE(grid >= 1, "%s: grid=%d", who, grid);
  return 0;
}

PSB_API int psb_darlin_grad_push(const psb_darlin_worker* wk, const psb_darlin_ctrl* ctrl,
                                 const int32_t* order, int32_t step, int32_t bps, int32_t grid,
                                 void* stream) {
  PSB_REQUIRE(wk && ctrl && order, "psb_darlin_grad_push: NULL argument");
  if (int rc = check_step_args("psb_darlin_grad_push", bps, grid)) return rc;
  PSB_REQUIRE(wk->n_servers <= PSB_MAX_PEERS, "n_servers > PSB_MAX_PEERS");
  if (wk->sval)
    k_grad_push<false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(*wk, ctrl, order, step, bps);
  else
    k_grad_push<true><<<grid, kThreads, 0, (cudaStream_t)stream>>>(*wk, ctrl, order, step, bps);
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int psb_darlin_server_update(const psb_darlin_server* sv, const psb_darlin_ctrl* ctrl,
                                     const int32_t* order, int32_t step, int32_t bps,
                                     int32_t grid, void* stream) {
  PSB_REQUIRE(sv && ctrl && order, "psb_darlin_server_update: NULL argument");
  if (int rc = check_step_args("psb_darlin_server_update", bps, grid)) return rc;
  PSB_REQUIRE(sv->n_workers <= PSB_MAX_PEERS, "n_workers > PSB_MAX_PEERS");
  k_server_update<<<grid, kThreads, 0, (cudaStream_t)stream>>>(*sv, ctrl, order, step, bps);
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int psb_darlin_pull_update(const psb_darlin_worker* wk, const psb_darlin_ctrl* ctrl,
                                   const int32_t* order, int32_t step, int32_t bps, int32_t grid,
                                   void* stream) {
  PSB_REQUIRE(wk && ctrl && order, "psb_darlin_pull_update: NULL argument");
  if (int rc = check_step_args("psb_darlin_pull_update", bps, grid)) return rc;
  if (wk->sval)
    k_pull_update<false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(*wk, ctrl, order, step, bps);
  else
    k_pull_update<true><<<grid, kThreads, 0, (cudaStream_t)stream>>>(*wk, ctrl, order, step, bps);
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int64_t psb_darlin_evaluate_scratch_bytes(int32_t grid) {
  return (int64_t)grid * 4 * sizeof(double) + 16;
}

PSB_API int psb_darlin_evaluate(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                                const psb_darlin_ctrl* ctrl, psb_darlin_progress* out_local,
                                psb_darlin_progress* const* out_peer, int32_t n_peers,
                                uint64_t seq, double* scratch, int32_t grid, void* stream) {
  PSB_REQUIRE(ctrl && scratch && grid >= 1, "psb_darlin_evaluate: bad argument");
  k_evaluate<<<grid, kThreads, 0, (cudaStream_t)stream>>>(
      wk ? wk->z : nullptr, wk ? wk->n_rows : 0, sv ? sv->w : nullptr,
      sv ? sv->active : nullptr, sv ? sv->n_keys : 0, sv ? sv->violation : nullptr, ctrl,
      out_local, out_peer, out_peer ? n_peers : 0, seq, scratch);
  PSB_LAUNCH_CHECK();
  return 0;
}

PSB_API int psb_darlin_enqueue_pass(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                                    const psb_darlin_ctrl* ctrl, const int32_t* order,
                                    int32_t n_steps, int32_t bps, int32_t grid_worker,
                                    int32_t grid_server, psb_darlin_progress* out_local,
                                    psb_darlin_progress* const* out_peer, int32_t n_peers,
                                    double* eval_scratch, int32_t grid_eval, void* stream) {
  if (int rc = psb_darlin_pass_begin(wk, sv, ctrl, stream)) return rc;
  for (int i = 0; i < n_steps; ++i) {
    if (wk) if (int rc = psb_darlin_grad_push(wk, ctrl, order, i, bps, grid_worker, stream)) return rc;
    if (sv) if (int rc = psb_darlin_server_update(sv, ctrl, order, i, bps, grid_server, stream)) return rc;
    if (wk) if (int rc = psb_darlin_pull_update(wk, ctrl, order, i, bps, grid_worker, stream)) return rc;
  }
  return psb_darlin_evaluate(wk, sv, ctrl, out_local, out_peer, n_peers, 0, eval_scratch,
                             grid_eval, stream);
}

// Whole pass in ONE cooperative launch (see k_darlin_pass).  `bar` is a zero-initialised
// uint32 in device memory owned by the caller; `grid` must not exceed
// psb_darlin_fused_max_grid() so that all CTAs are co-resident.
PSB_API int psb_darlin_fused_max_grid(int32_t binary) {
  int per_sm = 0;
  cudaError_t e = binary
      ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_darlin_pass<true, false>, kThreads, 0)
      : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_darlin_pass<false, false>, kThreads, 0);
  if (e != cudaSuccess || per_sm < 1) return 0;
  return per_sm * sm_count();
}

PSB_API int psb_darlin_fused_pass(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                                  const psb_darlin_ctrl* ctrl, const int32_t* order,
                                  const int64_t* blk_tab, int32_t n_steps, int32_t bps,
                                  int32_t grid, uint32_t* bar, psb_darlin_progress* out_local,
                                  psb_darlin_progress* const* out_peer, int32_t n_peers,
                                  double* scratch, uint64_t* prof, void* stream) {
  PSB_REQUIRE(wk && sv && ctrl && order && blk_tab && bar && scratch, "psb_darlin_fused_pass: NULL argument");
  if (int rc = check_step_args("psb_darlin_fused_pass", bps, grid)) return rc;
  PSB_REQUIRE(grid <= psb_darlin_fused_max_grid(wk->sval == nullptr),
              "psb_darlin_fused_pass: grid=%d exceeds the co-resident maximum", grid);
  int np = out_peer ? n_peers : 0;
  void* args[] = {(void*)wk, (void*)sv, (void*)&ctrl, (void*)&order, (void*)&blk_tab, (void*)&n_steps,
                  (void*)&bps, (void*)&bar, (void*)&out_local, (void*)&out_peer, (void*)&np,
                  (void*)&scratch, (void*)&prof};
  const bool single = wk->n_workers == 1 && wk->n_servers == 1;
  if (!single) {
    PSB_REQUIRE(wk->n_workers <= PSB_MAX_PEERS && wk->n_workers <= 32, "psb_darlin_fused_pass: too many ranks");
    PSB_REQUIRE(wk->col_has && wk->stash && wk->done_local && wk->land2[wk->rank],
                "psb_darlin_fused_pass: the multi-rank pass needs the landing tables, the holder masks, "
                "the stash and the flow-control flags (worker.land2, col_has, stash, done_*)");
  }
  const void* fn = wk->sval
      ? (single ? (const void*)k_darlin_pass<false, true> : (const void*)k_darlin_pass<false, false>)
      : (single ? (const void*)k_darlin_pass<true, true> : (const void*)k_darlin_pass<true, false>);
  PSB_CUDA(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kThreads), args, 0, (cudaStream_t)stream));
  return 0;
}

// Emulation of `n_ranks` ranks of the multi-rank pass on ONE GPU in one cooperative launch (see
// k_darlin_pass_emulated).  `ranks_dev`: device array of n_ranks argument records whose peer
// pointers point into the same device; grid_per_rank * n_ranks must not exceed
// psb_darlin_fused_max_grid().
PSB_API int psb_darlin_fused_pass_emulated(const psb_darlin_rank_args* ranks_dev, int32_t n_ranks,
                                           int32_t binary, int32_t n_steps, int32_t bps,
                                           int32_t grid_per_rank, int32_t n_peers, void* stream) {
  PSB_REQUIRE(ranks_dev && n_ranks >= 2 && n_ranks <= PSB_MAX_PEERS, "psb_darlin_fused_pass_emulated: bad ranks");
  if (int rc = check_step_args("psb_darlin_fused_pass_emulated", bps, grid_per_rank)) return rc;
  int per_sm = 0;
  PSB_CUDA(binary ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_darlin_pass_emulated<true>, kThreads, 0)
                  : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_darlin_pass_emulated<false>, kThreads, 0));
  PSB_REQUIRE((int64_t)grid_per_rank * n_ranks <= (int64_t)per_sm * sm_count(),
              "psb_darlin_fused_pass_emulated: %d x %d CTAs exceed the co-resident maximum %d",
              n_ranks, grid_per_rank, per_sm * sm_count());
  void* args[] = {(void*)&ranks_dev, (void*)&grid_per_rank, (void*)&n_steps, (void*)&bps, (void*)&n_peers};
  const void* fn = binary ? (const void*)k_darlin_pass_emulated<true> : (const void*)k_darlin_pass_emulated<false>;
  PSB_CUDA(cudaLaunchCooperativeKernel(fn, dim3(grid_per_rank * n_ranks), dim3(kThreads), args, 0,
                                       (cudaStream_t)stream));
  return 0;
}

// -------------------------------------------------------------------------------------------
// Progress evaluation for callers that must not allocate.
//
// psb_darlin_evaluate needs a scratch buffer (one set of partial sums per CTA plus an arrival
// counter) that the caller allocates.  Some callers cannot do that: a monitoring thread inside
// a training job that has already handed all of HBM to its own pools (a cudaMalloc there fails
// or fragments the shard layout), or a latency-sensitive loop that must not take the
// allocator's lock and the device-wide synchronisation cudaMalloc / cudaFree may imply.  For
// them the scratch lives in the library image: one fixed array per GPU that loads this
// library, held until the process exits.
//
// In host code the array's name is only a handle: passing it to a kernel would hand over a
// host-side address.  Its device address is looked up once per device with
// cudaGetSymbolAddress while that device is current, and that address is what the kernel gets.
// -------------------------------------------------------------------------------------------
#define PSB_DARLIN_EVAL_NOALLOC_MAX_GRID 1024
namespace psb {
constexpr size_t kEvalNoallocBytes =
    (size_t)PSB_DARLIN_EVAL_NOALLOC_MAX_GRID * 4 * sizeof(double) + 256;
static __device__ __align__(256) unsigned char g_eval_noalloc_scratch[kEvalNoallocBytes];

static std::mutex g_eval_noalloc_mu;
static double* g_eval_noalloc_ptr[64] = {nullptr};
}  // namespace psb

PSB_API int32_t psb_darlin_evaluate_noalloc_max_grid(void) { return PSB_DARLIN_EVAL_NOALLOC_MAX_GRID; }

PSB_API int psb_darlin_evaluate_noalloc(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                                        const psb_darlin_ctrl* ctrl, psb_darlin_progress* out_local,
                                        uint64_t seq, int32_t grid, void* stream) {
  PSB_REQUIRE(ctrl && out_local, "psb_darlin_evaluate_noalloc: ctrl and out_local are required");
  PSB_REQUIRE(grid >= 1, "psb_darlin_evaluate_noalloc: grid=%d", grid);
  PSB_REQUIRE(grid <= PSB_DARLIN_EVAL_NOALLOC_MAX_GRID,
              "psb_darlin_evaluate_noalloc: grid=%d exceeds the maximum of %d CTAs that the fixed "
              "%zu-byte scratch array can serve; use psb_darlin_evaluate with your own scratch",
              grid, PSB_DARLIN_EVAL_NOALLOC_MAX_GRID, kEvalNoallocBytes);
  int dev = 0;
  PSB_CUDA(cudaGetDevice(&dev));
  PSB_REQUIRE(dev >= 0 && dev < 64, "psb_darlin_evaluate_noalloc: device ordinal %d unsupported", dev);
  double* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_eval_noalloc_mu);
    if (!g_eval_noalloc_ptr[dev]) {
      // set-up, once per device and with that device current
      void* p = nullptr;
      PSB_CUDA(cudaGetSymbolAddress(&p, g_eval_noalloc_scratch));
      PSB_REQUIRE(p != nullptr, "psb_darlin_evaluate_noalloc: no device address for the scratch array");
      g_eval_noalloc_ptr[dev] = static_cast<double*>(p);
    }
    scratch = g_eval_noalloc_ptr[dev];
  }
  // the array keeps whatever the previous call left: the arrival counter behind the partial
  // sums has to start at zero (the partial sums themselves are fully rewritten by every call)
  PSB_CUDA(cudaMemsetAsync(scratch + 4 * (size_t)grid, 0, 16, (cudaStream_t)stream));
  k_evaluate<<<grid, kThreads, 0, (cudaStream_t)stream>>>(
      wk ? wk->z : nullptr, wk ? wk->n_rows : 0, sv ? sv->w : nullptr,
      sv ? sv->active : nullptr, sv ? sv->n_keys : 0, sv ? sv->violation : nullptr, ctrl,
      out_local, nullptr, 0, seq, scratch);
  PSB_LAUNCH_CHECK();
  return 0;
}
