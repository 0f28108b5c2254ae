/* psb200 — C ABI of the sm_100a CUDA library (libpsb200_cuda.so).
 *
 * Conventions
 *   - every entry point returns 0 on success, non-zero on failure;
 *     psb_last_error() returns the message of the last failure on this thread
 *   - `stream` is a cudaStream_t passed as void*; all work is asynchronous on it
 *   - pointers are device pointers unless the name ends in `_host`
 *   - nothing here allocates unless its comment says so
 *
 * Citations `ref:` name the CPU loop of immars/parameter_server whose behaviour the
 * kernel reproduces (paths relative to the reference tree).
 */
#ifndef PSB200_H_
#define PSB200_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define PSB_MAX_PEERS 16
#define PSB_ABI_VERSION 1

const char* psb_last_error(void);
int psb_abi_version(void);
int psb_device_sm_count(void);

/* ------------------------------------------------------------------------- *
 *  Darlin (L1 logistic regression, block coordinate descent)
 *  ref: src/app/linear_method/darlin.h, src/learner/bcd.h
 * ------------------------------------------------------------------------- */

/* Host-updated control words of one pass; lives in device memory so that a captured
 * CUDA graph can be replayed with new values (H2D copy before each replay). */
typedef struct psb_darlin_ctrl {
  double kkt_threshold;      /* ref: darlin.h:105-108 */
  double lambda;             /* L1 strength, penalty.lambda(0) */
  double eta;                /* learning_rate.alpha */
  double delta_max;          /* [PS.LM.delta_max_value] */
  uint64_t seq_base;         /* sequence number of the step before this pass' first step */
  int32_t reset_kkt_filter;  /* 1: refill the active sets at pass begin */
  int32_t n_steps;           /* steps in this pass */
  int32_t revive_on_reset;   /* 1: a reset also un-marks filtered weights (w=NaN -> 0); the
                                reference leaves them NaN forever, default 0 */
  int32_t server_cta_keys;   /* persistent kernel: steps with at most this many shard keys are
                                updated by one CTA (0 = default 4096); larger ones by the grid */
  int32_t flags;             /* bit 0: do not prefetch the next step's CSC ranges into L2 */
  int32_t reserved;
} psb_darlin_ctrl;

/* Result of EVALUATE_PROGRESS (ref: darlin.h:248-265,504-511; bcd.proto BCDProgress) */
typedef struct psb_darlin_progress {
  double loss;          /* sum_i log(1+exp(-z_i)) over this worker's rows */
  double penalty;       /* lambda * sum |w| over this server shard */
  double violation;     /* max KKT violation seen by this shard in the pass */
  uint64_t nnz_w;       /* non-zero, non-marked weights in this shard */
  uint64_t nnz_active;  /* popcount of the shard's active set */
  uint64_t seq;         /* sequence stamp, written last */
} psb_darlin_progress;

/* Worker side: the localized training matrix (CSC over local columns) plus routing. */
typedef struct psb_darlin_worker {
  const int64_t* col_ptr;   /* [n_cols+1] */
  const uint32_t* row_idx;  /* [nnz] */
  const float* sval;        /* [nnz] y_i*x_ij; NULL => binary features (value = y_i) */
  const float* y;           /* [n_rows] +-1 (used only when sval==NULL) */
  double* z;                /* [n_rows] margin y_i * <x_i,w>  (log of the reference's dual_) */
  double* w;                /* [n_cols] worker's cached weights */
  double* delta;            /* [n_cols] trust region */
  uint32_t* active;         /* bitmap over local columns */
  const int32_t* col_owner; /* [n_cols] owning server of each local column */
  const int64_t* col_off;   /* [n_cols] offset of that column inside the owner's shard */
  const int32_t* seg_col;   /* [n_seg] column of each work segment */
  const int64_t* seg_beg;   /* [n_seg] first nnz of the segment */
  const int64_t* col_seg;   /* [n_cols+1] first segment of each column */
  const int64_t* blk_col;   /* [n_blocks*2] local column range [lo,hi) of each feature block */
  const uint32_t* blk_srv_mask; /* [n_blocks] servers intersecting the block */
  double* part;             /* [2*max segments per step] scratch partial sums */
  int32_t* col_cnt;         /* [n_cols] scratch arrival counters (zero between launches) */
  uint32_t* done_ctr;       /* [4] scratch CTA arrival counters (zero between launches) */
  int64_t n_rows, n_cols;
  int32_t seg_len;          /* nnz per segment */
  int32_t rank, n_workers, n_servers;
  /* peer-mapped views of every server (index = server id) */
  double* land_g[PSB_MAX_PEERS];      /* landing G  [n_workers][land_cap] */
  double* land_u[PSB_MAX_PEERS];      /* landing U  [n_workers][land_cap] */
  int64_t land_cap[PSB_MAX_PEERS];
  uint64_t* push_flag[PSB_MAX_PEERS]; /* on the server: [n_workers] */
  const double* srv_w[PSB_MAX_PEERS]; /* server shard weights (NaN-marked when inactive) */
  const double* srv_dw[PSB_MAX_PEERS];/* server shard last step delta-w */
  const uint64_t* ver_flag;           /* local: [n_servers], written by the servers */
  const void* seg_meta;               /* [n_seg] int4 {col, len | first<<16 | single<<17, beg lo,
                                         beg hi}: one 16-byte load per segment (fused kernel) */
} psb_darlin_worker;

/* Server side: one key-range shard of the model. */
typedef struct psb_darlin_server {
  double* w;            /* [n_keys] NaN-marked when KKT-filtered */
  double* dw;           /* [n_keys] w_new - w_old of the last update of each coordinate */
  double* delta;        /* [n_keys] */
  uint32_t* active;     /* bitmap */
  double* land_g;       /* [n_workers][land_cap] */
  double* land_u;
  int64_t land_cap;
  const uint64_t* push_flag;   /* local: [n_workers], written by the workers */
  const int64_t* blk_rng;      /* [n_blocks*2] shard-relative [lo,hi) of each block (empty if none) */
  double* violation;           /* [1] */
  uint32_t* done_ctr;          /* [4] */
  int64_t n_keys;
  int32_t rank, n_workers, n_servers;
  uint64_t* ver_flag[PSB_MAX_PEERS];  /* on worker r: &ver_flag_r[rank] */
} psb_darlin_server;

/* Pass begin: violation=0 and, if ctrl->reset_kkt_filter, refill the active sets.
 * Either struct pointer may be NULL. ref: darlin.h:176-182,297-300 */
int psb_darlin_pass_begin(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                          const psb_darlin_ctrl* ctrl, void* stream);

/* K1 + push: gradient / curvature of the blocks order[step*bps .. +bps) written straight
 * into the owners' landing slots over peer memory, then push flags raised to seq.
 * ref: darlin.h:344-427 (computeAndPushGradient / computeGradient), C1 in SURVEY §2.9 */
int psb_darlin_grad_push(const psb_darlin_worker* wk, const psb_darlin_ctrl* ctrl,
                         const int32_t* order, int32_t step, int32_t blocks_per_step,
                         int32_t grid, void* stream);

/* K2+K3: wait for all workers' pushes, reduce landing slots in fixed worker order, proximal
 * Newton step with trust region + KKT filter, publish w/dw, raise version flags.
 * ref: kv_buffered_vector.h:41-79 + darlin.h:175-246 */
int psb_darlin_server_update(const psb_darlin_server* sv, const psb_darlin_ctrl* ctrl,
                             const int32_t* order, int32_t step, int32_t blocks_per_step,
                             int32_t grid, void* stream);

/* K4 + pull: wait for the owners' version flags, read w/dw of the step's columns from the
 * owner shards in peer memory, update cached w / delta / active set and the margins z.
 * ref: darlin.h:317-342,429-502 (pullAndUpdateDual / updateDual), C2 in SURVEY §2.9 */
int psb_darlin_pull_update(const psb_darlin_worker* wk, const psb_darlin_ctrl* ctrl,
                           const int32_t* order, int32_t step, int32_t blocks_per_step,
                           int32_t grid, void* stream);

/* K5: progress of this rank (worker loss + server penalty/nnz/active/violation), written to
 * `out_local` and to out_peer[r] (r < n_peers) when non-NULL; `scratch` holds
 * 4*grid doubles of block partials and 1 uint32 counter after them.
 * ref: darlin.h:248-265,504-511 */
int psb_darlin_evaluate(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                        const psb_darlin_ctrl* ctrl, psb_darlin_progress* out_local,
                        psb_darlin_progress* const* out_peer, int32_t n_peers, uint64_t seq,
                        double* scratch, int32_t grid, void* stream);
int64_t psb_darlin_evaluate_scratch_bytes(int32_t grid);

/* One whole pass (pass_begin, n_steps x {grad_push, server_update, pull_update}, evaluate)
 * enqueued on `stream`; this is the sequence bench/engine capture into a CUDA graph. */
int psb_darlin_enqueue_pass(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                            const psb_darlin_ctrl* ctrl, const int32_t* order,
                            int32_t n_steps, int32_t blocks_per_step, int32_t grid_worker,
                            int32_t grid_server, psb_darlin_progress* out_local,
                            psb_darlin_progress* const* out_peer, int32_t n_peers,
                            double* eval_scratch, int32_t grid_eval, void* stream);

/* Progress evaluation that allocates nothing (no counterpart in the reference, whose nodes are
 * ordinary CPU processes free to malloc).  Same result as psb_darlin_evaluate, but the scratch
 * (per-CTA partial sums + arrival counter) is a fixed array inside the library image instead of
 * a caller buffer.  Meant for callers that cannot allocate device memory: code running inside a
 * job whose HBM is already fully committed to its own pools, or a latency-sensitive thread that
 * must avoid the allocator lock and the device-wide synchronisation an allocation can imply.
 *   - limit: grid <= psb_darlin_evaluate_noalloc_max_grid() (1024 CTAs, 33 KB of scratch that
 *     every GPU loading the library holds until exit); larger grids are rejected with an error;
 *   - the first call on a device looks up the array's device address (one-time set-up with that
 *     device current); make that first call before starting any stream capture;
 *   - all calls on a device share the one array: calls that overlap on the GPU (different
 *     streams) race on it, so serialise them; nothing is written to peers (out_local only). */
int32_t psb_darlin_evaluate_noalloc_max_grid(void);
int psb_darlin_evaluate_noalloc(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                                const psb_darlin_ctrl* ctrl, psb_darlin_progress* out_local,
                                uint64_t seq, int32_t grid, void* stream);

/* Whole pass in ONE cooperative launch: pass_begin, every step and the evaluation run inside
 * a persistent kernel whose phases are separated by grid barriers; with one rank the shard
 * update is applied by the warp that completes a column's gradient.  `bar`: 4 zeroed uint32
 * owned by the caller.  blk_tab[b] = {first segment, #segments, first shard key, #shard keys,
 * server mask, 0}.  grid <= psb_darlin_fused_max_grid().  prof (optional): 5 uint64 that
 * accumulate ns spent in {phase 1, plan+prefetch, barrier A wait, phase 3, barrier C}. */
int psb_darlin_fused_max_grid(int32_t binary);
int psb_darlin_fused_pass(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                          const psb_darlin_ctrl* ctrl, const int32_t* order,
                          const int64_t* blk_tab, int32_t n_steps, int32_t blocks_per_step,
                          int32_t grid, uint32_t* bar, psb_darlin_progress* out_local,
                          psb_darlin_progress* const* out_peer, int32_t n_peers, double* scratch,
                          uint64_t* prof, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PSB200_H_ */
