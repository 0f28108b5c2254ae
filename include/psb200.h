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
  int32_t reserved0;
  int32_t flags;             /* persistent pass kernel only.  1: do not prefetch the next step's CSC
                              * ranges into L2; 4: fill CTAs one after the other with segments instead
                              * of dealing segments round robin over all CTAs. */
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
  /* fused multi-rank pass (replicated update, see darlin.cu): land2[q] = rank q's landing table
   * of 32-byte tagged entries [2 parities][n_workers][n_cols]; col_has[j] bit r = worker r holds
   * nonzeros of column j; stash = [2*max segments per step] phase-1 state parked for phase 3;
   * done_local[q] = last step rank q finished (written by q), done_peer[q] = &done_local of rank
   * q at index `rank` */
  char* land2[PSB_MAX_PEERS];
  const uint32_t* col_has;
  double* stash;
  const uint64_t* done_local;
  uint64_t* done_peer[PSB_MAX_PEERS];
  int32_t push_fp32;   /* 1: (g, u) are pushed rounded to fp32 in one 16-byte store (lossy push filter) */
  int32_t reserved1;
  double* acc;         /* fused one-rank pass: [2 parities][n_cols][2] accumulators of (g, u), all zero
                          between launches */
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
 *   - the array is chosen by the CURRENT device (cudaGetDevice), so the current device must be
 *     the one `stream` belongs to; otherwise the kernel would be handed another GPU's array;
 *   - all calls on a device share the one array: calls that overlap on the GPU (different
 *     streams) race on it, so serialise them; nothing is written to peers (out_local only). */
int32_t psb_darlin_evaluate_noalloc_max_grid(void);
int psb_darlin_evaluate_noalloc(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                                const psb_darlin_ctrl* ctrl, psb_darlin_progress* out_local,
                                uint64_t seq, int32_t grid, void* stream);

/* Whole pass in ONE cooperative launch: pass_begin, every step and the evaluation run inside
 * a persistent kernel whose phases are separated by grid barriers; with one rank the segments
 * accumulate (g, u) into worker.acc with fp64 reductions and every segment applies the column's
 * update itself after the barrier (worker.w is the replica of the shard, NaN-marked like it).  `bar`: 4 zeroed uint32
 * owned by the caller.  blk_tab[b] = {first segment, #segments, first shard key, #shard keys,
 * server mask, 0}.  grid <= psb_darlin_fused_max_grid().  prof (optional): 5 uint64 that
 * accumulate ns spent in {phase 1, plan+prefetch, barrier A wait, phase 3, barrier C}. */
/* Several ranks: every rank keeps a replica of (w, delta, active) for all columns; a push is one
 * tagged entry per column stored into the landing table of every rank that holds or owns the
 * column (one NVLink flight per step), phase 3 sums the holders' entries in worker order and
 * applies the update on every rank identically; the owner also writes its shard.  Needs
 * worker.land2 / col_has / stash / done_local / done_peer; worker.w must start as a copy of the
 * shard weights. */
int psb_darlin_fused_max_grid(int32_t binary);
int psb_darlin_fused_pass(const psb_darlin_worker* wk, const psb_darlin_server* sv,
                          const psb_darlin_ctrl* ctrl, const int32_t* order,
                          const int64_t* blk_tab, int32_t n_steps, int32_t blocks_per_step,
                          int32_t grid, uint32_t* bar, psb_darlin_progress* out_local,
                          psb_darlin_progress* const* out_peer, int32_t n_peers, double* scratch,
                          uint64_t* prof, void* stream);


/* The multi-rank pass of `n_ranks` ranks emulated on ONE GPU in one cooperative launch (CTAs
 * [r*grid_per_rank, (r+1)*grid_per_rank) are rank r; peers are pointers into the same device).
 * Ranks that wait on each other must not be separate launches on one GPU; this entry lets a
 * single-GPU box run the exact multi-GPU code path.  ranks_dev: device array of n_ranks records. */
typedef struct psb_darlin_rank_args {
  psb_darlin_worker wk;
  psb_darlin_server sv;
  const psb_darlin_ctrl* ctrl;
  const int32_t* order;
  const int64_t* blk_tab;
  uint32_t* bar;
  psb_darlin_progress* out_local;
  psb_darlin_progress* const* out_peer;
  double* scratch;
} psb_darlin_rank_args;
int psb_darlin_fused_pass_emulated(const psb_darlin_rank_args* ranks_dev, int32_t n_ranks,
                                   int32_t binary, int32_t n_steps, int32_t blocks_per_step,
                                   int32_t grid_per_rank, int32_t n_peers, void* stream);

/* ------------------------------------------------------------------------- *
 *  Asynchronous minibatch SGD (FTRL / AdaGrad / SGD) on a sharded hash store
 *  ref: src/app/linear_method/async_sgd.h, src/parameter/kv_store.h, src/learner/sgd.h
 * ------------------------------------------------------------------------- */

/* One minibatch on a worker: CSR over the minibatch's local columns (= its sorted unique keys),
 * the (owner, slot) of every column in the owners' tables, and the peer views needed to pull
 * weights and to push gradients.  A worker owns a ring of `depth` landing regions on every
 * server (depth = max_delay + 1): push `seq` goes to region seq % depth once the server
 * acknowledged push seq - depth. */
typedef struct psb_sgd_worker {
  const int64_t* row_ptr;      /* [n_rows+1] */
  const uint32_t* col;         /* [nnz] local column */
  const float* val;            /* [nnz] or NULL (binary features) */
  const float* y;              /* [n_rows] */
  int64_t n_rows, n_cols;
  const int32_t* col_owner;    /* [n_cols] */
  const int64_t* col_slot;     /* [n_cols] slot in the owner's table (-1: unknown key, weight 0) */
  const int64_t* owner_begin;  /* [n_servers+1] columns are sorted by key, so owners are segments */
  float* w_local;              /* [n_cols] pulled weights */
  float* g_local;              /* [n_cols] gradient */
  float* xw;                   /* [n_rows] margins (kept for AUC) */
  double* stats;               /* [3] += loss, correct predictions, rows */
  const float* tab_w[PSB_MAX_PEERS];    /* server s: weight array of its table */
  int32_t* land_slot[PSB_MAX_PEERS];    /* on server s, this worker's ring: [depth][land_cap] */
  float* land_val[PSB_MAX_PEERS];
  int8_t* land_q[PSB_MAX_PEERS];
  float* land_scale[PSB_MAX_PEERS];     /* [depth][land_cap/128+1] */
  int64_t* land_cnt[PSB_MAX_PEERS];     /* on server s: [depth] counts of this worker */
  uint64_t* land_flag[PSB_MAX_PEERS];   /* on server s: &flag[rank] */
  const uint64_t* ack;                  /* local: [n_servers], written by the servers */
  int64_t land_cap;
  int32_t rank, n_workers, n_servers, depth;
  int32_t loss;    /* LossConfig::Type: 1 SQUARE, 2 LOGIT, 4 SQUARE_HINGE */
  int32_t quant;   /* 0: fp32 push; 1: int8, 2: fp8 e4m3, both with one fp32 scale per 128 values */
  const int64_t* n_cols_dev;  /* not NULL: the column count is read from here (minibatches localised by
                                 psb_sgd_localize, whose sizes never reach the host); n_cols is then only
                                 an upper bound, and col may hold 0xffffffff = filtered key, skipped */
} psb_sgd_worker;

typedef struct psb_sgd_server {
  float* w;                  /* table arrays, [cap] */
  float* z;                  /* FTRL: z; AdaGrad: sum of squared gradients */
  float* n;                  /* FTRL: sqrt(n) */
  int64_t cap;
  const int32_t* land_slot;  /* [n_workers][depth][land_cap] */
  const float* land_val;
  const int8_t* land_q;
  const float* land_scale;   /* [n_workers][depth][land_cap/128+1] */
  const int64_t* land_cnt;   /* [n_workers][depth] */
  const uint64_t* land_flag; /* [n_workers] */
  uint64_t* ack_out[PSB_MAX_PEERS];  /* on worker w: &ack_w[rank] */
  int64_t land_cap;
  double* stats;             /* [3] += change of nnz(w), sum w^2, sum (w_new-w_old)^2 of touched keys */
  int32_t rank, n_workers, depth;
  int32_t algo;              /* 0 FTRL, 1 AdaGrad, 2 SGD */
  int32_t quant;
  float alpha, beta, l1, l2;
  float sgd_t;               /* sqrt(iteration) for the plain SGD step size */
  int32_t lr_constant;       /* learning_rate.type CONSTANT: eta = alpha (ref learning_rate.h:15-22);
                                0 = DECAY alpha / (x + beta) */
} psb_sgd_server;

/* table set-up and key resolution (K8a).  Capacities are powers of two. */
int psb_kv_table_clear(uint64_t* keys, int64_t cap, void* stream);
int psb_kv_lookup_insert(const uint64_t* keys, int64_t n, const int32_t* owner,
                         uint64_t* const* table_keys_host, const int64_t* cap_host, int32_t n_servers,
                         int64_t* slot_out, int32_t* overflow_dev, void* stream);
/* Whole minibatch localisation as one enqueue without a host synchronisation (sgd_localize.cu):
 * sort + run-length -> CountMin tail filter (sketch == NULL: none) -> kept keys `cols` -> local
 * column of every non-zero (`col`, 0xffffffff for a filtered key) -> owner segments -> table slots.
 * Device buffers of capacity nnz: sorted, uniq, cnt, flag, cols, col, col_owner, col_slot;
 * counts_dev[2] <- {unique keys, kept keys}; owner_begin[n_servers+1]; bounds_dev[n_servers-1]: first
 * key of servers 1..; *overflow_dev: 1 = a table is full, 2 = an owner segment exceeds land_cap. */
int64_t psb_sgd_localize_temp_bytes(int64_t max_nnz);
/* PSB_LOC_TIMING=1 (diagnosis; synchronises after every stage): mean ms per call of {sort, run-length,
 * sketch, select, columns, owner boundaries, table slots}; returns the calls averaged (0: off). */
int64_t psb_sgd_localize_timing(double* out7);
/* the two stages on their own (stage B touches the sketch and must run in minibatch order; stage A of
 * the next minibatch may overlap it on another stream, with its own sorted / uniq / cnt / temp) */
int psb_sgd_localize_sort(const uint64_t* keys, int64_t nnz, uint64_t* sorted, uint64_t* uniq, uint32_t* cnt,
                          int64_t* counts_dev, void* temp, int64_t temp_bytes, void* stream);
int psb_sgd_localize_finish(const uint64_t* keys, int64_t nnz, const uint64_t* uniq, const uint32_t* cnt,
                            uint8_t* flag, uint64_t* cols, uint32_t* col, int64_t* counts_dev,
                            uint8_t* sketch, int32_t cm_n, int32_t cm_k, int32_t freq,
                            const uint64_t* bounds_dev, int32_t n_servers, int64_t land_cap, int64_t* owner_begin,
                            uint64_t* const* table_keys_host, const int64_t* cap_host, int32_t* col_owner,
                            int64_t* col_slot, int32_t* overflow_dev, void* temp, int64_t temp_bytes, void* stream);
int psb_sgd_localize(const uint64_t* keys, int64_t nnz, uint64_t* sorted, uint64_t* uniq, uint32_t* cnt,
                     uint8_t* flag, uint64_t* cols, uint32_t* col, int64_t* counts_dev,
                     uint8_t* sketch, int32_t cm_n, int32_t cm_k, int32_t freq,
                     const uint64_t* bounds_dev, int32_t n_servers, int64_t land_cap, int64_t* owner_begin,
                     uint64_t* const* table_keys_host, const int64_t* cap_host, int32_t* col_owner,
                     int64_t* col_slot, int32_t* overflow_dev, void* temp, int64_t temp_bytes, void* stream);
int psb_kv_export_nonzero(const uint64_t* keys, const float* w, int64_t cap, uint64_t* out_keys,
                          float* out_w, uint64_t* count_dev, void* stream);
/* K6+K7+C3+C4: pull, margins / loss / gradient, push of one minibatch in one cooperative launch.
 * `bar`: one zeroed uint32.  seq starts at 1 and increases by one per step. */
/* Waits on peer flags give up after seconds and record {code, index, wanted, seen} here
 * (4 zeroed uint64 in pinned host memory; NULL disables) instead of hanging the GPU. */
int psb_sgd_set_fault_sink(uint64_t* host_pinned);
int psb_sgd_max_grid(void);
int psb_sgd_worker_step(const psb_sgd_worker* wk, uint64_t seq, int32_t grid, uint32_t* bar,
                        void* stream);
/* K8: wait (on the device) for push `seq` of every worker, apply them in worker order, then
 * acknowledge to the workers. */
int psb_sgd_server_update(const psb_sgd_server* sv, uint64_t seq, int32_t grid, void* stream);
/* The same in ONE cooperative launch (workers in rank order behind grid barriers, then the
 * acknowledgements): 2 launches per minibatch step instead of 2 + n_workers.  bar: zeroed uint32. */
int psb_sgd_server_update_fused(const psb_sgd_server* sv, uint64_t seq, int32_t grid, uint32_t* bar,
                                void* stream);


/* Streaming runner (sgd_stream.cu): the whole minibatch loop of a worker + its server in native
 * code.  Per minibatch, on one of the runner's copy streams (round robin): H2D of the host arrays into
 * one set of a ring of device buffers, stage A of the localisation, then - after stage B of the previous
 * minibatch on another stream - stage B; on the step stream: worker step, server update.  Events
 * order a set's reuse after the step that read it; the host never waits.  All device memory belongs to the
 * caller (who keeps it alive until psb_sgd_runner_destroy). */
#define PSB_SGD_MAX_RING 12
#define PSB_SGD_MAX_LANES 8
typedef struct psb_sgd_host_minibatch {  /* pinned host memory; row_ptr starts at 0 */
  const int64_t* row_ptr;
  const uint64_t* keys;
  const float* vals;   /* NULL: binary */
  const float* y;
  int64_t n_rows, nnz; /* 0, 0: this rank ran out of data and pushes nothing */
} psb_sgd_host_minibatch;
typedef struct psb_sgd_ring_set {
  int64_t* row_ptr;    /* [max_rows+1] */
  uint64_t* keys;      /* [max_nnz] */
  float* vals;         /* [max_nnz] or NULL */
  float* y;            /* [max_rows] */
  uint64_t* cols;      /* [max_nnz] localisation outputs ... */
  uint32_t* col;
  int32_t* col_owner;
  int64_t* col_slot;
  int64_t* counts;     /* [2] */
  int64_t* owner_begin;/* [n_servers+1] */
} psb_sgd_ring_set;
typedef struct psb_sgd_scratch {  /* work arrays of one localisation in flight */
  uint64_t* sorted;    /* [max_nnz] */
  uint64_t* uniq;      /* [max_nnz] */
  uint32_t* cnt;       /* [max_nnz] */
  uint8_t* flag;       /* [max_nnz] */
  void* temp;          /* [temp_bytes] psb_sgd_localize_temp_bytes(max_nnz) */
} psb_sgd_scratch;
typedef struct psb_sgd_runner_config {
  psb_sgd_worker worker;  /* peer views, w_local / g_local [max_nnz], xw [max_rows], stats, loss, quant */
  psb_sgd_server server;
  psb_sgd_ring_set set[PSB_SGD_MAX_RING];
  int32_t n_sets;
  int32_t fused_server;
  int64_t max_rows, max_nnz;
  psb_sgd_scratch scratch[PSB_SGD_MAX_LANES];  /* minibatch k localises with scratch[k % n_lanes] on copy_stream[k % n_lanes] */
  int64_t temp_bytes;
  uint8_t* sketch;     /* tail filter (NULL: none) */
  int32_t cm_n, cm_k, freq, reserved;
  const uint64_t* bounds_dev;
  uint64_t* table_keys[PSB_MAX_PEERS];
  int64_t table_cap[PSB_MAX_PEERS];
  int32_t* overflow_dev;
  int32_t* overflow_host;  /* pinned: receives *overflow_dev at the end of every submit */
  uint32_t* worker_bar;
  uint32_t* server_bar;
  int32_t grid_worker, grid_server;
  void* copy_stream[PSB_SGD_MAX_LANES];  /* several, so that upload and sort of one minibatch overlap the rest of those before */
  int32_t n_lanes;       /* 1 .. PSB_SGD_MAX_LANES, and < n_sets */
  int32_t reserved2;
} psb_sgd_runner_config;
/* returns NULL on failure (psb_last_error) */
void* psb_sgd_runner_create(const psb_sgd_runner_config* cfg);
void psb_sgd_runner_destroy(void* runner);
/* enqueue n minibatch steps with sequence numbers seq_first, seq_first + 1, ...; returns once
 * everything is enqueued.  The host arrays must stay valid until the copy stream has passed. */
/* PSB_SGD_TIMING=1 in the environment at creation: mean ms per step of {H2D, localise, worker kernel,
 * server kernel, copy stage, step stage, step period}; returns the steps averaged (0: off). */
int psb_sgd_runner_timing(void* runner, double* out7);
int psb_sgd_runner_submit(void* runner, const psb_sgd_host_minibatch* mb, int32_t n, uint64_t seq_first,
                          void* step_stream);

/* ------------------------------------------------------------------------- *
 *  Evaluation, sketches, codecs, fills (eval.cu)
 * ------------------------------------------------------------------------- */
/* K14 exact AUC (ref src/util/evaluation.h:22-48): sort by prediction, area = sum over negatives of
 * the positives ranked below.  work_f: float[2n], work_i: int64[n], out_dev: double[2] zeroed by the
 * caller -> {area, #positives}; AUC = area / (pos * (n - pos)), flipped when < 0.5. CUB temp convention. */
int psb_auc_exact(const float* label, const float* predict, int64_t n, float* work_f, int64_t* work_i,
                  double* out_dev, void* temp, int64_t* temp_bytes, void* stream);
/* out_dev: double[2] zeroed by the caller -> {#correct, total logloss} (ref evaluation.h:52-63) */
int psb_accuracy_logloss(const float* label, const float* predict, int64_t n, float threshold,
                         double* out_dev, void* stream);
/* K13 CountMin with saturating 8-bit counters, hash and probe sequence identical to the host
 * sketch (ref src/util/countmin.h:21-41, sketch.h:20-33, frequency_filter.h:26-43) */
int psb_countmin_insert(const uint64_t* keys, const uint32_t* counts, int64_t n, uint8_t* table,
                        int32_t cm_n, int32_t cm_k, void* stream);
int psb_countmin_query(const uint64_t* keys, int64_t n, const uint8_t* table, int32_t cm_n, int32_t cm_k,
                       int32_t freq, uint8_t* keep_out, void* stream);
/* K9 FIXING_FLOAT wire codec (ref src/filter/fixing_float.h:44-95), stochastic rounding */
int psb_fixfloat_encode_f32(const float* in, int64_t n, int32_t nbytes, double min_v, double max_v,
                            uint64_t seed, uint8_t* code, void* stream);
int psb_fixfloat_decode_f32(const uint8_t* code, int64_t n, int32_t nbytes, double min_v, double max_v,
                            float* out, void* stream);
int psb_fixfloat_encode_f64(const double* in, int64_t n, int32_t nbytes, double min_v, double max_v,
                            uint64_t seed, uint8_t* code, void* stream);
int psb_fixfloat_decode_f64(const uint8_t* code, int64_t n, int32_t nbytes, double min_v, double max_v,
                            double* out, void* stream);
/* K19 ParameterInitConfig: type 1 ZERO, 2 CONSTANT, 3 GAUSSIAN (ref shared_array_inl.h:103-130) */
int psb_fill_f64(double* p, int64_t n, int32_t type, double constant, double mean, double std,
                 uint64_t seed, void* stream);
/* K18 popcount of a bitmap; *out_dev zeroed by the caller (ref src/util/bitmap.h:35-75) */
int psb_bitmap_popcount(const uint32_t* words, int64_t n_words, uint64_t* out_dev, void* stream);
/* K6/K17 margins of CSR rows with 64-bit keys against a sorted (key, weight) model
 * (ref src/app/linear_method/model_evaluation.h:40-66) */
int psb_spmv_lookup(const int64_t* row_ptr, const uint64_t* keys, const float* val, int64_t n_rows,
                    const uint64_t* dict, const double* w, int64_t n_dict, float* out, void* stream);


/* ------------------------------------------------------------------------- *
 *  Generic sharded KV push / pull over peer memory (kv.cu)
 *  ref: SharedParameter::push/pull, KVVector::setValue/getValue, sliceKeyOrderedMsg
 * ------------------------------------------------------------------------- */
/* (owner, slot) of keys whose dictionary positions are known (idx1 = 1 + position, 0 = unknown);
 * shard_begin_dev[s] = first dictionary position of server s */
int psb_kv_route_dense(const uint32_t* idx1, int64_t n, const int64_t* shard_begin_dev,
                       int32_t n_servers, int32_t* owner, int64_t* slot, void* stream);
/* push: vals[i] is added into server owner[i]'s array at slot[i] (reductions over NVLink, 128-bit
 * wide on contiguous runs); pull: the reverse gather.  peers_host[s] = mapping of server s' array */
int psb_kv_push_add_f32(const float* vals, const int32_t* owner, const int64_t* slot, int64_t n,
                        float* const* peers_host, int32_t n_servers, void* stream);
int psb_kv_pull_f32(float* out, const int32_t* owner, const int64_t* slot, int64_t n,
                    float* const* peers_host, int32_t n_servers, void* stream);
/* bulk (bandwidth-bound) path: only contiguous value segments cross NVLink; the owner applies /
 * gathers locally with the slot lists it cached when the handle was registered */
int psb_kv_bulk_send(const float* src, const int64_t* seg_begin_host, float* const* dst_host,
                     uint64_t* const* flag_host, int32_t n_servers, uint64_t seq, uint32_t* done_ctr,
                     void* stream);
int psb_kv_bulk_apply_add(float* val, const float* landing, const int64_t* slots, int64_t n,
                          const uint64_t* flag, uint64_t seq, void* stream);
int psb_kv_bulk_gather_send(const float* val, const int64_t* const* slots_host, const int64_t* count_host,
                            float* const* dst_host, uint64_t* const* flag_host, int32_t n_workers,
                            uint64_t seq, uint32_t* done_ctr, void* stream);
int psb_kv_wait_flags(const uint64_t* flags, int32_t n, uint64_t seq, void* stream);
/* receiver-driven bulk push: ONE cooperative launch per owner reads every worker's segment from that
 * worker's staging buffer in peer memory (128-bit loads, workers in a staggered order, a grid barrier
 * between two workers: deterministic) and adds it into `val` at 32-bit slots; bar: zeroed uint32 */
int psb_kv_bulk_pull_apply(float* val, const float* const* src_host, const int32_t* const* slots_host,
                           const int64_t* count_host, int32_t n_workers, int32_t rank, uint32_t* bar,
                           void* stream);
/* bulk pull with 32-bit slots, 128-bit peer stores and staggered worker order */
int psb_kv_bulk_gather_send32(const float* val, const int32_t* const* slots_host, const int64_t* count_host,
                              float* const* dst_host, uint64_t* const* flag_host, int32_t n_workers,
                              int32_t rank, uint64_t seq, uint32_t* done_ctr, void* stream);


/* ------------------------------------------------------------------------- *
 *  Device storage of the public KV containers (devkv.cu), used by the host runtime's device
 *  backend (csrc/host/parameter/kv.h) through dlopen.
 *  ref: KVVector::setValue/getValue (kv_vector.h:46-93), KVMap (kv_map.h:4-45),
 *       compAssOp (shared_parameter.h:173-193)
 * ------------------------------------------------------------------------- */
enum { PSB_DT_F32 = 0, PSB_DT_F64 = 1, PSB_DT_I32 = 2, PSB_DT_I64 = 3, PSB_DT_U8 = 4 };
/* ASSIGN plus the reference's Operator enum (param.proto) */
enum { PSB_OP_ASSIGN = 0, PSB_OP_PLUS = 1, PSB_OP_MINUS = 2, PSB_OP_TIMES = 3, PSB_OP_DIVIDE = 4,
       PSB_OP_AND = 5, PSB_OP_OR = 6, PSB_OP_XOR = 7 };
int psb_device_count(int32_t* out);
int psb_set_device(int32_t dev);
int psb_stream_create(void** out);
int psb_stream_destroy(void* stream);
/* dst[pos[i]*k + c] (op)= src[i*k + c]; pos NULL = contiguous, pos[i] < 0 skipped.  `src` may be a
 * CUDA-IPC mapping of the sender's staging arena: the payload is read over NVLink while applied. */
int psb_dkv_apply(void* dst, const int64_t* pos, const void* src, int64_t n, int32_t k,
                  int32_t dtype, int32_t op, void* stream);
/* dst[i] = src[pos[i]] for entries of entry_bytes bytes (dst may be a peer mapping); pos[i] < 0 -> 0 */
int psb_dkv_gather(void* dst, const void* src, const int64_t* pos, int64_t n, int32_t entry_bytes,
                   void* stream);
/* pos[i] = index of keys[i] in the sorted unique my_keys, or -1 */
int psb_dkv_positions(const uint64_t* my_keys, int64_t n_my, const uint64_t* keys, int64_t n,
                      int64_t* pos, void* stream);
/* hash map (KVMap): open-addressing key table, capacity a power of two; values live in a caller
 * array indexed by slot */
int psb_dmap_clear(uint64_t* table_keys, int64_t cap, void* stream);
int psb_dmap_slots(uint64_t* table_keys, int64_t cap, const uint64_t* keys, int64_t n, int32_t insert,
                   int64_t* slot, int32_t* overflow_dev, void* stream);
int psb_dmap_export(const uint64_t* table_keys, int64_t cap, uint64_t* out_keys, int64_t* out_slot,
                    uint64_t* count_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PSB200_H_ */
