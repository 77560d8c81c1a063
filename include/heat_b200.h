/* include/heat_b200.h -- C ABI of the B200-native HEAT MF_CCL hot path.
 *
 * Drop-in boundary: the reference sits behind the pybind11 module `cf_c`
 * (cf_cpu/src/pybind/init_modules.cpp:13-152).  The host C++ classes that module binds
 * (heat_b200/csrc/host/) keep the reference's names and semantics and reach the GPU only
 * through the entry points declared here: plain pointers and sizes, no torch / pybind /
 * CUDA types.  Each entry point names the reference interface it replaces.
 *
 * Conventions: every function returns 0 on success, non-zero on failure;
 * heat_last_error() returns a thread-local message for the last failure.  A context is
 * bound to one CUDA device (one process per GPU); calls on one context are not
 * thread-safe.  There is NO CPU fallback: without a usable CUDA device heat_create fails.
 */
#ifndef HEAT_B200_H
#define HEAT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HEAT_ABI_VERSION 2

typedef struct heat_ctx heat_ctx;

/* Mirrors cf::modules::CFConfig (cf_cpu/src/modules/cf_config.hpp:12-35; python kwargs at
 * init_modules.cpp:17-30) plus the knobs the GPU path adds.  Zero in an extension field
 * selects the documented default, so a memset-0 struct with the reference fields filled
 * in is a valid configuration. */
typedef struct heat_config {
    /* --- reference fields --- */
    uint64_t emb_dim;           /* D: 32, 64 or 128 */
    uint64_t num_negs;          /* K */
    uint64_t num_users;         /* rows of the (local) user table */
    uint64_t num_items;
    uint64_t train_size;        /* N local samples */
    uint64_t neg_sampler;       /* 0 uniform, 1 random-tile: in the reference's shipped loop both draw uniformly
                                   (SURVEY 8.1 a4), and so they do here unless tile_sampler (below) is set */
    uint64_t tile_size;
    uint64_t refresh_interval;
    uint64_t num_subepochs;     /* fork-only; accepted and ignored */
    float l2;                   /* read, never applied (matrix_factorization.cpp:126,146,165,168) */
    float clip_val;
    float l_r;
    const uint64_t *milestones; /* copied */
    uint64_t n_milestones;
    /* --- extensions --- */
    uint64_t seed;              /* Philox key; default 2022 (the YAMLs' seed) */
    uint64_t sample_offset;     /* global index of local sample 0 (multi-GPU shards) */
    uint32_t chunk;             /* samples a warp takes per scheduling step; default 64
                                   (the reference's omp schedule(dynamic,512), engine.cpp:327) */
    uint32_t deterministic;     /* 1: one warp walks the samples in file order (== 1 OpenMP thread) */
    uint32_t w0_tail_mode;      /* 0 default: W0 grads left over in a thread block's ring at launch end (< 32 samples) are
                                   applied (scaled 1/32); 1: dropped like the reference's per-thread aggregator at epoch
                                   end (behavior_aggregators.cpp:141-146).  A multi-GPU epoch is several launches (one per
                                   sync segment), so the leftover is flushed (or dropped) once per segment and block */
    uint32_t warps_per_block;   /* 0 = tuned default */
    uint32_t blocks_per_sm;     /* 0 = tuned default */
    uint32_t update_mode;       /* item-row write-back in the Hogwild grid of the warp-per-sample kernel: 0 = default = 1: the grad rows
                                   are read right before the step and the weight rows stepped by L2 reductions (short
                                   read-modify-write window, no lost weight step); 2: grad and weight rows both by L2
                                   reductions; 3: the stepped snapshot is stored (the reference's own sequence; loses updates
                                   under thousands of concurrent samples).  deterministic=1 forces 3 */
    uint32_t sync_average;      /* multi-GPU item/W0 reconciliation: 0 = default = 1: every replica becomes the MEAN of the
                                   replicas (== base + mean of the rank deltas: the fork's averaging, engine.cpp:354-375);
                                   2: base + SUM of the rank deltas.  The mean keeps Recall@20 on the one-GPU run's for every
                                   exchange interval from 1/74 to 1/10 of an epoch; the sum does not */
    uint32_t history_mode;      /* 0 default: the history column sums are gathered once per run of a user's
                                   consecutive samples and patched with the exact change of the rows the warp
                                   rewrites; 1: re-gathered for every sample (the reference's literal sequence) */
    uint32_t engine;            /* 0 auto (see heat_last_engine): the warp-per-sample kernel on its full grid where it covers
                                   (emb_dim, num_negs) -- Recall@20 within 0.005 of the reference, per-epoch loss trailing it in the
                                   first epochs; 1 the same kernel, forced; 2 CTA-per-sample kernel
                                   (train_kernel_cta.cuh): one sample in flight per thread block, per-epoch loss within 0.03 AND
                                   Recall@20 within 0.005 of the reference -- the strict choice; 3 large-negative kernel */
    uint32_t max_blocks;        /* cap on thread blocks = bound on Hogwild concurrency.  0 = default: the warp-per-sample kernel
                                   runs one block per SM, at most one warp in flight per 32 item rows; the CTA kernel (engine 2)
                                   min(#SMs, max(8, num_items/600)) blocks */
    uint32_t exchange_mode;     /* multi-GPU item/W0 reconciliation: 0 auto -- ONE kernel per rank that reads every rank's
                                   replica over NVLink peer memory (CUDA IPC), reduces its slice of the rows and writes
                                   the result back into every replica (comm.cu); NCCL all-reduce when peer mapping is
                                   not available; 1 = always NCCL (delta kernel + ncclAllReduce + apply kernel) */
    uint32_t tile_sampler;      /* 1 with neg_sampler 1: hot-tile policy (random_tile_negative_sampler.cpp:22-45, paper
                                   Alg. 1): the samples of one window of refresh_interval samples draw their negatives
                                   from a Philox-defined tile of tile_size item ids (philox.cuh).  0 (default): as the
                                   reference's shipped loop, the tile is ignored */
    uint32_t sync_grads;        /* multi-GPU: 1 = the persistent item grad rows are reconciled together with the item table
                                   (g = clamp(g_base + sum of rank deltas)), so that every rank steps with the running sum of
                                   ALL ranks' touches, as one GPU training the union of the shards would; 0 = grad rows stay
                                   local to a rank.  Only matters with more than one exchange per epoch (the rows are zeroed
                                   at every epoch end) */
    uint32_t eval_precision;    /* fused evaluation (heat_eval_topk / heat_eval_metrics): 0 = default = 3: three TF32 passes per tile
                                   (A_lo.B_hi + A_hi.B_lo + A_hi.B_hi: fp32-level scores, 2e-6 relative); 1: one TF32 pass (1e-3 relative
                                   scores; the top-50 sets differ in a few near-ties).  heat_eval_dense always uses 3 */
} heat_config;

/* ---- life cycle -------------------------------------------------------------------- */
int heat_abi_version(void);
const char *heat_last_error(void);
/* replaces: cf_c.modules.CFConfig(...) + Engine ctor (engine.cpp:9-20) */
int heat_create(heat_ctx **out, const heat_config *cfg, int device);
int heat_destroy(heat_ctx *ctx);

/* ---- inputs ------------------------------------------------------------------------ */
/* replaces: cf_c.modules.datasets.ClickDataset(click_dataset, historical_items, masks)
 * (init_modules.cpp:45-61; dataset.cpp:10-19).  Host uint64 arrays in the reference's
 * layout: clicks [n,2], his_items [num_users,max_his], masks [num_users,1].  Copied to the
 * device and narrowed to int32 there; the host arrays are not retained. */
int heat_set_dataset(heat_ctx *ctx, const uint64_t *clicks, uint64_t n,
                     const uint64_t *his_items, uint64_t num_users, uint64_t max_his,
                     const uint64_t *masks);
/* replaces: cf_c.modules.models.MatrixFactorization(cf_config, user_weights, item_weights)
 * and behavior_aggregators.AggregatorWeights(aggregator_weights0) (init_modules.cpp:74-84,
 * 95-98).  The reference aliases these host buffers and trains them in place; here they
 * are uploaded, the pointers are retained, and heat_sync_to_host writes the trained
 * tables back into them. */
int heat_bind_tables(heat_ctx *ctx, float *user_w, float *item_w, float *w0);
int heat_upload_tables(heat_ctx *ctx);   /* host -> device again (after the caller edited them) */
int heat_sync_to_host(heat_ctx *ctx);    /* device -> the bound host buffers */

/* ---- training ---------------------------------------------------------------------- */
/* replaces: Engine::train_one_epoch() (upstream loop, engine.cpp:156-160,293-347,378-385):
 * LR schedule, all N local samples in file order through the fused kernel, zero_grad,
 * epoch += 1.  *mean_loss = sum(loss) / N. */
int heat_train_epoch(heat_ctx *ctx, double *mean_loss);
/* The same kernel over local samples [begin,end) without the epoch bookkeeping (no LR
 * schedule, no zero_grad, epoch counter unchanged).  loss_per_sample may be NULL or a
 * host array of (end-begin) floats. */
int heat_train_range(heat_ctx *ctx, uint64_t begin, uint64_t end, double *loss_sum,
                     float *loss_per_sample);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same contract and same
 * kernels as heat_train_range, but the call allocates nothing: the chunk counter, the loss sum and the per-sample loss
 * staging live in ONE file-scope `static __device__` array of the library (train_launch.cu; 256 B +
 * 4 B x HEAT_NOALLOC_MAX_SAMPLES = 4 MiB held by every GPU that loads the library until exit).  A range of more than
 * HEAT_NOALLOC_MAX_SAMPLES samples is refused with an error (split it).  All calls on one device share the array, so two
 * calls that overlap on the GPU (different contexts / host threads) race; the words that must start at zero are cleared
 * by every call.  The first call on a device looks the array's device address up (cudaGetSymbolAddress). */
#define HEAT_NOALLOC_MAX_SAMPLES (1u << 20)
int heat_train_range_noalloc(heat_ctx *ctx, uint64_t begin, uint64_t end, double *loss_sum,
                             float *loss_per_sample);
int heat_zero_grad(heat_ctx *ctx);                       /* embedding.cpp:41-45 */
int heat_get_grads(heat_ctx *ctx, float *user_g, float *item_g);   /* either may be NULL */
int heat_set_lr(heat_ctx *ctx, float l_r);
int heat_get_lr(heat_ctx *ctx, float *l_r);
int heat_get_epoch(heat_ctx *ctx, uint64_t *epoch);
int heat_set_epoch(heat_ctx *ctx, uint64_t epoch);
/* device time of the last heat_train_epoch / heat_train_range kernel, CUDA events, ms */
int heat_last_kernel_ms(heat_ctx *ctx, float *ms);
/* device time of the whole last heat_train_epoch (kernel(s), zero_grad, item syncs), CUDA events, ms */
int heat_last_epoch_ms(heat_ctx *ctx, float *ms);
/* kernel family the last heat_train_epoch / heat_train_range ran: 1 warp-per-sample (train_kernel.cuh), 2 CTA-per-sample
 * (train_kernel_cta.cuh), 3 large-negative (train_kernel_bigk.cuh); 0 before the first launch.  With engine 0 the
 * library picks: deterministic -> 2 (1 where 2 does not cover the shape); large num_negs -> 3; otherwise 1. */
int heat_last_engine(heat_ctx *ctx, int *engine);
/* kernels launched by this context since creation (bench.py's gpu_launches) */
int heat_launch_count(heat_ctx *ctx, uint64_t *n);

/* ---- dataset ingest (host code) ----------------------------------
 * replaces: the per-line Python parser of cf_cpu/cf/datasets.py:31-40.  LightGCN format, one line per
 * user: `user item item ...` (fields separated by `separator` and/or blanks, empty lines skipped).
 * heat_clickfile_open parses the whole file with one thread per core and reports the row and
 * interaction counts; heat_clickfile_read fills caller arrays in FILE order: row_users [num_rows],
 * row_indptr [num_rows+1], items [nnz]; heat_clickfile_close frees the handle. */
typedef struct heat_clickfile heat_clickfile;
int heat_clickfile_open(const char *path, char separator, heat_clickfile **out, uint64_t *num_rows, uint64_t *nnz);
int heat_clickfile_read(heat_clickfile *f, int64_t *row_users, int64_t *row_indptr, int64_t *items);
int heat_clickfile_close(heat_clickfile *f);

/* ---- checkpoint / resume (the reference has none: a killed job restarts) ----
 * One binary file: a 64-byte header (magic "HEATB200", version, emb_dim, num_users, num_items,
 * epoch, scheduled LR) followed by user_w, item_w, w0, user_g, item_g as raw fp32 in that order.
 * heat_save_state reads the DEVICE state (no heat_sync_to_host needed); heat_load_state checks the
 * shapes against the context's config, restores tables, grad rows, epoch and LR on the device
 * and also refreshes the bound host buffers.  Training resumed from a file continues exactly
 * as the uninterrupted run would (the Philox counter is a function of epoch and sample index). */
int heat_save_state(heat_ctx *ctx, const char *path);
int heat_load_state(heat_ctx *ctx, const char *path);

/* ---- index contracts (bit-exact tests) --------------------------------------------- */
/* Negative ids the device sampler draws for local samples [begin,end) of `epoch`:
 * out[(s-begin)*K + k].  replaces: UniformRandomNegativeSampler::ignore_pos_sampling
 * (uniform_random_negative_sampler.cpp:26-36). */
int heat_sample_negatives(heat_ctx *ctx, uint64_t epoch, uint64_t begin, uint64_t end, uint64_t *out);
/* Item ids the history aggregation gathers for local samples [begin,end):
 * out_ids[(s-begin)*max_his + j] for j < out_count[s-begin], rest = UINT64_MAX.
 * replaces: BehaviorAggregator::forward index reads (behavior_aggregators.cpp:60-62,96-99). */
int heat_history_indices(heat_ctx *ctx, uint64_t begin, uint64_t end, uint64_t *out_ids, uint64_t *out_count);
/* Host twin of the sampler (no GPU needed): one raw draw / one slot after look-back. */
uint64_t heat_philox_draw(uint64_t seed, uint64_t epoch, uint64_t sample, uint32_t k, uint64_t num_items);
/* The same with the hot-tile mapping (tile_size 0 = uniform): what the device draws when tile_sampler is on. */
uint64_t heat_tile_draw(uint64_t seed, uint64_t epoch, uint64_t sample, uint32_t k, uint64_t num_items,
                        uint64_t tile_size, uint64_t refresh_interval);

/* ---- evaluation -------------------------------------------------------------------- */
/* replaces: Engine::evaluate0() + the copying binding lambda (engine.cpp:388-400,
 * init_modules.cpp:122-129): out[U*I] = U . I^T (row-major, fp32 result). */
int heat_eval_dense(heat_ctx *ctx, float *out);
/* replaces: metrics.evaluate_metrics masking + argpartition/argsort (cf/metrics.py:21-29):
 * score GEMM fused with a per-user masked top-k; the score matrix is never stored.
 * train_indptr [U+1] / train_items: CSR of items to mask per user (may be NULL = no mask).
 * topk_ids [U*k] int32 (sorted by score desc, ties to the smaller id), topk_scores [U*k]. */
int heat_eval_topk(heat_ctx *ctx, uint32_t k, const int64_t *train_indptr, const int32_t *train_items,
                   int32_t *topk_ids, float *topk_scores);
/* replaces: Recall / NDCG / HitRate (cf/metrics.py:39-48,87-113,131-139) at k = 20 and 50,
 * averaged over users with at least one test item.  out[6] = Recall@20, Recall@50,
 * NDCG@20, NDCG@50, HitRate@20, HitRate@50; sums_out (may be NULL) receives the six
 * un-normalised sums followed by the evaluated-user count (for multi-GPU reduction). */
int heat_eval_metrics(heat_ctx *ctx, const int64_t *train_indptr, const int32_t *train_items,
                      const int64_t *test_indptr, const int32_t *test_items,
                      double out[6], double sums_out[7]);
int heat_last_eval_ms(heat_ctx *ctx, float *ms);

/* ---- multi-GPU (one process per GPU; NCCL over NVLink) -------------------------------- */
/* 128-byte ncclUniqueId made by rank 0 and handed to every rank by the launcher. */
int heat_comm_unique_id(uint8_t out[128]);
int heat_comm_init(heat_ctx *ctx, int rank, int world, const uint8_t unique_id[128]);
/* replaces: the fork's per-row MPI_Bcast / MPI_Allreduce of item rows and the W0 average
 * (engine.cpp:148-152,262-286,354-375): one all-reduce of the item-table delta and one of
 * the W0 delta since the last sync; every rank ends with base + sum of deltas. */
int heat_comm_sync_items(heat_ctx *ctx);
/* samples a rank trains between two exchanges.  0 (the initial value) = auto: max(65536, 8 x num_items) -- every item row
 * is touched ~140 times per replica between exchanges, 4 exchanges per epoch at the Amazon-Books shape (Recall@20 does not
 * move between 3 and 74 exchanges per epoch, one per epoch is too few: DESIGN.md section 7); any value >= the shard size =
 * once per epoch.  Every rank must set the same value. */
int heat_set_sync_interval(heat_ctx *ctx, uint64_t samples);
/* exchanges of the last heat_train_epoch and their total device time (CUDA events on the context's stream), ms */
int heat_last_exchange(heat_ctx *ctx, uint64_t *count, float *ms);
/* Multi-rank emulation on ONE device, for convergence tests and interval studies on a single-GPU box: n contexts
 * created on the same device, context r holding rank r's user shard and its own replica of the item table and W0, train
 * one epoch exactly as n ranks would -- every rank's segment, then the reconciliation of heat_comm_sync_items (same
 * arithmetic, same rank order, same sync_average / sync_grads / sync interval, taken from ctxs[0]) -- only one rank after
 * the other instead of side by side.  heat_local_group_init snapshots the base copies (call it after heat_bind_tables).
 * *mean_loss = sum of all ranks' losses / all ranks' samples (what rank 0 of a real run returns). */
int heat_local_group_init(heat_ctx **ctxs, int n);
int heat_local_group_train_epoch(heat_ctx **ctxs, int n, double *mean_loss);
/* Experiment knob of the emulation group (what would a finer-grained exchange buy?): the members SHARE one physical
 * array instead of reconciling copies -- mask bit 0: W0, bit 1: the item grad rows, bit 2: the item table.  Call once,
 * right after heat_local_group_init. */
int heat_local_group_share(heat_ctx **ctxs, int n, uint32_t mask);
/* how heat_comm_sync_items exchanges: 0 no communicator, 1 NCCL all-reduce, 2 fused peer-memory kernel */
int heat_comm_mode(heat_ctx *ctx, int *mode);
/* replaces: MPI_Allreduce of loss / iteration count (engine.cpp:380-383) for caller-side sums, e.g. the
 * seven metric sums of heat_eval_metrics on user-sharded evaluation: vals[n] (n <= 16) summed over ranks. */
int heat_comm_allreduce_f64(heat_ctx *ctx, double *vals, int n);

/* ---- measurement helpers ------------------------------------------------------------ */
/* L2-resident row gather/scatter micro-benchmark used as the L2 roofline denominator:
 * random 256-byte rows of a `bytes`-sized buffer are read and rewritten. GB/s (read+write). */
int heat_probe_l2_bandwidth(int device, uint64_t bytes, int iters, double *gbps);
/* Dense TF32 tensor peak of this GPU, measured: nothing but back-to-back tcgen05.mma.kind::tf32 (M128 N256 K8, operands in
 * shared memory, accumulators in TMEM) on every SM; the roofline denominator of the evaluation kernel. TFLOP/s. */
int heat_probe_tf32_peak(int device, int iters, double *tflops);
/* page-locked host memory for the end-to-end path */
int heat_host_alloc(void **ptr, uint64_t bytes);
int heat_host_free(void *ptr);

#ifdef __cplusplus
}
#endif
#endif /* HEAT_B200_H */
