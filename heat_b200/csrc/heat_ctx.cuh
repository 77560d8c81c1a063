// heat_b200/csrc/heat_ctx.cuh -- device-resident state behind the C ABI (include/heat_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <algorithm>
#include <string>
#include <vector>
#include "../../include/heat_b200.h"

namespace heat {

// HBM layout (all allocations 256-byte aligned by cudaMalloc):
//   clicks  int2 [N]            8 B/sample          his    int32 [U, max_his]
//   masks   int32 [U]                               user_w fp32 [U, D]   user_g fp32 [U, D]
//   item_w  fp32 [I, D]   item_g fp32 [I, D]        w0     fp32 [D, D]
//   item_base / w0_base (multi-GPU only): copies taken at the last sync, for the delta all-reduce
struct Ctx {
    heat_config cfg{};
    std::vector<uint64_t> milestones;
    int device = 0;
    int num_sms = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_e0 = nullptr, ev_e1 = nullptr;

    int2 *clicks = nullptr;
    int *his = nullptr, *masks = nullptr;
    uint64_t n_samples = 0, max_his = 0;
    unsigned long long *stage = nullptr;                  // uint64 staging of heat_set_dataset (+ 3 range flags)
    size_t clicks_cap = 0, his_cap = 0, masks_cap = 0, stage_cap = 0;
    float *user_w = nullptr, *item_w = nullptr, *user_g = nullptr, *item_g = nullptr, *w0 = nullptr;
    float *h_user_w = nullptr, *h_item_w = nullptr, *h_w0 = nullptr;   // bound host buffers
    unsigned long long *chunk_counter = nullptr;
    double *loss_sum = nullptr;

    float l_r = 0.f;          // optimizer LR (scheduled); cfg.l_r stays the aggregator's LR
    uint64_t epoch = 0;
    float last_kernel_ms = 0.f, last_eval_ms = 0.f, last_epoch_ms = 0.f, last_exchange_ms = 0.f;
    uint64_t last_exchanges = 0;
    std::vector<cudaEvent_t> seg_ev;           // 3 events per sync segment of an epoch: train start, train end, exchange end
    uint64_t nmax_for = ~0ull, nmax = 0;       // largest shard over the ranks (cached per dataset size)
    uint64_t launches = 0;
    int last_engine = 0;      // kernel family of the last training launch: 1 warp per sample, 2 CTA per sample, 3 large-negative

    // multi-GPU
    int rank = 0, world = 1;
    void *nccl_comm = nullptr;
    float *item_base = nullptr, *w0_base = nullptr, *delta = nullptr;
    float *item_g_base = nullptr;              // sync_grads: item grad rows at the last exchange
    bool local_group = false;                  // member of a single-device emulation group (heat_local_group_init)
    uint32_t shared_mask = 0;                  // heat_local_group_share: arrays this member borrows from member 0 (not freed here)
    uint32_t group_shared = 0;                 // ... and the arrays the group shares at all (skipped by the reconciliation)
    double *loss_pair = nullptr;
    uint64_t sync_interval = 0, syncs = 0;
    // fused exchange over NVLink peer memory (comm.cu): every rank's tables and flags mapped through CUDA IPC
    bool peer_ok = false;
    float *peer_item_w[8] = {}, *peer_item_base[8] = {}, *peer_w0[8] = {}, *peer_w0_base[8] = {};
    float *peer_item_g[8] = {}, *peer_item_g_base[8] = {};
    unsigned long long *peer_flags[8] = {};    // [2][8] arrival counters per rank: [0] start barrier, [1] end barrier
    unsigned long long *xflags = nullptr;      // this rank's flags
    unsigned *grid_sync = nullptr;             // intra-grid release flag and done counter of peer_sync_kernel

    // evaluation scratch (lazily sized)
    void *eval_scratch = nullptr;
    size_t eval_scratch_bytes = 0;
};

void set_error(const std::string &msg);
int fail(const std::string &msg);

#define HEAT_CUDA(expr)                                                                          \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess)                                                                   \
            return ::heat::fail(std::string(#expr) + ": " + cudaGetErrorString(_e));             \
    } while (0)

// multi-GPU exchange policy (heat_config.sync_average, heat_set_sync_interval): the documented defaults
inline bool sync_mean(const Ctx *c) { return c->cfg.sync_average != 2; }
inline uint64_t sync_interval_of(const Ctx *c)
{
    return c->sync_interval ? c->sync_interval : std::max<uint64_t>(65536, 8 * c->cfg.num_items);
}

// hot-tile sampler fields of a kernel parameter block from the config (tile_sampler && neg_sampler == 1), else uniform
template <class Params> inline void set_sampler(const Ctx *c, Params &P)
{
    const bool tile = c->cfg.tile_sampler && c->cfg.neg_sampler == 1 && c->cfg.tile_size > 0;
    P.tile_size = tile ? (unsigned)c->cfg.tile_size : 0u;
    P.refresh_interval = c->cfg.refresh_interval ? c->cfg.refresh_interval : 1ull;
}

// train_launch.cu
int launch_train(Ctx *c, uint64_t begin, uint64_t end, double *loss_sum_host, float *loss_per_sample_host);
// the same launch queued on the context's stream without any synchronisation; the loss is ADDED to c->loss_sum
int queue_train(Ctx *c, uint64_t begin, uint64_t end);
// train_launch_tile.cu: the CTA kernel's hot-tile-sampler instantiations (-1: (emb_dim, num_negs) not covered)
struct TrainParams;
int dispatch_cta_tile(Ctx *c, const TrainParams &P, int blocks);
// train_bigk.cu
int dispatch_bigk(Ctx *c, const TrainParams &P, bool deterministic);
// eval_kernel.cu
int eval_dense(Ctx *c, float *out_host);
int eval_topk(Ctx *c, uint32_t k, const int64_t *train_indptr, const int32_t *train_items,
              int32_t *topk_ids_host, float *topk_scores_host);
int eval_metrics(Ctx *c, const int64_t *train_indptr, const int32_t *train_items,
                 const int64_t *test_indptr, const int32_t *test_items, double out[6], double sums[7]);
void eval_free(Ctx *c);
// comm.cu
int comm_unique_id(uint8_t out[128]);
int comm_init(Ctx *c, int rank, int world, const uint8_t uid[128]);
int comm_sync_items(Ctx *c);
int local_group_init(Ctx **cs, int n);
int local_group_sync(Ctx **cs, int n);
int local_group_share(Ctx **cs, int n, uint32_t mask);
int comm_mode(const Ctx *c);
int comm_reduce_loss(Ctx *c, double *loss_sum, double *samples);
int comm_allreduce_f64(Ctx *c, double *vals, int n);
int comm_max_u64(Ctx *c, uint64_t *v);
void comm_destroy(Ctx *c);
// probe.cu
int probe_l2(int device, uint64_t bytes, int iters, double *gbps);
// probe_tf32.cu
int probe_tf32(int device, int iters, double *tflops);

}  // namespace heat
