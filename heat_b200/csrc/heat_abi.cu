// heat_b200/csrc/heat_abi.cu -- the extern "C" surface of include/heat_b200.h.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>
#include "heat_ctx.cuh"
#include "index_kernels.cuh"

namespace heat {

static thread_local std::string g_last_error;
void set_error(const std::string &msg) { g_last_error = msg; }
int fail(const std::string &msg) { g_last_error = msg; return 1; }

namespace {

// uint64 host layout -> int32 device layout; flags ids that do not fit their table.
__global__ void narrow_ids_kernel(const unsigned long long *src, int *dst, size_t n, unsigned long long limit_even,
                                  unsigned long long limit_odd, int *bad)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const unsigned long long v = src[i];
        const unsigned long long lim = (i & 1) ? limit_odd : limit_even;
        if (v >= lim) atomicExch(bad, 1);
        dst[i] = (int)v;
    }
}

// Queue copy + narrow of one uint64 host array on the context's stream (no synchronisation).  `stage` is a slice of
// the context's grow-only staging buffer; `bad` is the device flag of the whole heat_set_dataset call.
int queue_narrow(Ctx *c, const uint64_t *host, size_t n, unsigned long long *stage, int *dst, uint64_t lim_even,
                 uint64_t lim_odd, int *bad)
{
    if (!n) return 0;
    HEAT_CUDA(cudaMemcpyAsync(stage, host, n * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
    const int blocks = (int)std::min<size_t>((n + 255) / 256, (size_t)c->num_sms * 8);
    narrow_ids_kernel<<<blocks > 0 ? blocks : 1, 256, 0, c->stream>>>(stage, dst, n, lim_even, lim_odd, bad);
    c->launches += 1;
    HEAT_CUDA(cudaGetLastError());
    return 0;
}

template <class T> int grow(T **p, size_t *have, size_t need)
{
    if (*p && *have >= need) return 0;
    if (*p) cudaFree(*p);
    *p = nullptr; *have = 0;
    HEAT_CUDA(cudaMalloc(p, (need ? need : 1) * sizeof(T)));
    *have = need;
    return 0;
}

void free_dataset(Ctx *c)
{
    cudaFree(c->clicks); cudaFree(c->his); cudaFree(c->masks); cudaFree(c->stage);
    c->clicks = nullptr; c->his = nullptr; c->masks = nullptr; c->stage = nullptr;
    c->clicks_cap = c->his_cap = c->masks_cap = c->stage_cap = 0;
}
void free_tables(Ctx *c)
{
    // arrays borrowed from member 0 of an emulation group (heat_local_group_share) are not this context's to free
    if (c->shared_mask & 1u) c->w0 = nullptr;
    if (c->shared_mask & 2u) c->item_g = nullptr;
    if (c->shared_mask & 4u) c->item_w = nullptr;
    cudaFree(c->user_w); cudaFree(c->item_w); cudaFree(c->user_g); cudaFree(c->item_g); cudaFree(c->w0);
    cudaFree(c->item_base); cudaFree(c->w0_base);
    c->user_w = c->item_w = c->user_g = c->item_g = c->w0 = c->item_base = c->w0_base = nullptr;
}

TrainParams index_params(Ctx *c, uint64_t epoch, uint64_t begin, uint64_t end)
{
    TrainParams P{};
    P.clicks = c->clicks; P.his = c->his; P.masks = c->masks;
    P.begin = (long long)begin; P.end = (long long)end;
    P.sample_offset = c->cfg.sample_offset; P.seed = c->cfg.seed; P.epoch = (unsigned)epoch;
    set_sampler(c, P);
    P.max_his = (int)c->max_his; P.num_items = (int)c->cfg.num_items; P.K = (int)c->cfg.num_negs;
    return P;
}

}  // namespace
}  // namespace heat

using namespace heat;

extern "C" {

int heat_abi_version(void) { return HEAT_ABI_VERSION; }
const char *heat_last_error(void) { return g_last_error.c_str(); }

int heat_create(heat_ctx **out, const heat_config *cfg, int device)
{
    if (!out || !cfg) return fail("heat_create: null argument");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(std::string("heat_create: no CUDA device (") + cudaGetErrorString(e) +
                    "); this library has no CPU fallback");
    if (device < 0 || device >= ndev) return fail("heat_create: device index out of range");
    HEAT_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop{};
    HEAT_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(std::string("heat_create: built for sm_100a, found ") + prop.name);
    if (cfg->emb_dim == 0 || cfg->num_negs == 0 || cfg->num_items == 0 || cfg->num_users == 0)
        return fail("heat_create: emb_dim, num_negs, num_users and num_items must be non-zero");
    if (cfg->num_items >= (1ull << 31) || cfg->num_users >= (1ull << 31))
        return fail("heat_create: ids must fit int32");

    Ctx *c = new (std::nothrow) Ctx();
    if (!c) return fail("heat_create: out of host memory");
    c->cfg = *cfg;
    c->milestones.assign(cfg->milestones, cfg->milestones + (cfg->milestones ? cfg->n_milestones : 0));
    c->cfg.milestones = nullptr;
    if (c->cfg.seed == 0) c->cfg.seed = 2022;
    if (c->cfg.chunk == 0) c->cfg.chunk = 64;
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    c->l_r = cfg->l_r;
    HEAT_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    HEAT_CUDA(cudaEventCreate(&c->ev0));
    HEAT_CUDA(cudaEventCreate(&c->ev1));
    HEAT_CUDA(cudaEventCreate(&c->ev_e0));
    HEAT_CUDA(cudaEventCreate(&c->ev_e1));
    HEAT_CUDA(cudaMalloc(&c->chunk_counter, sizeof(unsigned long long)));
    HEAT_CUDA(cudaMalloc(&c->loss_sum, sizeof(double)));
    *out = reinterpret_cast<heat_ctx *>(c);
    return 0;
}

int heat_destroy(heat_ctx *ctx)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return 0;
    cudaSetDevice(c->device);
    comm_destroy(c);
    free_dataset(c); free_tables(c);
    eval_free(c);
    cudaFree(c->chunk_counter); cudaFree(c->loss_sum);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->ev_e0) cudaEventDestroy(c->ev_e0);
    if (c->ev_e1) cudaEventDestroy(c->ev_e1);
    for (cudaEvent_t e : c->seg_ev) cudaEventDestroy(e);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return 0;
}

int heat_set_dataset(heat_ctx *ctx, const uint64_t *clicks, uint64_t n, const uint64_t *his_items,
                     uint64_t num_users, uint64_t max_his, const uint64_t *masks)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !clicks || !his_items || !masks) return fail("heat_set_dataset: null argument");
    if (num_users != c->cfg.num_users) return fail("heat_set_dataset: num_users differs from the config");
    if (n >= (1ull << 40)) return fail("heat_set_dataset: too many samples");
    HEAT_CUDA(cudaSetDevice(c->device));
    // device buffers and the uint64 staging area are grow-only: a caller that re-submits a dataset of the same shape
    // every epoch (bench.py's end-to-end arm) pays three copies, three narrowing launches and ONE synchronisation
    const size_t nh = num_users * max_his, n_ids = 2 * n + nh + num_users;
    if (int rc = grow(&c->clicks, &c->clicks_cap, (size_t)n)) return rc;
    if (int rc = grow(&c->his, &c->his_cap, nh)) return rc;
    if (int rc = grow(&c->masks, &c->masks_cap, (size_t)num_users)) return rc;
    if (int rc = grow(&c->stage, &c->stage_cap, n_ids + 2)) return rc;
    c->n_samples = n; c->max_his = max_his; c->cfg.train_size = n;
    c->nmax_for = ~0ull;                                // the largest shard over the ranks is agreed on again
    int *bad = reinterpret_cast<int *>(c->stage + n_ids);
    HEAT_CUDA(cudaMemsetAsync(bad, 0, 3 * sizeof(int), c->stream));
    if (int rc = queue_narrow(c, clicks, 2 * n, c->stage, reinterpret_cast<int *>(c->clicks), c->cfg.num_users, c->cfg.num_items, bad)) return rc;
    if (int rc = queue_narrow(c, his_items, nh, c->stage + 2 * n, c->his, c->cfg.num_items, c->cfg.num_items, bad + 1)) return rc;
    if (int rc = queue_narrow(c, masks, num_users, c->stage + 2 * n + nh, c->masks, max_his + 1, max_his + 1, bad + 2)) return rc;
    int h_bad[3] = {0, 0, 0};
    HEAT_CUDA(cudaMemcpyAsync(h_bad, bad, sizeof(h_bad), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    static const char *what[3] = {"clicks", "historical_items", "masks"};
    for (int i = 0; i < 3; ++i)
        if (h_bad[i]) {
            c->n_samples = 0;
            return fail(std::string(what[i]) + ": id out of range for its table");
        }
    return 0;
}

int heat_bind_tables(heat_ctx *ctx, float *user_w, float *item_w, float *w0)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !user_w || !item_w || !w0) return fail("heat_bind_tables: null argument");
    if (c->nccl_comm || c->local_group)
        return fail("heat_bind_tables: the tables are already shared with other ranks (peers hold mappings of them); create a new context");
    HEAT_CUDA(cudaSetDevice(c->device));
    free_tables(c);
    const size_t D = c->cfg.emb_dim, ub = c->cfg.num_users * D * sizeof(float), ib = c->cfg.num_items * D * sizeof(float);
    HEAT_CUDA(cudaMalloc(&c->user_w, ub)); HEAT_CUDA(cudaMalloc(&c->user_g, ub));
    HEAT_CUDA(cudaMalloc(&c->item_w, ib)); HEAT_CUDA(cudaMalloc(&c->item_g, ib));
    HEAT_CUDA(cudaMalloc(&c->w0, D * D * sizeof(float)));
    c->h_user_w = user_w; c->h_item_w = item_w; c->h_w0 = w0;
    int rc = heat_upload_tables(ctx);
    if (!rc) rc = heat_zero_grad(ctx);
    return rc;
}

int heat_upload_tables(heat_ctx *ctx)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !c->user_w) return fail("heat_upload_tables: no tables bound");
    HEAT_CUDA(cudaSetDevice(c->device));
    const size_t D = c->cfg.emb_dim;
    HEAT_CUDA(cudaMemcpyAsync(c->user_w, c->h_user_w, c->cfg.num_users * D * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    HEAT_CUDA(cudaMemcpyAsync(c->item_w, c->h_item_w, c->cfg.num_items * D * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    HEAT_CUDA(cudaMemcpyAsync(c->w0, c->h_w0, D * D * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    if (c->item_base) {
        HEAT_CUDA(cudaMemcpyAsync(c->item_base, c->item_w, c->cfg.num_items * D * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
        HEAT_CUDA(cudaMemcpyAsync(c->w0_base, c->w0, D * D * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
    }
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int heat_sync_to_host(heat_ctx *ctx)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !c->user_w) return fail("heat_sync_to_host: no tables bound");
    HEAT_CUDA(cudaSetDevice(c->device));
    const size_t D = c->cfg.emb_dim;
    HEAT_CUDA(cudaMemcpyAsync(c->h_user_w, c->user_w, c->cfg.num_users * D * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaMemcpyAsync(c->h_item_w, c->item_w, c->cfg.num_items * D * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaMemcpyAsync(c->h_w0, c->w0, D * D * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int heat_zero_grad(heat_ctx *ctx)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !c->user_g) return fail("heat_zero_grad: no tables bound");
    HEAT_CUDA(cudaSetDevice(c->device));
    const size_t D = c->cfg.emb_dim;
    HEAT_CUDA(cudaMemsetAsync(c->user_g, 0, c->cfg.num_users * D * sizeof(float), c->stream));
    HEAT_CUDA(cudaMemsetAsync(c->item_g, 0, c->cfg.num_items * D * sizeof(float), c->stream));
    if (c->item_g_base) HEAT_CUDA(cudaMemsetAsync(c->item_g_base, 0, c->cfg.num_items * D * sizeof(float), c->stream));
    return 0;
}

int heat_train_range(heat_ctx *ctx, uint64_t begin, uint64_t end, double *loss_sum, float *loss_per_sample)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail("heat_train_range: null context");
    HEAT_CUDA(cudaSetDevice(c->device));
    return launch_train(c, begin, end, loss_sum, loss_per_sample);
}

namespace {

// LR schedule, engine.cpp:156-160 -> optimizer.cpp:24-38 (decay 0.1)
void lr_schedule(Ctx *c)
{
    if (c->milestones.size() > 1) {
        for (uint64_t m : c->milestones)
            if (m == c->epoch) { c->l_r = c->l_r * 0.1f; break; }
    } else if (c->milestones.size() == 1 && c->milestones[0] != 0) {
        if (c->epoch > 0 && c->epoch % c->milestones[0] == 0) c->l_r = c->l_r * 0.1f;
    }
}

int need_seg_events(Ctx *c, uint64_t nseg)
{
    while (c->seg_ev.size() < 3 * nseg) {
        cudaEvent_t e = nullptr;
        HEAT_CUDA(cudaEventCreate(&e));
        c->seg_ev.push_back(e);
    }
    return 0;
}

// segment i of nseg over N local samples
inline void seg_range(uint64_t N, uint64_t i, uint64_t nseg, uint64_t *b, uint64_t *e)
{
    *b = (uint64_t)((unsigned __int128)N * i / nseg);
    *e = (uint64_t)((unsigned __int128)N * (i + 1) / nseg);
}

// device times of the epoch just synchronised: training kernels and exchanges, from the segment events
void seg_times(Ctx *c, uint64_t nseg, bool exchanged)
{
    float k = 0.f, x = 0.f, t = 0.f;
    for (uint64_t i = 0; i < nseg; ++i) {
        if (cudaEventElapsedTime(&t, c->seg_ev[3 * i], c->seg_ev[3 * i + 1]) == cudaSuccess) k += t;
        if (exchanged && cudaEventElapsedTime(&t, c->seg_ev[3 * i + 1], c->seg_ev[3 * i + 2]) == cudaSuccess) x += t;
    }
    c->last_kernel_ms = k; c->last_exchange_ms = x; c->last_exchanges = exchanged ? nseg : 0;
}

}  // namespace

int heat_train_epoch(heat_ctx *ctx, double *mean_loss)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail("heat_train_epoch: null context");
    if (c->local_group) return fail("heat_train_epoch: the context belongs to a local emulation group (heat_local_group_train_epoch)");
    HEAT_CUDA(cudaSetDevice(c->device));
    lr_schedule(c);
    const uint64_t N = c->n_samples;
    // Multi-GPU: every rank runs the same number of segments (a collective follows each one), so
    // the count comes from the largest shard; a rank's segment i is its own [N*i/nseg, N*(i+1)/nseg).
    uint64_t nseg = 1;
    if (c->world > 1) {
        if (c->nmax_for != N) {
            uint64_t nmax = N;
            int rc = comm_max_u64(c, &nmax);
            if (rc) return rc;
            c->nmax = nmax; c->nmax_for = N;
        }
        const uint64_t iv = sync_interval_of(c);
        nseg = std::max<uint64_t>(1, (c->nmax + iv - 1) / iv);
    }
    if (int rc = need_seg_events(c, nseg)) return rc;
    // The whole epoch is queued on the stream -- segments and exchanges back to back -- and synchronised once: the
    // loss accumulates on the device and is read at the end.
    HEAT_CUDA(cudaEventRecord(c->ev_e0, c->stream));
    HEAT_CUDA(cudaMemsetAsync(c->loss_sum, 0, sizeof(double), c->stream));
    for (uint64_t i = 0; i < nseg; ++i) {
        uint64_t b, e;
        seg_range(N, i, nseg, &b, &e);
        HEAT_CUDA(cudaEventRecord(c->seg_ev[3 * i], c->stream));
        int rc = queue_train(c, b, e);
        if (rc) return rc;
        HEAT_CUDA(cudaEventRecord(c->seg_ev[3 * i + 1], c->stream));
        if (c->world > 1) {
            rc = comm_sync_items(c);
            if (rc) return rc;
            HEAT_CUDA(cudaEventRecord(c->seg_ev[3 * i + 2], c->stream));
        }
    }
    int rc = heat_zero_grad(ctx);                       // engine.cpp:345-347
    if (rc) return rc;
    double total = 0.0;
    HEAT_CUDA(cudaMemcpyAsync(&total, c->loss_sum, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaEventRecord(c->ev_e1, c->stream));
    cudaError_t se = cudaStreamSynchronize(c->stream);
    if (se != cudaSuccess) return fail(std::string("training epoch: ") + cudaGetErrorString(se));
    cudaEventElapsedTime(&c->last_epoch_ms, c->ev_e0, c->ev_e1);
    seg_times(c, nseg, c->world > 1);
    c->epoch += 1;                                      // engine.cpp:378
    double samples = (double)N;
    rc = comm_reduce_loss(c, &total, &samples);         // engine.cpp:380-383
    if (rc) return rc;
    if (mean_loss) *mean_loss = samples > 0 ? total / samples : 0.0;
    return 0;
}

int heat_local_group_init(heat_ctx **ctxs, int n)
{
    if (!ctxs || n < 1) return fail("heat_local_group_init: bad argument");
    Ctx **cs = reinterpret_cast<Ctx **>(ctxs);
    if (!cs[0]) return fail("heat_local_group_init: null context");
    HEAT_CUDA(cudaSetDevice(cs[0]->device));
    return local_group_init(cs, n);
}

int heat_local_group_train_epoch(heat_ctx **ctxs, int n, double *mean_loss)
{
    if (!ctxs || n < 1 || n > 8) return fail("heat_local_group_train_epoch: bad argument");
    Ctx **cs = reinterpret_cast<Ctx **>(ctxs);
    for (int q = 0; q < n; ++q)
        if (!cs[q] || !cs[q]->local_group || cs[q]->world != n || cs[q]->rank != q)
            return fail("heat_local_group_train_epoch: not the group heat_local_group_init was called with");
    Ctx *c0 = cs[0];
    HEAT_CUDA(cudaSetDevice(c0->device));
    uint64_t nmax = 0, total_n = 0;
    for (int q = 0; q < n; ++q) { lr_schedule(cs[q]); nmax = std::max(nmax, cs[q]->n_samples); total_n += cs[q]->n_samples; }
    const uint64_t iv = sync_interval_of(c0);
    const uint64_t nseg = std::max<uint64_t>(1, (nmax + iv - 1) / iv);
    if (int rc = need_seg_events(c0, nseg)) return rc;
    for (int q = 0; q < n; ++q) HEAT_CUDA(cudaMemsetAsync(cs[q]->loss_sum, 0, sizeof(double), cs[q]->stream));
    float k_ms = 0.f, x_ms = 0.f;
    for (uint64_t i = 0; i < nseg; ++i) {
        for (int q = 0; q < n; ++q) {                   // "ranks" one after the other: replicas are independent between exchanges
            uint64_t b, e;
            seg_range(cs[q]->n_samples, i, nseg, &b, &e);
            HEAT_CUDA(cudaEventRecord(cs[q]->ev_e0, cs[q]->stream));
            int rc = queue_train(cs[q], b, e);
            if (rc) return rc;
            HEAT_CUDA(cudaEventRecord(cs[q]->ev_e1, cs[q]->stream));
            HEAT_CUDA(cudaStreamSynchronize(cs[q]->stream));
            float t = 0.f;
            cudaEventElapsedTime(&t, cs[q]->ev_e0, cs[q]->ev_e1);
            if (q == 0) k_ms += t;                      // rank 0's kernels, as a real rank would report
        }
        HEAT_CUDA(cudaEventRecord(c0->seg_ev[3 * i + 1], c0->stream));
        int rc = local_group_sync(cs, n);
        if (rc) return rc;
        HEAT_CUDA(cudaEventRecord(c0->seg_ev[3 * i + 2], c0->stream));
        HEAT_CUDA(cudaStreamSynchronize(c0->stream));
        float t = 0.f;
        cudaEventElapsedTime(&t, c0->seg_ev[3 * i + 1], c0->seg_ev[3 * i + 2]);
        x_ms += t;
    }
    double total = 0.0;
    for (int q = 0; q < n; ++q) {
        int rc = heat_zero_grad(reinterpret_cast<heat_ctx *>(cs[q]));
        if (rc) return rc;
        double part = 0.0;
        HEAT_CUDA(cudaMemcpyAsync(&part, cs[q]->loss_sum, sizeof(double), cudaMemcpyDeviceToHost, cs[q]->stream));
        HEAT_CUDA(cudaStreamSynchronize(cs[q]->stream));
        total += part;
        cs[q]->epoch += 1;
        cs[q]->last_kernel_ms = k_ms; cs[q]->last_exchange_ms = x_ms; cs[q]->last_exchanges = nseg;
    }
    if (mean_loss) *mean_loss = total_n ? total / (double)total_n : 0.0;
    return 0;
}

int heat_local_group_share(heat_ctx **ctxs, int n, uint32_t mask)
{
    if (!ctxs || n < 1 || n > 8) return fail("heat_local_group_share: bad argument");
    Ctx **cs = reinterpret_cast<Ctx **>(ctxs);
    if (!cs[0]) return fail("heat_local_group_share: null context");
    HEAT_CUDA(cudaSetDevice(cs[0]->device));
    return local_group_share(cs, n, mask);
}

int heat_last_exchange(heat_ctx *ctx, uint64_t *count, float *ms)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail("heat_last_exchange: null context");
    if (count) *count = c->last_exchanges;
    if (ms) *ms = c->last_exchange_ms;
    return 0;
}

int heat_get_grads(heat_ctx *ctx, float *user_g, float *item_g)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !c->user_g) return fail("heat_get_grads: no tables bound");
    HEAT_CUDA(cudaSetDevice(c->device));
    const size_t D = c->cfg.emb_dim;
    if (user_g) HEAT_CUDA(cudaMemcpyAsync(user_g, c->user_g, c->cfg.num_users * D * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    if (item_g) HEAT_CUDA(cudaMemcpyAsync(item_g, c->item_g, c->cfg.num_items * D * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

namespace {
struct StateHeader {
    char magic[8];
    uint32_t version, emb_dim;
    uint64_t num_users, num_items, epoch;
    float l_r;
    uint32_t pad[5];
};
static_assert(sizeof(StateHeader) == 64, "checkpoint header is 64 bytes");
struct FileCloser { FILE *f; ~FileCloser() { if (f) fclose(f); } };
}  // namespace

int heat_save_state(heat_ctx *ctx, const char *path)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !path) return fail("heat_save_state: null argument");
    if (!c->user_w) return fail("heat_save_state: no tables bound");
    HEAT_CUDA(cudaSetDevice(c->device));
    const size_t D = c->cfg.emb_dim, U = c->cfg.num_users, I = c->cfg.num_items;
    const std::string tmp = std::string(path) + ".tmp";
    FileCloser fc{fopen(tmp.c_str(), "wb")};
    if (!fc.f) return fail(std::string("heat_save_state: cannot open ") + tmp);
    StateHeader h{};
    memcpy(h.magic, "HEATB200", 8);
    h.version = 1; h.emb_dim = (uint32_t)D; h.num_users = U; h.num_items = I; h.epoch = c->epoch; h.l_r = c->l_r;
    if (fwrite(&h, sizeof(h), 1, fc.f) != 1) return fail("heat_save_state: write failed");
    const struct { const float *p; size_t n; } parts[5] = {{c->user_w, U * D}, {c->item_w, I * D}, {c->w0, D * D}, {c->user_g, U * D}, {c->item_g, I * D}};
    const size_t chunk = (size_t)16 << 20;                       // floats per staging copy (64 MiB, pinned)
    float *stage = nullptr;
    HEAT_CUDA(cudaMallocHost(&stage, chunk * sizeof(float)));
    int rc = 0;
    for (int i = 0; i < 5 && !rc; ++i)
        for (size_t off = 0; off < parts[i].n && !rc; off += chunk) {
            const size_t n = std::min(chunk, parts[i].n - off);
            if (cudaMemcpyAsync(stage, parts[i].p + off, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
                cudaStreamSynchronize(c->stream) != cudaSuccess)
                rc = fail("heat_save_state: device read failed");
            else if (fwrite(stage, sizeof(float), n, fc.f) != n)
                rc = fail("heat_save_state: write failed");
        }
    cudaFreeHost(stage);
    if (rc) return rc;
    if (fclose(fc.f) != 0) { fc.f = nullptr; return fail("heat_save_state: close failed"); }
    fc.f = nullptr;
    if (rename(tmp.c_str(), path) != 0) return fail("heat_save_state: rename failed");   // a crash never leaves a torn file
    return 0;
}

int heat_load_state(heat_ctx *ctx, const char *path)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !path) return fail("heat_load_state: null argument");
    if (!c->user_w) return fail("heat_load_state: no tables bound");
    HEAT_CUDA(cudaSetDevice(c->device));
    const size_t D = c->cfg.emb_dim, U = c->cfg.num_users, I = c->cfg.num_items;
    FileCloser fc{fopen(path, "rb")};
    if (!fc.f) return fail(std::string("heat_load_state: cannot open ") + path);
    StateHeader h{};
    if (fread(&h, sizeof(h), 1, fc.f) != 1 || memcmp(h.magic, "HEATB200", 8) != 0 || h.version != 1)
        return fail("heat_load_state: not a heat_b200 checkpoint");
    if (h.emb_dim != D || h.num_users != U || h.num_items != I)
        return fail("heat_load_state: checkpoint shape differs from the context's config");
    const struct { float *p; size_t n; } parts[5] = {{c->user_w, U * D}, {c->item_w, I * D}, {c->w0, D * D}, {c->user_g, U * D}, {c->item_g, I * D}};
    const size_t chunk = (size_t)16 << 20;
    float *stage = nullptr;
    HEAT_CUDA(cudaMallocHost(&stage, chunk * sizeof(float)));
    int rc = 0;
    for (int i = 0; i < 5 && !rc; ++i)
        for (size_t off = 0; off < parts[i].n && !rc; off += chunk) {
            const size_t n = std::min(chunk, parts[i].n - off);
            if (fread(stage, sizeof(float), n, fc.f) != n)
                rc = fail("heat_load_state: file is truncated");
            else if (cudaMemcpyAsync(parts[i].p + off, stage, n * sizeof(float), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
                     cudaStreamSynchronize(c->stream) != cudaSuccess)
                rc = fail("heat_load_state: device write failed");
        }
    cudaFreeHost(stage);
    if (rc) return rc;
    c->epoch = h.epoch; c->l_r = h.l_r;
    if (c->item_base) {
        HEAT_CUDA(cudaMemcpyAsync(c->item_base, c->item_w, I * D * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
        HEAT_CUDA(cudaMemcpyAsync(c->w0_base, c->w0, D * D * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
    }
    return heat_sync_to_host(ctx);
}

int heat_set_lr(heat_ctx *ctx, float l_r) { Ctx *c = reinterpret_cast<Ctx *>(ctx); if (!c) return fail("null context"); c->l_r = l_r; return 0; }
int heat_get_lr(heat_ctx *ctx, float *l_r) { Ctx *c = reinterpret_cast<Ctx *>(ctx); if (!c || !l_r) return fail("null argument"); *l_r = c->l_r; return 0; }
int heat_get_epoch(heat_ctx *ctx, uint64_t *epoch) { Ctx *c = reinterpret_cast<Ctx *>(ctx); if (!c || !epoch) return fail("null argument"); *epoch = c->epoch; return 0; }
int heat_set_epoch(heat_ctx *ctx, uint64_t epoch) { Ctx *c = reinterpret_cast<Ctx *>(ctx); if (!c) return fail("null context"); c->epoch = epoch; return 0; }
int heat_last_kernel_ms(heat_ctx *ctx, float *ms) { Ctx *c = reinterpret_cast<Ctx *>(ctx); if (!c || !ms) return fail("null argument"); *ms = c->last_kernel_ms; return 0; }
int heat_last_epoch_ms(heat_ctx *ctx, float *ms) { Ctx *c = reinterpret_cast<Ctx *>(ctx); if (!c || !ms) return fail("null argument"); *ms = c->last_epoch_ms; return 0; }
int heat_last_eval_ms(heat_ctx *ctx, float *ms) { Ctx *c = reinterpret_cast<Ctx *>(ctx); if (!c || !ms) return fail("null argument"); *ms = c->last_eval_ms; return 0; }
int heat_last_engine(heat_ctx *ctx, int *engine) { Ctx *c = reinterpret_cast<Ctx *>(ctx); if (!c || !engine) return fail("null argument"); *engine = c->last_engine; return 0; }
int heat_launch_count(heat_ctx *ctx, uint64_t *n) { Ctx *c = reinterpret_cast<Ctx *>(ctx); if (!c || !n) return fail("null argument"); *n = c->launches; return 0; }

int heat_sample_negatives(heat_ctx *ctx, uint64_t epoch, uint64_t begin, uint64_t end, uint64_t *out)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !out) return fail("heat_sample_negatives: null argument");
    if (!c->clicks) return fail("heat_sample_negatives: heat_set_dataset has not been called");
    if (end > c->n_samples || begin > end) return fail("heat_sample_negatives: range out of bounds");
    HEAT_CUDA(cudaSetDevice(c->device));
    const size_t n = (end - begin) * c->cfg.num_negs;
    if (!n) return 0;
    unsigned long long *d = nullptr;
    HEAT_CUDA(cudaMalloc(&d, n * sizeof(uint64_t)));
    const TrainParams P = index_params(c, epoch, begin, end);
    sample_negatives_kernel<<<c->num_sms * 4, 256, 0, c->stream>>>(P, d);
    c->launches += 1;
    HEAT_CUDA(cudaGetLastError());
    HEAT_CUDA(cudaMemcpyAsync(out, d, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(d);
    return 0;
}

int heat_history_indices(heat_ctx *ctx, uint64_t begin, uint64_t end, uint64_t *out_ids, uint64_t *out_count)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !out_ids || !out_count) return fail("heat_history_indices: null argument");
    if (!c->clicks) return fail("heat_history_indices: heat_set_dataset has not been called");
    if (end > c->n_samples || begin > end) return fail("heat_history_indices: range out of bounds");
    if (c->max_his > 128) return fail("max_his > 128 is not supported");
    HEAT_CUDA(cudaSetDevice(c->device));
    const size_t n = end - begin;
    if (!n) return 0;
    unsigned long long *d_ids = nullptr, *d_cnt = nullptr;
    HEAT_CUDA(cudaMalloc(&d_ids, n * c->max_his * sizeof(uint64_t)));
    HEAT_CUDA(cudaMalloc(&d_cnt, n * sizeof(uint64_t)));
    const TrainParams P = index_params(c, 0, begin, end);
    history_indices_kernel<<<c->num_sms * 4, 256, 0, c->stream>>>(P, d_ids, d_cnt);
    c->launches += 1;
    HEAT_CUDA(cudaGetLastError());
    HEAT_CUDA(cudaMemcpyAsync(out_ids, d_ids, n * c->max_his * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaMemcpyAsync(out_count, d_cnt, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(d_ids); cudaFree(d_cnt);
    return 0;
}

uint64_t heat_philox_draw(uint64_t seed, uint64_t epoch, uint64_t sample, uint32_t k, uint64_t num_items)
{
    return philox_draw(seed, (uint32_t)epoch, sample, k, (uint32_t)num_items);
}
uint64_t heat_tile_draw(uint64_t seed, uint64_t epoch, uint64_t sample, uint32_t k, uint64_t num_items, uint64_t tile_size,
                        uint64_t refresh_interval)
{
    return philox_draw(seed, (uint32_t)epoch, sample, k, (uint32_t)num_items, (uint32_t)tile_size, refresh_interval ? refresh_interval : 1);
}

int heat_eval_dense(heat_ctx *ctx, float *out)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !out) return fail("heat_eval_dense: null argument");
    HEAT_CUDA(cudaSetDevice(c->device));
    return eval_dense(c, out);
}
int heat_eval_topk(heat_ctx *ctx, uint32_t k, const int64_t *train_indptr, const int32_t *train_items,
                   int32_t *topk_ids, float *topk_scores)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !topk_ids) return fail("heat_eval_topk: null argument");
    HEAT_CUDA(cudaSetDevice(c->device));
    return eval_topk(c, k, train_indptr, train_items, topk_ids, topk_scores);
}
int heat_eval_metrics(heat_ctx *ctx, const int64_t *train_indptr, const int32_t *train_items,
                      const int64_t *test_indptr, const int32_t *test_items, double out[6], double sums_out[7])
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !test_indptr || !test_items || !out) return fail("heat_eval_metrics: null argument");
    HEAT_CUDA(cudaSetDevice(c->device));
    return eval_metrics(c, train_indptr, train_items, test_indptr, test_items, out, sums_out);
}

int heat_comm_unique_id(uint8_t out[128]) { return comm_unique_id(out); }
int heat_comm_init(heat_ctx *ctx, int rank, int world, const uint8_t unique_id[128])
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail("heat_comm_init: null context");
    HEAT_CUDA(cudaSetDevice(c->device));
    return comm_init(c, rank, world, unique_id);
}
int heat_comm_sync_items(heat_ctx *ctx)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail("heat_comm_sync_items: null context");
    HEAT_CUDA(cudaSetDevice(c->device));
    return comm_sync_items(c);
}
int heat_comm_allreduce_f64(heat_ctx *ctx, double *vals, int n)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !vals || n < 0) return fail("heat_comm_allreduce_f64: bad argument");
    HEAT_CUDA(cudaSetDevice(c->device));
    return comm_allreduce_f64(c, vals, n);
}
int heat_comm_mode(heat_ctx *ctx, int *mode)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !mode) return fail("heat_comm_mode: null argument");
    *mode = comm_mode(c);
    return 0;
}

int heat_set_sync_interval(heat_ctx *ctx, uint64_t samples)
{
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail("heat_set_sync_interval: null context");
    c->sync_interval = samples;
    return 0;
}

int heat_probe_l2_bandwidth(int device, uint64_t bytes, int iters, double *gbps) { return probe_l2(device, bytes, iters, gbps); }
int heat_probe_tf32_peak(int device, int iters, double *tflops) { return probe_tf32(device, iters, tflops); }

int heat_host_alloc(void **ptr, uint64_t bytes)
{
    if (!ptr) return fail("heat_host_alloc: null argument");
    HEAT_CUDA(cudaHostAlloc(ptr, bytes, cudaHostAllocDefault));
    return 0;
}
int heat_host_free(void *ptr) { HEAT_CUDA(cudaFreeHost(ptr)); return 0; }

}  // extern "C"
