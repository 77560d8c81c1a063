// heat_b200/csrc/comm.cu -- multi-GPU exchange step: one process per GPU, NCCL over NVLink/NVSwitch.
//
// Partitioning: users, their samples, user rows and user grad rows are split by
// contiguous user-id range and never communicated.  The item table and W0 are replicated; every
// rank trains its shard against its replica and, every `sync_interval` samples, the replicas are
// reconciled by ONE all-reduce of the item-table delta and one of the W0 delta:
//     delta = w - base;  all-reduce(sum) delta;  base += delta (or delta / world);  w = base
// This replaces the fork's per-row MPI_Bcast / MPI_Allreduce of item rows and its W0 average
// (train/engine.cpp:148-152, 262-286, 354-375).  Item grad rows stay local to a rank.
//
// Exchange kernel (default): the all-reduce is not a library call between two elementwise kernels but ONE kernel per
// rank over NVLink peer memory (`peer_sync_kernel`): every rank maps every other rank's item table, base copy and W0
// through CUDA IPC at heat_comm_init; rank r owns a 1/world slice of the rows, reads that slice of every replica
// straight from the peers' HBM (system-scope loads), forms base + scale * sum_q (w_q - base) in rank order (so all
// replicas receive the same bits) and stores the result into every replica's table and base copy.  Two flag
// barriers in peer memory bracket it (all ranks' training stores done / all remote stores landed).  Per rank that is
// (world-1)/world of the table read and 2(world-1)/world written over NVLink instead of a delta kernel, a ring
// all-reduce and an apply kernel (three passes over HBM plus the ring's 2(world-1)/world).  NCCL remains the
// bootstrap (handle all-gather), the scalar reductions, and the fallback when peer mapping is unavailable.
//
// NCCL is resolved at run time (dlopen "libnccl.so.2"): inside a PyTorch process that is the
// library torch already loaded, so both share one NCCL; in a plain C++ host it is the system one.
#include <dlfcn.h>
#include <nccl.h>
#include <cstring>
#include <vector>
#include "heat_ctx.cuh"

namespace heat {
namespace {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
};

NcclApi *nccl()
{
    static NcclApi api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (api.handle) {
#define HEAT_SYM(field, sym) *reinterpret_cast<void **>(&api.field) = dlsym(api.handle, sym)
            HEAT_SYM(GetUniqueId, "ncclGetUniqueId");
            HEAT_SYM(CommInitRank, "ncclCommInitRank");
            HEAT_SYM(CommDestroy, "ncclCommDestroy");
            HEAT_SYM(AllReduce, "ncclAllReduce");
            HEAT_SYM(AllGather, "ncclAllGather");
            HEAT_SYM(GetErrorString, "ncclGetErrorString");
            HEAT_SYM(GetVersion, "ncclGetVersion");
#undef HEAT_SYM
        }
    }
    if (!api.handle || !api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy) return nullptr;
    return &api;
}

#define HEAT_NCCL(expr)                                                                           \
    do {                                                                                          \
        ncclResult_t _r = (expr);                                                                 \
        if (_r != ncclSuccess)                                                                    \
            return fail(std::string(#expr) + ": " + (api->GetErrorString ? api->GetErrorString(_r) : "NCCL error")); \
    } while (0)

// delta = w - base (float4 grid-stride; n4 = elements / 4)
__global__ void delta_kernel(const float4 *__restrict__ w, const float4 *__restrict__ base, float4 *__restrict__ delta, size_t n4)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = w[i], b = base[i];
        delta[i] = make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w);
    }
}
__device__ __forceinline__ float clampc(float v, float c) { return fmaxf(fminf(v, c), -c); }
// base += scale * delta_sum (clamped to +-clip when clip > 0: the persistent grad rows);  w = base
__global__ void apply_kernel(float4 *__restrict__ w, float4 *__restrict__ base, const float4 *__restrict__ delta, float scale, float clip, size_t n4)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 d = delta[i];
        float4 b = base[i];
        b.x = fmaf(scale, d.x, b.x); b.y = fmaf(scale, d.y, b.y); b.z = fmaf(scale, d.z, b.z); b.w = fmaf(scale, d.w, b.w);
        if (clip > 0.f) { b.x = clampc(b.x, clip); b.y = clampc(b.y, clip); b.z = clampc(b.z, clip); b.w = clampc(b.w, clip); }
        base[i] = b;
        w[i] = b;
    }
}

// w = scale * sum (mean mode), clamped to +-clip when clip > 0
__global__ void scale_kernel(float4 *__restrict__ w, const float4 *__restrict__ sum, float scale, float clip, size_t n4)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 d = sum[i];
        float4 b = make_float4(d.x * scale, d.y * scale, d.z * scale, d.w * scale);
        if (clip > 0.f) { b.x = clampc(b.x, clip); b.y = clampc(b.y, clip); b.z = clampc(b.z, clip); b.w = clampc(b.w, clip); }
        w[i] = b;
    }
}

// ---- fused exchange over peer memory ------------------------------------------------------------------------
constexpr int MAX_PEERS = 8;
struct PeerSync {
    int rank, world;
    int slice, slices;                             // this launch reduces rows [slice/slices, (slice+1)/slices) of every array
    float4 *w[MAX_PEERS], *base[MAX_PEERS];        // item table / base copy of every rank (peer-mapped; [rank] = local)
    float4 *w0[MAX_PEERS], *w0b[MAX_PEERS];
    float4 *g[MAX_PEERS], *gb[MAX_PEERS];          // sync_grads: persistent item grad rows / their base copy (null = off)
    float clip;                                    // the grad rows are clamped to +-clip after the reduction
    unsigned long long *flags[MAX_PEERS];          // [2][8] arrival counters of every rank
    unsigned long long seq;                        // this exchange's number (1, 2, ...)
    float scale;
    int mean;                                      // 1: new = scale * sum_q w_q (no base copies involved)
    size_t n4_items, n4_w0;
    unsigned *grid_sync;                           // [0] release flag of the start barrier, [1] blocks done
};

__device__ __forceinline__ float4 ld_sys4(const float4 *p)
{
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_sys4(float4 *p, const float4 &v)
{
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned *p, unsigned v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// all ranks: flags[q][which][rank] = seq on every rank q, then wait until every rank has written mine
__device__ __forceinline__ void peer_barrier(const PeerSync &S, int which)
{
    if ((int)threadIdx.x < S.world) {
        const int q = (int)threadIdx.x;
        __threadfence_system();
        st_release_sys(S.flags[q] + which * MAX_PEERS + S.rank, S.seq);
        while (ld_acquire_sys(S.flags[S.rank] + which * MAX_PEERS + q) < S.seq) { }
    }
    __syncthreads();
}

// One slice of one replicated array.  U float4 per thread and replica are in flight at a time (U x world 16-byte loads,
// world-1 of them over NVLink): the loop is bound by the number of bytes in flight, not by instruction issue.
//   sum  mode (scale 1):      new = base + sum_q (w_q - base)      -> every replica and every base copy
//   mean mode (scale 1/world): new = (1/world) sum_q w_q            -> every replica; algebraically base + mean of the
//                              deltas, so the base copies are neither read nor written (half the remote stores)
template <int U>
__device__ __forceinline__ void peer_reduce_range(const PeerSync &S, float4 *const *w, float4 *const *base, size_t n4, float clip)
{
    const size_t per = (n4 + S.slices - 1) / S.slices;
    const size_t lo = per * S.slice, hi = lo + per < n4 ? lo + per : n4;
    const bool mean = S.mean != 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i0 = lo + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i0 < hi; i0 += stride * U) {
        float4 a[U][MAX_PEERS], b[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = i0 + u * stride;
            if (i < hi) {
                if (!mean) b[u] = base[S.rank][i];
#pragma unroll
                for (int q = 0; q < MAX_PEERS; ++q)
                    if (q < S.world) a[u][q] = ld_sys4(w[q] + i);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = i0 + u * stride;
            if (i >= hi) continue;
            float4 nb;
            if (mean) {
                float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int q = 0; q < MAX_PEERS; ++q)
                    if (q < S.world) { d.x += a[u][q].x; d.y += a[u][q].y; d.z += a[u][q].z; d.w += a[u][q].w; }
                nb = make_float4(d.x * S.scale, d.y * S.scale, d.z * S.scale, d.w * S.scale);
            } else {
                float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int q = 0; q < MAX_PEERS; ++q)
                    if (q < S.world) { d.x += a[u][q].x - b[u].x; d.y += a[u][q].y - b[u].y; d.z += a[u][q].z - b[u].z; d.w += a[u][q].w - b[u].w; }
                nb = make_float4(b[u].x + d.x, b[u].y + d.y, b[u].z + d.z, b[u].w + d.w);
            }
            if (clip > 0.f) { nb.x = clampc(nb.x, clip); nb.y = clampc(nb.y, clip); nb.z = clampc(nb.z, clip); nb.w = clampc(nb.w, clip); }
#pragma unroll
            for (int q = 0; q < MAX_PEERS; ++q)
                if (q < S.world) { st_sys4(w[q] + i, nb); if (!mean) st_sys4(base[q] + i, nb); }
        }
    }
}

// Single-device emulation group (heat_local_group_*): all replicas are plain device pointers of this GPU, one launch
// reduces every row of every array; same arithmetic and rank order as peer_sync_kernel / the NCCL path.
__global__ void __launch_bounds__(256) local_sync_kernel(const PeerSync S)
{
    if (S.w[0]) peer_reduce_range<2>(S, S.w, S.base, S.n4_items, 0.f);
    if (S.w0[0]) peer_reduce_range<1>(S, S.w0, S.w0b, S.n4_w0, 0.f);
    if (S.g[0]) peer_reduce_range<2>(S, S.g, S.gb, S.n4_items, S.clip);
}

__global__ void __launch_bounds__(256) peer_sync_kernel(const PeerSync S)
{
    const unsigned seq32 = (unsigned)S.seq;
    // start barrier: every rank's training kernel has finished (this kernel follows it on the rank's stream)
    if (blockIdx.x == 0) {
        peer_barrier(S, 0);
        if (threadIdx.x == 0) st_release_gpu(S.grid_sync, seq32);
    } else {
        if (threadIdx.x == 0) while (ld_acquire_gpu(S.grid_sync) != seq32) { }
        __syncthreads();
    }
    peer_reduce_range<2>(S, S.w, S.base, S.n4_items, 0.f);
    peer_reduce_range<1>(S, S.w0, S.w0b, S.n4_w0, 0.f);
    if (S.g[0]) peer_reduce_range<2>(S, S.g, S.gb, S.n4_items, S.clip);
    // end barrier: my stores have landed in every replica and every peer's stores have landed in mine before any
    // later kernel of this stream starts; the last block to finish carries it
    __threadfence_system();
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) last = atomicAdd(S.grid_sync + 1, 1u) + 1u == seq32 * gridDim.x;
    __syncthreads();
    if (last) peer_barrier(S, 1);
}

// Map every rank's buffers (CUDA IPC handles all-gathered with NCCL).  Any failure leaves peer_ok false: NCCL path.
int peer_setup(Ctx *c, NcclApi *api)
{
    if (c->world > MAX_PEERS || c->cfg.exchange_mode == 1 || !api->AllGather) return 0;
    // A local failure must not return before the collectives below (the other ranks would wait in them for ever): it is
    // recorded in mine.ok and vetoes the fused path for every rank through the agreement all-reduce.
    bool local_ok = cudaMalloc(&c->xflags, 2 * MAX_PEERS * sizeof(unsigned long long)) == cudaSuccess &&
                    cudaMemset(c->xflags, 0, 2 * MAX_PEERS * sizeof(unsigned long long)) == cudaSuccess &&
                    cudaMalloc(&c->grid_sync, 2 * sizeof(unsigned)) == cudaSuccess &&
                    cudaMemset(c->grid_sync, 0, 2 * sizeof(unsigned)) == cudaSuccess;
    if (!local_ok) cudaGetLastError();
    constexpr int NH = 7;
    struct Export { cudaIpcMemHandle_t h[NH]; int ok; int pad[15]; };
    Export mine{};
    void *ptrs[NH] = {c->item_w, c->item_base, c->w0, c->w0_base, c->xflags, c->item_g, c->item_g_base};
    const int nh = c->item_g_base ? NH : 5;        // the grad rows travel only with sync_grads
    mine.ok = local_ok ? 1 : 0;
    for (int i = 0; i < nh && local_ok; ++i)
        if (cudaIpcGetMemHandle(&mine.h[i], ptrs[i]) != cudaSuccess) { mine.ok = 0; cudaGetLastError(); }
    Export *d_all = nullptr;
    HEAT_CUDA(cudaMalloc(&d_all, sizeof(Export) * c->world));
    HEAT_CUDA(cudaMemcpyAsync(d_all + c->rank, &mine, sizeof(Export), cudaMemcpyHostToDevice, c->stream));
    HEAT_NCCL(api->AllGather(d_all + c->rank, d_all, sizeof(Export), ncclChar, static_cast<ncclComm_t>(c->nccl_comm), c->stream));
    std::vector<Export> all(c->world);
    HEAT_CUDA(cudaMemcpyAsync(all.data(), d_all, sizeof(Export) * c->world, cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(d_all);
    bool ok = true;
    for (int q = 0; q < c->world; ++q) ok = ok && all[q].ok;
    for (int q = 0; q < c->world && ok; ++q) {
        void *m[NH] = {};
        if (q == c->rank) {
            for (int i = 0; i < nh; ++i) m[i] = ptrs[i];
        } else {
            for (int i = 0; i < nh && ok; ++i)
                if (cudaIpcOpenMemHandle(&m[i], all[q].h[i], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = false; cudaGetLastError(); }
        }
        if (!ok) break;
        c->peer_item_g[q] = static_cast<float *>(m[5]); c->peer_item_g_base[q] = static_cast<float *>(m[6]);
        c->peer_item_w[q] = static_cast<float *>(m[0]); c->peer_item_base[q] = static_cast<float *>(m[1]);
        c->peer_w0[q] = static_cast<float *>(m[2]); c->peer_w0_base[q] = static_cast<float *>(m[3]);
        c->peer_flags[q] = static_cast<unsigned long long *>(m[4]);
    }
    // every rank must take the same path: agree on the outcome (a rank that failed to map vetoes)
    double agree = ok ? 0.0 : 1.0;
    HEAT_CUDA(cudaMemcpyAsync(c->loss_pair, &agree, sizeof(double), cudaMemcpyHostToDevice, c->stream));
    HEAT_NCCL(api->AllReduce(c->loss_pair, c->loss_pair, 1, ncclDouble, ncclSum, static_cast<ncclComm_t>(c->nccl_comm), c->stream));
    HEAT_CUDA(cudaMemcpyAsync(&agree, c->loss_pair, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    c->peer_ok = agree == 0.0;
    return 0;
}

int peer_sync(Ctx *c)
{
    PeerSync S{};
    S.rank = c->rank; S.world = c->world; S.slice = c->rank; S.slices = c->world;
    for (int q = 0; q < c->world; ++q) {
        S.w[q] = reinterpret_cast<float4 *>(c->peer_item_w[q]); S.base[q] = reinterpret_cast<float4 *>(c->peer_item_base[q]);
        S.w0[q] = reinterpret_cast<float4 *>(c->peer_w0[q]); S.w0b[q] = reinterpret_cast<float4 *>(c->peer_w0_base[q]);
        S.g[q] = reinterpret_cast<float4 *>(c->peer_item_g[q]); S.gb[q] = reinterpret_cast<float4 *>(c->peer_item_g_base[q]);
        S.flags[q] = c->peer_flags[q];
    }
    S.clip = c->cfg.clip_val;
    const size_t D = c->cfg.emb_dim;
    S.seq = c->syncs + 1;
    S.scale = sync_mean(c) ? 1.0f / (float)c->world : 1.0f;
    S.mean = sync_mean(c) ? 1 : 0;
    S.n4_items = c->cfg.num_items * D / 4; S.n4_w0 = D * D / 4;
    S.grid_sync = c->grid_sync;
    peer_sync_kernel<<<c->num_sms, 256, 0, c->stream>>>(S);       // one resident block per SM: the spin barriers need co-residency
    HEAT_CUDA(cudaGetLastError());
    c->launches += 1;
    c->syncs += 1;
    return 0;
}

}  // namespace

int comm_mode(const Ctx *c) { return (c->world <= 1 || !c->nccl_comm) ? 0 : (c->peer_ok ? 2 : 1); }

int comm_unique_id(uint8_t out[128])
{
    NcclApi *api = nccl();
    if (!api) return fail("NCCL (libnccl.so.2) could not be loaded");
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    HEAT_NCCL(api->GetUniqueId(&id));
    std::memcpy(out, &id, 128);
    return 0;
}

int comm_init(Ctx *c, int rank, int world, const uint8_t uid[128])
{
    NcclApi *api = nccl();
    if (!api) return fail("NCCL (libnccl.so.2) could not be loaded");
    if (world < 1 || rank < 0 || rank >= world) return fail("heat_comm_init: bad rank/world");
    if (!c->item_w) return fail("heat_comm_init: bind the tables first");
    if (c->nccl_comm) return fail("heat_comm_init: already initialised");
    if (c->local_group) return fail("heat_comm_init: the context belongs to a local emulation group");
    ncclUniqueId id;
    std::memcpy(&id, uid, 128);
    ncclComm_t comm;
    HEAT_NCCL(api->CommInitRank(&comm, world, id, rank));
    c->nccl_comm = comm;
    c->rank = rank; c->world = world;
    const size_t D = c->cfg.emb_dim, ib = c->cfg.num_items * D * sizeof(float), wb = D * D * sizeof(float);
    HEAT_CUDA(cudaMalloc(&c->item_base, ib));
    HEAT_CUDA(cudaMalloc(&c->w0_base, wb));
    if (c->cfg.sync_grads) {
        HEAT_CUDA(cudaMalloc(&c->item_g_base, ib));
        HEAT_CUDA(cudaMemcpyAsync(c->item_g_base, c->item_g, ib, cudaMemcpyDeviceToDevice, c->stream));
    }
    HEAT_CUDA(cudaMalloc(&c->delta, ib + wb + (c->cfg.sync_grads ? ib : 0)));      // [item delta | W0 delta | grad-row delta]: one all-reduce
    HEAT_CUDA(cudaMalloc(&c->loss_pair, 2 * sizeof(double)));
    HEAT_CUDA(cudaMemcpyAsync(c->item_base, c->item_w, ib, cudaMemcpyDeviceToDevice, c->stream));
    HEAT_CUDA(cudaMemcpyAsync(c->w0_base, c->w0, wb, cudaMemcpyDeviceToDevice, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    return peer_setup(c, api);
}

// Reconcile the replicated item table and W0 across ranks (see the header comment).
int comm_sync_items(Ctx *c)
{
    if (c->world <= 1 || !c->nccl_comm) return 0;
    if (c->peer_ok) return peer_sync(c);
    NcclApi *api = nccl();
    const size_t D = c->cfg.emb_dim, in4 = c->cfg.num_items * D / 4, wn4 = D * D / 4;
    float4 *delta = reinterpret_cast<float4 *>(c->delta);
    const int blocks = c->num_sms * 8;
    const size_t gn4 = c->item_g_base ? in4 : 0;
    const ncclComm_t comm = static_cast<ncclComm_t>(c->nccl_comm);
    if (sync_mean(c)) {
        // mean mode: new = (1/world) sum_q w_q (== base + mean of the deltas): all-reduce the arrays themselves
        HEAT_CUDA(cudaMemcpyAsync(delta, c->item_w, in4 * 16, cudaMemcpyDeviceToDevice, c->stream));
        HEAT_CUDA(cudaMemcpyAsync(delta + in4, c->w0, wn4 * 16, cudaMemcpyDeviceToDevice, c->stream));
        if (gn4) HEAT_CUDA(cudaMemcpyAsync(delta + in4 + wn4, c->item_g, gn4 * 16, cudaMemcpyDeviceToDevice, c->stream));
        HEAT_NCCL(api->AllReduce(c->delta, c->delta, (in4 + wn4 + gn4) * 4, ncclFloat, ncclSum, comm, c->stream));
        const float scale = 1.0f / (float)c->world;
        scale_kernel<<<blocks, 256, 0, c->stream>>>(reinterpret_cast<float4 *>(c->item_w), delta, scale, 0.f, in4);
        scale_kernel<<<8, 256, 0, c->stream>>>(reinterpret_cast<float4 *>(c->w0), delta + in4, scale, 0.f, wn4);
        if (gn4) scale_kernel<<<blocks, 256, 0, c->stream>>>(reinterpret_cast<float4 *>(c->item_g), delta + in4 + wn4, scale, c->cfg.clip_val, gn4);
        HEAT_CUDA(cudaGetLastError());
        c->launches += gn4 ? 3 : 2;
    } else {
        delta_kernel<<<blocks, 256, 0, c->stream>>>(reinterpret_cast<const float4 *>(c->item_w), reinterpret_cast<const float4 *>(c->item_base), delta, in4);
        delta_kernel<<<8, 256, 0, c->stream>>>(reinterpret_cast<const float4 *>(c->w0), reinterpret_cast<const float4 *>(c->w0_base), delta + in4, wn4);
        if (gn4) delta_kernel<<<blocks, 256, 0, c->stream>>>(reinterpret_cast<const float4 *>(c->item_g), reinterpret_cast<const float4 *>(c->item_g_base), delta + in4 + wn4, gn4);
        HEAT_CUDA(cudaGetLastError());
        HEAT_NCCL(api->AllReduce(c->delta, c->delta, (in4 + wn4 + gn4) * 4, ncclFloat, ncclSum, comm, c->stream));
        apply_kernel<<<blocks, 256, 0, c->stream>>>(reinterpret_cast<float4 *>(c->item_w), reinterpret_cast<float4 *>(c->item_base), delta, 1.0f, 0.f, in4);
        apply_kernel<<<8, 256, 0, c->stream>>>(reinterpret_cast<float4 *>(c->w0), reinterpret_cast<float4 *>(c->w0_base), delta + in4, 1.0f, 0.f, wn4);
        if (gn4) apply_kernel<<<blocks, 256, 0, c->stream>>>(reinterpret_cast<float4 *>(c->item_g), reinterpret_cast<float4 *>(c->item_g_base), delta + in4 + wn4, 1.0f, c->cfg.clip_val, gn4);
        HEAT_CUDA(cudaGetLastError());
        c->launches += gn4 ? 6 : 4;
    }
    c->syncs += 1;
    return 0;
}

// ---- single-device emulation group ---------------------------------------------------------------------------
int local_group_init(Ctx **cs, int n)
{
    if (n < 1 || n > MAX_PEERS) return fail("heat_local_group_init: 1..8 contexts");
    for (int q = 0; q < n; ++q) {
        Ctx *c = cs[q];
        if (!c || !c->item_w) return fail("heat_local_group_init: bind the tables of every context first");
        if (c->nccl_comm) return fail("heat_local_group_init: a context already has a communicator");
        if (c->device != cs[0]->device || c->cfg.emb_dim != cs[0]->cfg.emb_dim || c->cfg.num_items != cs[0]->cfg.num_items)
            return fail("heat_local_group_init: the contexts must share device, emb_dim and num_items");
    }
    const size_t D = cs[0]->cfg.emb_dim, ib = cs[0]->cfg.num_items * D * sizeof(float), wb = D * D * sizeof(float);
    for (int q = 0; q < n; ++q) {
        Ctx *c = cs[q];
        if (!c->item_base) { HEAT_CUDA(cudaMalloc(&c->item_base, ib)); HEAT_CUDA(cudaMalloc(&c->w0_base, wb)); }
        if (cs[0]->cfg.sync_grads && !c->item_g_base) HEAT_CUDA(cudaMalloc(&c->item_g_base, ib));
        HEAT_CUDA(cudaMemcpyAsync(c->item_base, c->item_w, ib, cudaMemcpyDeviceToDevice, c->stream));
        HEAT_CUDA(cudaMemcpyAsync(c->w0_base, c->w0, wb, cudaMemcpyDeviceToDevice, c->stream));
        if (c->item_g_base) HEAT_CUDA(cudaMemcpyAsync(c->item_g_base, c->item_g, ib, cudaMemcpyDeviceToDevice, c->stream));
        HEAT_CUDA(cudaStreamSynchronize(c->stream));
        c->local_group = true; c->rank = q; c->world = n;
    }
    return 0;
}

int local_group_share(Ctx **cs, int n, uint32_t mask)
{
    for (int q = 0; q < n; ++q)
        if (!cs[q] || !cs[q]->local_group) return fail("heat_local_group_share: call heat_local_group_init first");
    if (cs[0]->group_shared) return fail("heat_local_group_share: already shared");
    for (int q = 1; q < n; ++q) {
        Ctx *c = cs[q];
        if (mask & 1u) { cudaFree(c->w0); c->w0 = cs[0]->w0; }
        if (mask & 2u) { cudaFree(c->item_g); c->item_g = cs[0]->item_g; }
        if (mask & 4u) { cudaFree(c->item_w); c->item_w = cs[0]->item_w; }
        c->shared_mask = mask;
    }
    for (int q = 0; q < n; ++q) cs[q]->group_shared = mask;
    return 0;
}

// all member streams are idle when this is called (heat_local_group_train_epoch synchronises them); runs on cs[0]'s stream
int local_group_sync(Ctx **cs, int n)
{
    Ctx *c0 = cs[0];
    PeerSync S{};
    S.rank = 0; S.world = n; S.slice = 0; S.slices = 1;
    const uint32_t sh = c0->group_shared;          // arrays the members share physically need no reconciliation
    const bool grads = c0->cfg.sync_grads && !(sh & 2u);
    for (int q = 0; q < n; ++q) {
        S.w[q] = (sh & 4u) ? nullptr : reinterpret_cast<float4 *>(cs[q]->item_w); S.base[q] = reinterpret_cast<float4 *>(cs[q]->item_base);
        S.w0[q] = (sh & 1u) ? nullptr : reinterpret_cast<float4 *>(cs[q]->w0); S.w0b[q] = reinterpret_cast<float4 *>(cs[q]->w0_base);
        S.g[q] = grads ? reinterpret_cast<float4 *>(cs[q]->item_g) : nullptr;
        S.gb[q] = grads ? reinterpret_cast<float4 *>(cs[q]->item_g_base) : nullptr;
    }
    const size_t D = c0->cfg.emb_dim;
    S.scale = sync_mean(c0) ? 1.0f / (float)n : 1.0f;
    S.mean = sync_mean(c0) ? 1 : 0;
    S.clip = c0->cfg.clip_val;
    S.n4_items = c0->cfg.num_items * D / 4; S.n4_w0 = D * D / 4;
    local_sync_kernel<<<c0->num_sms * 4, 256, 0, c0->stream>>>(S);
    HEAT_CUDA(cudaGetLastError());
    c0->launches += 1;
    for (int q = 0; q < n; ++q) cs[q]->syncs += 1;
    return 0;
}

// sum over ranks of (loss_sum, samples): engine.cpp:380-383
int comm_reduce_loss(Ctx *c, double *loss_sum, double *samples)
{
    if (c->world <= 1 || !c->nccl_comm) return 0;
    NcclApi *api = nccl();
    double pair[2] = {*loss_sum, *samples};
    HEAT_CUDA(cudaMemcpyAsync(c->loss_pair, pair, sizeof(pair), cudaMemcpyHostToDevice, c->stream));
    HEAT_NCCL(api->AllReduce(c->loss_pair, c->loss_pair, 2, ncclDouble, ncclSum, static_cast<ncclComm_t>(c->nccl_comm), c->stream));
    HEAT_CUDA(cudaMemcpyAsync(pair, c->loss_pair, sizeof(pair), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    *loss_sum = pair[0]; *samples = pair[1];
    return 0;
}

// max over ranks of one count (so that every rank runs the same number of sync segments per epoch)
int comm_max_u64(Ctx *c, uint64_t *v)
{
    if (c->world <= 1 || !c->nccl_comm) return 0;
    NcclApi *api = nccl();
    double x = (double)*v;
    HEAT_CUDA(cudaMemcpyAsync(c->loss_pair, &x, sizeof(x), cudaMemcpyHostToDevice, c->stream));
    HEAT_NCCL(api->AllReduce(c->loss_pair, c->loss_pair, 1, ncclDouble, ncclMax, static_cast<ncclComm_t>(c->nccl_comm), c->stream));
    HEAT_CUDA(cudaMemcpyAsync(&x, c->loss_pair, sizeof(x), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    *v = (uint64_t)x;
    return 0;
}

// sum over ranks of n doubles on the host (evaluation metric sums)
int comm_allreduce_f64(Ctx *c, double *vals, int n)
{
    if (c->world <= 1 || !c->nccl_comm) return 0;
    if (n > 16) return fail("heat_comm_allreduce_f64: at most 16 values");
    NcclApi *api = nccl();
    double *buf = nullptr;
    HEAT_CUDA(cudaMalloc(&buf, 16 * sizeof(double)));
    cudaError_t e = cudaMemcpyAsync(buf, vals, n * sizeof(double), cudaMemcpyHostToDevice, c->stream);
    ncclResult_t r = ncclSuccess;
    if (e == cudaSuccess) r = api->AllReduce(buf, buf, n, ncclDouble, ncclSum, static_cast<ncclComm_t>(c->nccl_comm), c->stream);
    if (e == cudaSuccess && r == ncclSuccess) e = cudaMemcpyAsync(vals, buf, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(buf);
    if (r != ncclSuccess) return fail("ncclAllReduce failed");
    if (e != cudaSuccess) return fail(std::string("heat_comm_allreduce_f64: ") + cudaGetErrorString(e));
    return 0;
}

void comm_destroy(Ctx *c)
{
    if (c->nccl_comm) {
        NcclApi *api = nccl();
        if (api) api->CommDestroy(static_cast<ncclComm_t>(c->nccl_comm));
        c->nccl_comm = nullptr;
    }
    if (c->peer_ok) {
        for (int q = 0; q < c->world; ++q) {
            if (q == c->rank) continue;
            if (c->peer_item_g[q]) { cudaIpcCloseMemHandle(c->peer_item_g[q]); cudaIpcCloseMemHandle(c->peer_item_g_base[q]); }
            cudaIpcCloseMemHandle(c->peer_item_w[q]); cudaIpcCloseMemHandle(c->peer_item_base[q]);
            cudaIpcCloseMemHandle(c->peer_w0[q]); cudaIpcCloseMemHandle(c->peer_w0_base[q]); cudaIpcCloseMemHandle(c->peer_flags[q]);
        }
        c->peer_ok = false;
    }
    cudaFree(c->delta); cudaFree(c->loss_pair); cudaFree(c->xflags); cudaFree(c->grid_sync); cudaFree(c->item_g_base);
    c->item_g_base = nullptr;
    c->delta = nullptr; c->loss_pair = nullptr; c->xflags = nullptr; c->grid_sync = nullptr;
}

}  // namespace heat
