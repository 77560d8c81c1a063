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
// NCCL is resolved at run time (dlopen "libnccl.so.2"): inside a PyTorch process that is the
// library torch already loaded, so both share one NCCL; in a plain C++ host it is the system one.
#include <dlfcn.h>
#include <nccl.h>
#include <cstring>
#include "heat_ctx.cuh"

namespace heat {
namespace {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
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
// base += scale * delta_sum;  w = base
__global__ void apply_kernel(float4 *__restrict__ w, float4 *__restrict__ base, const float4 *__restrict__ delta, float scale, size_t n4)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 d = delta[i];
        float4 b = base[i];
        b.x = fmaf(scale, d.x, b.x); b.y = fmaf(scale, d.y, b.y); b.z = fmaf(scale, d.z, b.z); b.w = fmaf(scale, d.w, b.w);
        base[i] = b;
        w[i] = b;
    }
}

}  // namespace

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
    ncclUniqueId id;
    std::memcpy(&id, uid, 128);
    ncclComm_t comm;
    HEAT_NCCL(api->CommInitRank(&comm, world, id, rank));
    c->nccl_comm = comm;
    c->rank = rank; c->world = world;
    const size_t D = c->cfg.emb_dim, ib = c->cfg.num_items * D * sizeof(float), wb = D * D * sizeof(float);
    HEAT_CUDA(cudaMalloc(&c->item_base, ib));
    HEAT_CUDA(cudaMalloc(&c->w0_base, wb));
    HEAT_CUDA(cudaMalloc(&c->delta, ib + wb));      // item delta followed by the W0 delta: one all-reduce
    HEAT_CUDA(cudaMalloc(&c->loss_pair, 2 * sizeof(double)));
    HEAT_CUDA(cudaMemcpyAsync(c->item_base, c->item_w, ib, cudaMemcpyDeviceToDevice, c->stream));
    HEAT_CUDA(cudaMemcpyAsync(c->w0_base, c->w0, wb, cudaMemcpyDeviceToDevice, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

// Reconcile the replicated item table and W0 across ranks (see the header comment).
int comm_sync_items(Ctx *c)
{
    if (c->world <= 1 || !c->nccl_comm) return 0;
    NcclApi *api = nccl();
    const size_t D = c->cfg.emb_dim, in4 = c->cfg.num_items * D / 4, wn4 = D * D / 4;
    float4 *delta = reinterpret_cast<float4 *>(c->delta);
    const int blocks = c->num_sms * 8;
    delta_kernel<<<blocks, 256, 0, c->stream>>>(reinterpret_cast<const float4 *>(c->item_w), reinterpret_cast<const float4 *>(c->item_base), delta, in4);
    delta_kernel<<<8, 256, 0, c->stream>>>(reinterpret_cast<const float4 *>(c->w0), reinterpret_cast<const float4 *>(c->w0_base), delta + in4, wn4);
    HEAT_CUDA(cudaGetLastError());
    HEAT_NCCL(api->AllReduce(c->delta, c->delta, (in4 + wn4) * 4, ncclFloat, ncclSum, static_cast<ncclComm_t>(c->nccl_comm), c->stream));
    const float scale = c->cfg.sync_average ? 1.0f / (float)c->world : 1.0f;
    apply_kernel<<<blocks, 256, 0, c->stream>>>(reinterpret_cast<float4 *>(c->item_w), reinterpret_cast<float4 *>(c->item_base), delta, scale, in4);
    apply_kernel<<<8, 256, 0, c->stream>>>(reinterpret_cast<float4 *>(c->w0), reinterpret_cast<float4 *>(c->w0_base), delta + in4, scale, wn4);
    HEAT_CUDA(cudaGetLastError());
    c->launches += 4;
    c->syncs += 1;
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
    cudaFree(c->delta); cudaFree(c->loss_pair);
    c->delta = nullptr; c->loss_pair = nullptr;
}

}  // namespace heat
