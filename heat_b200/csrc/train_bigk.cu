// heat_b200/csrc/train_bigk.cu -- launches the large-negative training kernel (train_kernel_bigk.cuh).
#include "heat_ctx.cuh"
#include "train_kernel_bigk.cuh"
#include <algorithm>

namespace heat {

namespace {

template <int D>
int launch_bigk(Ctx *c, TrainParams P, int blocks)
{
    using C = BigKCfg<D>;
    auto kern = train_bigk_kernel<D>;
    const BigKLayout L = bigk_layout(D, C::G, P.K, C::RING);
    const int max_smem = 232448;                              // 227 KB: the per-block opt-in limit of sm_100
    if (L.off_cache > max_smem) return fail("num_negs too large for the large-negative kernel's shared-memory tables");
    P.cache_rows = std::min(P.K, (max_smem - L.off_cache) / (D * 4));
    const size_t smem = (size_t)L.off_cache + (size_t)P.cache_rows * D * 4;
    HEAT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    HEAT_CUDA(cudaEventRecord(c->ev0, c->stream));
    kern<<<blocks, C::NT, smem, c->stream>>>(P);
    HEAT_CUDA(cudaGetLastError());
    HEAT_CUDA(cudaEventRecord(c->ev1, c->stream));
    c->launches += 1;
    return 0;
}

// Large-negative engine (train_kernel_bigk.cuh): any num_negs up to 16384, emb_dim 32/64/128.
}  // namespace

int dispatch_bigk(Ctx *c, const TrainParams &P, bool deterministic)
{
    const int D = (int)c->cfg.emb_dim;
    int blocks = deterministic ? 1 : c->num_sms;
    if (c->cfg.max_blocks && blocks > (int)c->cfg.max_blocks) blocks = (int)c->cfg.max_blocks;
    if (P.K > 16384) return fail("num_negs > 16384 is not supported");
    if (D == 128) return launch_bigk<128>(c, P, blocks);
    if (D == 64) return launch_bigk<64>(c, P, blocks);
    if (D == 32) return launch_bigk<32>(c, P, blocks);
    return fail("unsupported emb_dim " + std::to_string(D) + ": the training kernels cover emb_dim 32, 64 and 128");
}

}  // namespace heat
