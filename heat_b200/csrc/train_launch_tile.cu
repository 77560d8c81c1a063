// heat_b200/csrc/train_launch_tile.cu -- the CTA-per-sample kernel with the hot-tile negative sampler compiled in
// (heat_config.tile_sampler; philox.cuh).  A separate translation unit: the default instantiations in
// train_launch.cu stay free of the tile code (it cost the default path ~1 % when it was a run-time branch) and the
// two files compile in parallel.
#include "heat_ctx.cuh"
#include "train_kernel_cta.cuh"

namespace heat {

namespace {

template <int D, int NPS>
int launch_cta_tile(Ctx *c, const TrainParams &P, int blocks)
{
    using C = CtaCfg<D, NPS>;
    auto kern = train_cta_kernel<D, NPS, true>;
    const size_t smem = (size_t)C::FLOATS * sizeof(float);
    HEAT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    HEAT_CUDA(cudaEventRecord(c->ev0, c->stream));
    kern<<<blocks, C::NT, smem, c->stream>>>(P);
    HEAT_CUDA(cudaGetLastError());
    HEAT_CUDA(cudaEventRecord(c->ev1, c->stream));
    c->launches += 1;
    return 0;
}

}  // namespace

int dispatch_cta_tile(Ctx *c, const TrainParams &P, int blocks)
{
    const int D = (int)c->cfg.emb_dim, K = P.K;
    if (D == 64 && K <= 16) return launch_cta_tile<64, 1>(c, P, blocks);
    if (D == 64 && K <= 32) return launch_cta_tile<64, 2>(c, P, blocks);
    if (D == 64 && K <= 64) return launch_cta_tile<64, 4>(c, P, blocks);
    if (D == 32 && K <= 16) return launch_cta_tile<32, 1>(c, P, blocks);
    if (D == 32 && K <= 32) return launch_cta_tile<32, 2>(c, P, blocks);
    return -1;
}

}  // namespace heat
