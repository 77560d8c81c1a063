// heat_b200/csrc/probe_tf32.cu -- measured dense TF32 tensor peak: the roofline denominator of the evaluation kernel.
//
// MEASURED_PEAKS.json holds a bf16 GEMM figure only; the score GEMM of eval_kernel.cu runs `tcgen05.mma.kind::tf32`.
// This probe issues nothing but that instruction: one CTA per SM, one elected thread, M=128 x N=256 x K=8 MMAs from two
// resident shared-memory tiles (one 128-byte swizzle atom of A, one of B) into two alternating TMEM accumulators,
// back to back, one commit at the end.  flops = 2*128*256*8 per instruction.  It is the ceiling of the instruction
// stream the evaluation kernel uses (cta_group::1, operands in shared memory), not a vendor number.
#include "heat_ctx.cuh"
#include "sm100_ptx.cuh"
#include <cstdlib>
#include <string>

namespace heat {
namespace {

constexpr int PM = 128, PN_MAX = 256;
constexpr uint32_t P_TMEM_COLS = 512;

template <int PN, bool A_TMEM>
__global__ void __launch_bounds__(128, 1) tf32_peak_kernel(int iters, unsigned *timeouts)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    float *sA = reinterpret_cast<float *>(smem);                       // [128 rows][32 tf32]  16 KB
    float *sB = reinterpret_cast<float *>(smem + PM * 128);            // [PN rows][32 tf32]
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + (PM + PN_MAX) * 128);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar + 1);

    for (int i = threadIdx.x; i < (PM + PN_MAX) * 32; i += blockDim.x) {
        unsigned h = (unsigned)i * 2654435761u + blockIdx.x;
        h ^= h >> 15; h *= 0x2c1b3c6du; h ^= h >> 12;
        sA[i] = __uint_as_float(0x3f000000u | (h & 0x007fe000u)) - 0.75f;      // [-0.25, 0.25), already tf32-representable
    }
    if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
    if (threadIdx.x < 32) ptx::tmem_alloc<P_TMEM_COLS>(tmem_slot);
    ptx::fence_proxy_async();                                           // generic-proxy tile writes -> async proxy (MMA reads)
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    if (threadIdx.x == 0) {
        constexpr uint32_t IDESC = ptx::idesc_tf32(PM, PN);
        const uint32_t a = ptx::smem_u32(sA), b = ptx::smem_u32(sB);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int acc = 0; acc < 2; ++acc)
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                    if (A_TMEM) ptx::mma_tf32_ts(tmem + acc * PN, tmem + 2 * PN_MAX - 64 + kk * 8, ptx::smem_desc_k_sw128(b + kk * 32), IDESC, (it | kk) != 0);
                    else ptx::mma_tf32(tmem + acc * PN, ptx::smem_desc_k_sw128(a + kk * 32), ptx::smem_desc_k_sw128(b + kk * 32), IDESC, (it | kk) != 0);
        }
        ptx::mma_commit(bar);
        // bounded wait: a probe must never hang the GPU
        bool done = false;
        for (unsigned tries = 0; tries < (1u << 22) && !done; ++tries) {
            unsigned ok;
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\nselp.u32 %0, 1, 0, p;\n}\n"
                         : "=r"(ok) : "r"(ptx::smem_u32(bar)), "r"(0u), "r"(100000u) : "memory");
            done = ok != 0;
        }
        if (!done) atomicAdd(timeouts, 1u);
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) ptx::tmem_dealloc<P_TMEM_COLS>(tmem);
}

}  // namespace

template <int PN, bool A_TMEM> int probe_tf32_variant(int device, int iters, double *tflops);

int probe_tf32(int device, int iters, double *tflops)
{
    if (const char *e = std::getenv("HEAT_PROBE_TF32_VARIANT")) {      // experiment: N = 64 / 128 / 256, A from TMEM (suffix t)
        const std::string v(e);
        if (v == "64") return probe_tf32_variant<64, false>(device, iters, tflops);
        if (v == "128") return probe_tf32_variant<128, false>(device, iters, tflops);
        if (v == "64t") return probe_tf32_variant<64, true>(device, iters, tflops);
        if (v == "128t") return probe_tf32_variant<128, true>(device, iters, tflops);
        if (v == "256t") return probe_tf32_variant<224, true>(device, iters, tflops);
        if (v == "96t") return probe_tf32_variant<96, true>(device, iters, tflops);
        if (v == "192t") return probe_tf32_variant<192, true>(device, iters, tflops);
    }
    return probe_tf32_variant<256, false>(device, iters, tflops);
}

template <int PN, bool A_TMEM> int probe_tf32_variant(int device, int iters, double *tflops)
{
    if (!tflops) return fail("heat_probe_tf32_peak: null argument");
    HEAT_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop{};
    HEAT_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail("heat_probe_tf32_peak: needs an sm_100 device");
    if (iters <= 0) iters = 4000;
    unsigned *timeouts = nullptr;
    HEAT_CUDA(cudaMalloc(&timeouts, sizeof(unsigned)));
    HEAT_CUDA(cudaMemset(timeouts, 0, sizeof(unsigned)));
    const size_t smem = 1024 + (PM + PN_MAX) * 128 + 64;
    auto kern = tf32_peak_kernel<PN, A_TMEM>;
    HEAT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1;
    HEAT_CUDA(cudaEventCreate(&e0)); HEAT_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    int rc = 0;
    for (int rep = 0; rep < 4 && !rc; ++rep) {
        cudaEventRecord(e0);
        kern<<<prop.multiProcessorCount, 128, smem>>>(iters, timeouts);
        cudaEventRecord(e1);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) { rc = fail(std::string("heat_probe_tf32_peak: ") + cudaGetErrorString(e)); break; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = (double)prop.multiProcessorCount * iters * 8.0 * 2.0 * PM * PN * 8;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    unsigned h_to = 0;
    cudaMemcpy(&h_to, timeouts, sizeof(unsigned), cudaMemcpyDeviceToHost);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(timeouts);
    if (rc) return rc;
    if (h_to) return fail("heat_probe_tf32_peak: the MMA stream did not complete in time");
    *tflops = best;
    return 0;
}

}  // namespace heat
