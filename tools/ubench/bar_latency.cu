// Release latency of the block-level synchronisation primitives the CTA training kernel could use:
// named bar.sync over 10 of 11 warps, an mbarrier, and a shared-memory sequence flag (one producer warp).
// Each round every warp spins a warp-specific delay, stamps its arrival, synchronises, then stamps the
// time a dependent shared-memory load returns.  Reported: mean (release - last arrival) over the rounds.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/ubench/bar_latency.bin tools/ubench/bar_latency.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ long long clk() { long long t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ void spin(int c) { long long t0 = clk(); while (clk() - t0 < c) {} }

constexpr int NW = 10, ROUNDS = 2000;

__global__ void __launch_bounds__(352, 1) k_bar(int mode, float *out)
{
    __shared__ long long arrive[NW];
    __shared__ volatile int data[NW];
    __shared__ unsigned long long mbar;
    __shared__ volatile unsigned flag;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        unsigned a = (unsigned)__cvta_generic_to_shared(&mbar);
        asm volatile("mbarrier.init.shared.b64 [%0], %1;" ::"r"(a), "r"(NW * 32));
        flag = 0;
    }
    __syncthreads();
    if (w >= NW) return;
    double acc = 0.0;
    unsigned phase = 0;
    for (int r = 0; r < ROUNDS; ++r) {
        spin(50 + 37 * ((w * 7 + r) % NW));
        if (lane == 0) { data[w] = r; arrive[w] = clk(); }
        long long t1;
        if (mode == 0) {
            asm volatile("bar.sync 1, %0;" ::"n"(NW * 32) : "memory");
        } else if (mode == 1) {
            unsigned a = (unsigned)__cvta_generic_to_shared(&mbar);
            asm volatile("{ .reg .b64 s; mbarrier.arrive.shared.b64 s, [%0]; }" ::"r"(a) : "memory");
            asm volatile("{ .reg .pred p; W: mbarrier.try_wait.parity.shared.b64 p, [%0], %1; @!p bra W; }" ::"r"(a), "r"(phase) : "memory");
            phase ^= 1;
        } else if (mode == 2) {
            // single producer (warp 0, after a fixed long delay) -> consumers poll a sequence flag
            if (w == 0) { spin(400); if (lane == 0) { arrive[0] = clk(); __threadfence_block(); flag = r + 1; } __syncwarp(); }
            else { while (flag != (unsigned)(r + 1)) {} }
        } else {
            // single producer through bar.arrive / bar.sync (the kernel's mean barrier)
            if (w == 0) { spin(400); if (lane == 0) arrive[0] = clk(); __syncwarp(); asm volatile("bar.arrive 3, %0;" ::"n"(NW * 32) : "memory"); }
            else asm volatile("bar.sync 3, %0;" ::"n"(NW * 32) : "memory");
        }
        const int v = data[(w + 1) % NW];                      // dependent load: the deferred block fires here
        t1 = clk() + (v == -1);
        if (mode >= 2) {
            if (w != 0 && lane == 0) acc += (double)(t1 - arrive[0]);
        } else if (lane == 0) {
            long long last = 0;
            for (int i = 0; i < NW; ++i) last = max(last, arrive[i]);
            acc += (double)(t1 - last);
        }
        asm volatile("bar.sync 2, %0;" ::"n"(NW * 32) : "memory");   // round separator
    }
    if (lane == 0 && blockIdx.x == 0) out[mode * 16 + w] = (float)(acc / ROUNDS);
}

int main()
{
    float *d, h[64];
    cudaMalloc(&d, sizeof(h));
    cudaMemset(d, 0, sizeof(h));
    for (int m = 0; m < 4; ++m) k_bar<<<148, 352>>>(m, d);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    const char *names[4] = {"bar.sync(10 warps)", "mbarrier", "flag poll (1 producer)", "bar.arrive/sync (1 producer)"};
    for (int m = 0; m < 4; ++m) {
        printf("%-30s", names[m]);
        for (int w = 0; w < NW; ++w) printf(" %6.1f", h[m * 16 + w]);
        printf("\n");
    }
    return 0;
}
