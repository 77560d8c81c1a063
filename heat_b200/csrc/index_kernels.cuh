// heat_b200/csrc/index_kernels.cuh -- kernels behind heat_sample_negatives / heat_history_indices.
// They call the SAME device functions as the training kernel (sample_negative,
// load_history_ids + shuffle), so the bit-exact index tests exercise the training path's code.
#pragma once
#include "train_kernel.cuh"

namespace heat {

// ---- index-contract kernels (tests): same device functions as the training kernel ----
__global__ void sample_negatives_kernel(const TrainParams P, unsigned long long *out)
{
    const long long n = P.end - P.begin;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n * P.K; i += (long long)gridDim.x * blockDim.x) {
        const long long s = P.begin + i / P.K;
        const int k = (int)(i % P.K);
        out[i] = (unsigned long long)sample_negative(P, s, k, __ldg(&P.clicks[s]).y);
    }
}

// one warp per sample, the training kernel's id fetch (load_history_ids + shuffle)
__global__ void history_indices_kernel(const TrainParams P, unsigned long long *out_ids, unsigned long long *out_cnt)
{
    const int lane = threadIdx.x & 31;
    const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long s = P.begin + w; s < P.end; s += nw) {
        const int u = __ldg(&P.clicks[s]).x;
        const int H = __ldg(P.masks + u);
        int hid[4];
        load_history_ids(P, u, lane, hid);
        unsigned long long *row = out_ids + (s - P.begin) * P.max_his;
        for (int h0 = 0; h0 < P.max_his; h0 += 32) {
            const int hq = pick4(hid, h0 >> 5);
            const int h = h0 + lane;
            const int id = __shfl_sync(FULL, hq, h & 31);
            if (h < P.max_his) row[h] = (h < H) ? (unsigned long long)id : ~0ULL;
        }
        if (lane == 0) out_cnt[s - P.begin] = (unsigned long long)H;
    }
}

}  // namespace heat
