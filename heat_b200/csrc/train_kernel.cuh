// heat_b200/csrc/train_kernel.cuh -- fused forward + backward + clipped-SGD kernel, sm_100a.
//
// One WARP owns one sample at a time; a warp takes `chunk` consecutive samples per
// scheduling step from a global counter (the GPU form of the reference's
// `#pragma omp for schedule(dynamic,512)`, train/engine.cpp:327) and walks them in file
// order, so a user's consecutive samples are applied sequentially exactly as one OpenMP
// thread would apply them.  Different warps race on shared rows Hogwild-style, as the
// reference's threads do (ICS23-HEAT.pdf section 4.5).
//
// What one sample does (reference: models/matrix_factorization.cpp:15-181,
// behavior_aggregators/behavior_aggregators.cpp:51-153, optimizers/sgd.cpp:14-26):
//   gather U[u], I[p], the K negative rows, their persistent grad rows and the H history
//   rows (float4 per lane, D/4 lanes per row, 32/(D/4) rows per warp-wide load, L2-coherent
//   ld.global.cg); mean -> mean.W0 (W0 staged in shared memory) -> u = 0.4u + 0.6f;
//   warp-shuffle dots, cosine, temperature-0.07 softmax loss with the implicit zero logit;
//   closed-form cosine grads reusing the forward dots/norms; g = clamp(g_prev + dg);
//   w -= lr*g; both rows written back; W0 outer-product grads buffered per warp in a
//   32-slot shared-memory ring and applied every 32 samples (behavior_aggregators.cpp:141-146).
//
// Layout: tables are row-major fp32 [rows, D]; ids int32; one row = D*4 bytes = one or two
// 128-byte lines, so every warp-wide LDG.128 moves whole lines.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "philox.cuh"

namespace heat {

struct TrainParams {
    const int2 *clicks;      // [N] (user, item)
    const int *his;          // [U, max_his]
    const int *masks;        // [U]
    float *user_w, *item_w;  // [U,D], [I,D]
    float *user_g, *item_g;  // persistent grad rows, zero at epoch start
    float *w0;               // [D,D] row-major
    unsigned long long *chunk_counter;
    double *loss_sum;
    float *loss_out;         // optional [end-begin]
    long long begin, end;    // local sample range
    unsigned long long sample_offset, seed;
    unsigned epoch;
    int max_his, num_items, K, chunk, drop_w0_tail;
    long long tail_begin;    // guided scheduling: [begin, tail_begin) is handed out in chunks of `chunk` samples, [tail_begin, end)
    int tail_chunk;          // in chunks of `tail_chunk` (small): the warps / CTAs of a launch finish within a few samples of each other
    int upd;                 // item-row update mode, see UPD_* below
    int no_cache;            // 1: gather the history rows for every sample (history_mode 1)
    int cache_rows;          // train_kernel_bigk.cuh: negative rows kept in shared memory between the two passes
    float lr, lr_w0, clip;
    unsigned tile_size;                    // hot-tile sampler (philox.cuh): 0 = uniform over all items
    unsigned long long refresh_interval;   // samples per tile (>= 1)
};

constexpr unsigned FULL = 0xffffffffu;

// chunk number c of a launch -> its sample range (false: no such chunk).  Chunks are consecutive ranges in file order.
__device__ __forceinline__ bool chunk_range(const TrainParams &P, long long c, long long &s0, long long &s1)
{
    const long long nb = (P.tail_begin - P.begin + P.chunk - 1) / P.chunk;
    if (c < nb) {
        s0 = P.begin + c * P.chunk;
        s1 = min(P.tail_begin, s0 + (long long)P.chunk);
        return true;
    }
    s0 = P.tail_begin + (c - nb) * P.tail_chunk;
    s1 = min(P.end, s0 + (long long)P.tail_chunk);
    return s0 < P.end;
}

// How a sample's item rows (positive + negatives: weight row and persistent grad row) are written back.
//   UPD_STORE  the reference's own sequence (matrix_factorization.cpp:133-149,169-174): the row
//              snapshot taken at gather time is stepped and stored.  Exact in sequential order;
//              under thousands of concurrent warps the long read-modify-write window loses updates.
//   UPD_FRESH  the GRAD rows are read only right before the step (g = clamp(g_now + dg), stored) and the
//              weight rows are stepped by an L2 vector reduction (w += -lr*g): the read-modify-write
//              window of g shrinks from the whole sample to one L2 round trip and no weight step is ever
//              lost.  Same result as UPD_STORE when nothing races (up to the rounding of w - lr*g).
//   UPD_ATOMIC the step is applied as an L2 vector reduction (w += -lr*g_new, g += g_new - g_old):
//              no update is ever lost, whatever the concurrency.
enum { UPD_STORE = 0, UPD_FRESH = 1, UPD_ATOMIC = 2 };

__device__ __forceinline__ float4 ldcg4(const float *p) { return __ldcg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ void stcg4(float *p, const float4 &v) { __stcg(reinterpret_cast<float4 *>(p), v); }
// vector reduction at L2 (REDG.E.ADD.F32x4): the row update of the atomic Hogwild mode
__device__ __forceinline__ void red4(float *p, const float4 &v)
{
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 f4zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
// float4 arithmetic on Blackwell's packed fp32x2 pipe (FADD2 / FMUL2 / FFMA2: two IEEE fp32 results per
// issued instruction, same rounding as the scalar forms), so a row slice costs 2 issue slots, not 4.
__device__ __forceinline__ float2 lo2(const float4 &a) { return make_float2(a.x, a.y); }
__device__ __forceinline__ float2 hi2(const float4 &a) { return make_float2(a.z, a.w); }
__device__ __forceinline__ float4 cat4(const float2 &l, const float2 &h) { return make_float4(l.x, l.y, h.x, h.y); }
__device__ __forceinline__ float4 operator+(const float4 &a, const float4 &b) { return cat4(__fadd2_rn(lo2(a), lo2(b)), __fadd2_rn(hi2(a), hi2(b))); }
__device__ __forceinline__ float4 operator-(const float4 &a, const float4 &b)
{
    const float2 m1 = make_float2(-1.f, -1.f);      // a - b = fma(b, -1, a): exact product, one rounding
    return cat4(__ffma2_rn(lo2(b), m1, lo2(a)), __ffma2_rn(hi2(b), m1, hi2(a)));
}
__device__ __forceinline__ float4 operator*(float s, const float4 &a)
{
    const float2 s2 = make_float2(s, s);
    return cat4(__fmul2_rn(s2, lo2(a)), __fmul2_rn(s2, hi2(a)));
}
// x*x' + y*y' + z*z' + w*w' as (x*x' + z*z') + (y*y' + w*w')
__device__ __forceinline__ float dot4(const float4 &a, const float4 &b)
{
    const float2 t = __ffma2_rn(hi2(a), hi2(b), __fmul2_rn(lo2(a), lo2(b)));
    return t.x + t.y;
}
__device__ __forceinline__ void fma4(float4 &acc, float s, const float4 &a)
{
    const float2 s2 = make_float2(s, s);
    acc = cat4(__ffma2_rn(s2, lo2(a), lo2(acc)), __ffma2_rn(s2, hi2(a), hi2(acc)));
}
__device__ __forceinline__ float clampf(float g, float c) { return fmaxf(fminf(g, c), -c); }   // optimizer.cpp:17-22
__device__ __forceinline__ float4 clamp4(const float4 &g, float c) { return make_float4(clampf(g.x, c), clampf(g.y, c), clampf(g.z, c), clampf(g.w, c)); }

// sum over the VL lanes that hold one row (result in every lane of the group)
template <int VL> __device__ __forceinline__ float group_sum(float v, unsigned mask = FULL)
{
#pragma unroll
    for (int o = VL / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o);
    return v;
}
// sum over the 32/VL row groups of the warp
template <int VL> __device__ __forceinline__ float cross_sum(float v)
{
#pragma unroll
    for (int o = 16; o >= VL; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
template <int VL> __device__ __forceinline__ float cross_max(float v)
{
#pragma unroll
    for (int o = 16; o >= VL; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL, v, o));
    return v;
}
template <int VL> __device__ __forceinline__ float4 cross_sum4(float4 v)
{
    v.x = cross_sum<VL>(v.x); v.y = cross_sum<VL>(v.y); v.z = cross_sum<VL>(v.z); v.w = cross_sum<VL>(v.w);
    return v;
}

// Slot k of local sample s: Philox draw with the reference's "keep the previous value when the
// draw equals the positive" rule in stateless form (see philox.cuh).
__device__ __forceinline__ int sample_negative(const TrainParams &P, long long s, int k, int pos)
{
    long long t = s;
    for (;;) {
        const int d = (int)philox_draw(P.seed, P.epoch, P.sample_offset + (unsigned long long)t, (uint32_t)k,
                                       (uint32_t)P.num_items, P.tile_size, P.refresh_interval);
        if (d != pos) return d;
        if (t == 0) return 0;
        --t;
        pos = __ldg(&P.clicks[t]).y;
    }
}

// max(x, eps)^-1/2 (matrix_factorization.cpp:50-53 clamps the squared norms at eps before the root)
// as MUFU.RSQ + one Newton step (full fp32 accuracy without the division / square-root subroutines)
__device__ __forceinline__ float rsqrt_eps(float x)
{
    x = fmaxf(x, 1e-8f);
    const float y = rsqrtf(x);
    return y * fmaf(-0.5f * x * y, y, 1.5f);
}

// Two 11-bit hashes of an item id: the bit positions of the per-warp 2048-bit history filter (CACHE)
__device__ __forceinline__ unsigned hfilter_h1(int id) { return ((unsigned)id * 0x9E3779B1u) >> 21; }
__device__ __forceinline__ unsigned hfilter_h2(int id) { return ((unsigned)id * 0x85EBCA77u + 0x165667B1u) >> 21; }

// History ids of user u: lane holds ids lane, lane+32, lane+64, lane+96 (max_his <= 128).
__device__ __forceinline__ void load_history_ids(const TrainParams &P, int u, int lane, int (&hid)[4])
{
    const int *row = P.his + (size_t)u * P.max_his;
#pragma unroll
    for (int q = 0; q < 4; ++q) hid[q] = (lane + 32 * q < P.max_his) ? __ldg(row + lane + 32 * q) : 0;
}
__device__ __forceinline__ int pick4(const int (&hid)[4], int q)
{
    return q == 0 ? hid[0] : q == 1 ? hid[1] : q == 2 ? hid[2] : hid[3];
}

// Apply a warp's buffered W0 gradients: W0 -= lr0 * (sum_s mean_s^T (0.6 g_s)) / 32
// (behavior_aggregators.cpp:131-146).  ring slot s = [mean_s (D) | 0.6*g_s (D)].
template <int D>
__device__ __forceinline__ void w0_flush(const float *ring, float *W0s, float *w0g, int cnt, float lr0, int lane)
{
    constexpr int CPL = D / 32;   // columns per lane
    constexpr int IB = 16;        // rows per pass
    for (int ib = 0; ib < D; ib += IB) {
        float acc[IB][CPL];
#pragma unroll
        for (int i = 0; i < IB; ++i)
#pragma unroll
            for (int c = 0; c < CPL; ++c) acc[i][c] = 0.f;
        for (int s = 0; s < cnt; ++s) {
            const float *ms = ring + s * 2 * D;
            float gv[CPL];
#pragma unroll
            for (int c = 0; c < CPL; ++c) gv[c] = ms[D + lane * CPL + c];
#pragma unroll
            for (int i4 = 0; i4 < IB; i4 += 4) {
                const float4 m4 = *reinterpret_cast<const float4 *>(ms + ib + i4);
#pragma unroll
                for (int c = 0; c < CPL; ++c) {
                    acc[i4 + 0][c] = fmaf(m4.x, gv[c], acc[i4 + 0][c]);
                    acc[i4 + 1][c] = fmaf(m4.y, gv[c], acc[i4 + 1][c]);
                    acc[i4 + 2][c] = fmaf(m4.z, gv[c], acc[i4 + 2][c]);
                    acc[i4 + 3][c] = fmaf(m4.w, gv[c], acc[i4 + 3][c]);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < IB; ++i)
#pragma unroll
            for (int c = 0; c < CPL; ++c) {
                const int e = (ib + i) * D + lane * CPL + c;
                const float delta = -(lr0 * (acc[i][c] / 32.0f));
                const float old = atomicAdd(w0g + e, delta);   // returns the value before: picks up other warps' updates
                W0s[e] = old + delta;
            }
    }
}

// D: embedding dim (32/64/128).  NJ: negative rows held per row group (K <= NJ * 32/(D/4)).
// WARPS: warps per block.  PRELOAD_G: negative grad rows loaded up front with the weights.
template <int D, int NJ, int WARPS, bool PRELOAD_G, int UPD, bool CACHE>
__global__ void __launch_bounds__(WARPS * 32, 1) train_kernel(const TrainParams P)
{
    constexpr int VL = D / 4;        // lanes per row
    constexpr int R = 32 / VL;       // rows per warp-wide load
    constexpr int UNR = 8;           // history loads in flight per lane
    constexpr int IPG = D / R;       // W0 rows per group in the GEMV
    static_assert(NJ * R <= 32, "negative ids are exchanged by lane index");

    extern __shared__ __align__(16) float smem[];
    float *W0s = smem;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane / VL, sl = lane % VL;
    // W0 gradients: (mean, 0.6 g_u) pairs of the CTA's samples go into ONE ring of two 32-slot halves shared by its
    // warps; the warp that completes a half applies it (behavior_aggregators.cpp:141-146 with the CTA in the role of the
    // OpenMP thread).  Pending, unapplied gradients are then <= 64 per CTA whatever the number of warps, as in the
    // CTA-per-sample kernel -- with one ring per WARP a full grid holds ~28 000 of them, and the loss follows
    // (DESIGN.md section 6).
    float *ring = smem + D * D;                                   // [64][2D]
    float *wscr = ring + 64 * 2 * D + warp * D;                   // this warp's staging of the history mean (GEMV operand)
    unsigned *rctl = reinterpret_cast<unsigned *>(ring + 64 * 2 * D + WARPS * D);   // [0] slots claimed, [1] slots applied, [2..3] written per half
    unsigned *hfilter = rctl + 4 + warp * 64;                     // CACHE: 2048-bit filter
    if (threadIdx.x < 4) rctl[threadIdx.x] = 0u;

    for (int i = threadIdx.x * 4; i < D * D; i += WARPS * 32 * 4)
        *reinterpret_cast<float4 *>(W0s + i) = ldcg4(P.w0 + i);
    __syncthreads();

    const float gamma = 0.4f;                        // behavior_aggregators.cpp:37
    const float omg = 1.0f - gamma;
    const float score_mul = (float)(1.0 / 0.07);     // matrix_factorization.cpp:101
    const int K = P.K;
    double loss_acc = 0.0;
    // CACHE: state carried across the consecutive samples of one user (a "run").  The history column
    // sums are gathered once per run and then patched with the exact change of every history row this
    // warp itself rewrites; the user's weight and grad rows stay in registers.  In file order with one
    // warp this is the reference's arithmetic up to fp32 re-association of the column sums (the run is
    // re-gathered every REGATHER samples to bound that drift); under Hogwild, rows changed by OTHER
    // warps during the run are seen at the next gather instead of at the next sample.
    constexpr int REGATHER = 64;
    int cur_u = -1, H = 0, age = 0;
    int hid[4] = {0, 0, 0, 0};                       // the run's history ids: lane holds positions lane + 32q
    float4 acc = f4zero(), u_keep = f4zero(), gu_keep = f4zero();

    for (;;) {
        unsigned long long c = 0;
        if (lane == 0) c = atomicAdd(P.chunk_counter, 1ULL);
        c = __shfl_sync(FULL, c, 0);
        long long s0, s1;
        if (!chunk_range(P, (long long)c, s0, s1)) break;

        for (long long s = s0; s < s1; ++s) {
            __syncwarp();   // rows written by other lanes for the previous sample are visible
            const int2 up = __ldg(&P.clicks[s]);
            const int u = up.x, p = up.y;

            // ---- negatives (lane k draws slot k) and duplicate detection
            int negid = -1 - lane;
            if (lane < K) negid = sample_negative(P, s, lane, p);
            const unsigned same = __match_any_sync(FULL, negid);
            const bool dup = __any_sync(FULL, (same & (same - 1)) != 0);

            // ---- issue the row gathers
            const bool fresh = !CACHE || u != cur_u || age >= REGATHER;      // warp-uniform
            float4 uw, gu_old;
            if (fresh) {
                uw = ldcg4(P.user_w + (size_t)u * D + 4 * sl);
                gu_old = ldcg4(P.user_g + (size_t)u * D + 4 * sl);
            } else {
                uw = u_keep; gu_old = gu_keep;
            }
            float4 pw = ldcg4(P.item_w + (size_t)p * D + 4 * sl);
            const float4 gp_old = ldcg4(P.item_g + (size_t)p * D + 4 * sl);
            int nid[NJ];
            float4 nw[NJ];
            float4 gn[PRELOAD_G ? NJ : 1];
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const int k = j * R + g;
                nid[j] = __shfl_sync(FULL, negid, k & 31);
                const bool ok = k < K;
                nw[j] = ok ? ldcg4(P.item_w + (size_t)nid[j] * D + 4 * sl) : f4zero();
                if (PRELOAD_G) gn[j] = ok ? ldcg4(P.item_g + (size_t)nid[j] * D + 4 * sl) : f4zero();
            }

            // ---- history aggregation (behavior_aggregators.cpp:60-105)
            if (fresh) {
                H = __ldg(P.masks + u);
                load_history_ids(P, u, lane, hid);
                if (CACHE) {                                   // the run's history as a Bloom filter (two probes per id)
                    hfilter[lane] = 0u; hfilter[lane + 32] = 0u;
                    __syncwarp();
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        if (lane + 32 * q < H) {
                            const unsigned a = hfilter_h1(hid[q]), b = hfilter_h2(hid[q]);
                            atomicOr(&hfilter[a >> 5], 1u << (a & 31));
                            atomicOr(&hfilter[b >> 5], 1u << (b & 31));
                        }
                    }
                    __syncwarp();
                }
                acc = f4zero();
                for (int h0 = 0; h0 < H; h0 += R * UNR) {
                    const int hq = pick4(hid, h0 >> 5);
                    float4 t[UNR];
#pragma unroll
                    for (int i = 0; i < UNR; ++i) {
                        const int h = h0 + i * R + g;
                        const int id = __shfl_sync(FULL, hq, h & 31);
                        t[i] = (h < H) ? ldcg4(P.item_w + (size_t)id * D + 4 * sl) : f4zero();
                    }
#pragma unroll
                    for (int i = 0; i < UNR; ++i) acc = acc + t[i];
                }
                acc = cross_sum4<VL>(acc);
                cur_u = u; age = 0;
            }
            // CACHE: which of the item rows this sample rewrites are in the user's history list?
            // (p: with multiplicity, its change is patched into `acc`; a negative: rare, re-gather)
            int pmult = 0;
            bool regather = false;
            if (CACHE) {
                // the ids sit in registers since the run's gather: the positive's multiplicity is four ballots; the
                // negatives go through the filter first (two shared-memory probes per lane) and only a sample with a
                // filter hit (~7 % at H = 68) pays the exact comparison against the 128 register-resident ids
#pragma unroll
                for (int q = 0; q < 4; ++q) pmult += __popc(__ballot_sync(FULL, lane + 32 * q < H && hid[q] == p));
                bool maybe = false;
                if (lane < K) {
                    const unsigned a = hfilter_h1(negid), b = hfilter_h2(negid);
                    maybe = ((hfilter[a >> 5] >> (a & 31)) & (hfilter[b >> 5] >> (b & 31)) & 1u) != 0;
                }
                bool nhit = false;
                if (__any_sync(FULL, maybe)) {
                    for (int k = 0; k < K; ++k) {
                        const int nk = __shfl_sync(FULL, negid, k);
#pragma unroll
                        for (int q = 0; q < 4; ++q) nhit |= lane + 32 * q < H && hid[q] == nk;
                    }
                    nhit = __any_sync(FULL, nhit);
                }
                const bool pin = __any_sync(FULL, negid == p);
                regather = nhit || (pmult != 0 && (dup || pin));
            }
            const float rH = (float)(1.0 / (double)H);     // behavior_aggregators.cpp:63
            const float4 mean = rH * acc;
            float *ms = wscr;
            if (g == 0) *reinterpret_cast<float4 *>(ms + 4 * sl) = mean;
            __syncwarp();

            // ---- f = mean . W0 (behavior_aggregators.cpp:118), u <- gamma*u + (1-gamma)*f (:122)
            float4 f = f4zero();
#pragma unroll 4
            for (int i4 = 0; i4 < IPG; i4 += 4) {
                const float4 m4 = *reinterpret_cast<const float4 *>(ms + g * IPG + i4);
                const float *wr = W0s + (g * IPG + i4) * D + 4 * sl;
                fma4(f, m4.x, *reinterpret_cast<const float4 *>(wr));
                fma4(f, m4.y, *reinterpret_cast<const float4 *>(wr + D));
                fma4(f, m4.z, *reinterpret_cast<const float4 *>(wr + 2 * D));
                fma4(f, m4.w, *reinterpret_cast<const float4 *>(wr + 3 * D));
            }
            f = cross_sum4<VL>(f);
            float4 ua = make_float4(gamma * uw.x + omg * f.x, gamma * uw.y + omg * f.y,
                                    gamma * uw.z + omg * f.z, gamma * uw.w + omg * f.w);

            // ---- positive pair (matrix_factorization.cpp:45-63)
            const float uu = group_sum<VL>(dot4(ua, ua));
            const float pp = group_sum<VL>(dot4(pw, pw));
            const float upd = group_sum<VL>(dot4(ua, pw));
            //      1/|x| = max(x.x, eps)^-1/2 as MUFU.RSQ + one Newton step: the products below replace the reference's
            //      sqrt / divide sequence at fp32 accuracy (the single-step parity tests hold at 1e-5)
            const float ru = rsqrt_eps(uu), rp = rsqrt_eps(pp);
            const float r_u3_p = (ru * ru) * (ru * rp), r_u_p3 = (ru * rp) * (rp * rp);
            const float4 gu_p = r_u3_p * (uu * pw - upd * ua);
            const float4 gp_p = (-r_u_p3) * (pp * ua - upd * pw);
            const float cos_p = upd * (ru * rp);

            // ---- negatives: dots, scores, softmax loss (matrix_factorization.cpp:82-109)
            float un[NJ], nn[NJ], rn[NJ], sc[NJ];
            float m = -INFINITY;
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                un[j] = group_sum<VL>(dot4(ua, nw[j]));
                nn[j] = group_sum<VL>(dot4(nw[j], nw[j]));
                rn[j] = rsqrt_eps(nn[j]);
                sc[j] = (j * R + g < K) ? (un[j] * (ru * rn[j]) - cos_p) * score_mul : -INFINITY;
                m = fmaxf(m, sc[j]);
            }
            m = cross_max<VL>(m);
            float esum = 0.f;
#pragma unroll
            for (int j = 0; j < NJ; ++j) { sc[j] = expf(sc[j] - m); esum += sc[j]; }
            esum = cross_sum<VL>(esum);
            esum += expf(-m);                                  // the implicit zero logit (:107)
            const float loss = m + logf(esum);                 // :108
            const float inv_esum = 1.f / esum;

            // ---- backward (matrix_factorization.cpp:118-150)
            float4 gacc = f4zero();
            float lgsum = 0.f;
            if (UPD != UPD_STORE || !dup) {
                float4 dl[UPD == UPD_FRESH ? NJ : 1];
#pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    if (UPD == UPD_FRESH) dl[j] = f4zero();
                    if (j * R + g < K) {
                        const float lg = (sc[j] * inv_esum) * score_mul;            // :109
                        const float r_u3_n = (ru * ru) * (ru * rn[j]), r_u_n3 = (ru * rn[j]) * (rn[j] * rn[j]);
                        const float4 gu_n = r_u3_n * (uu * nw[j] - un[j] * ua);     // :138
                        const float4 gn_n = r_u_n3 * (nn[j] * ua - un[j] * nw[j]);  // :139
                        fma4(gacc, lg, gu_n - gu_p);                                 // :141
                        lgsum += lg;                                                 // :142
                        float *wrow = P.item_w + (size_t)nid[j] * D + 4 * sl;
                        float *grow = P.item_g + (size_t)nid[j] * D + 4 * sl;
                        if (UPD == UPD_FRESH) {
                            dl[j] = lg * gn_n;
                        } else {
                            const float4 gold = PRELOAD_G ? gn[j] : ldcg4(grow);
                            float4 gk = gold;
                            fma4(gk, lg, gn_n);                                      // :143
                            gk = clamp4(gk, P.clip);                                 // sgd.cpp:21-25
                            if (UPD == UPD_ATOMIC) {
                                red4(wrow, (-P.lr) * gk);
                                red4(grow, gk - gold);
                            } else {
                                fma4(nw[j], -P.lr, gk);
                                stcg4(wrow, nw[j]);                                  // :148
                                stcg4(grow, gk);                                     // :149
                            }
                        }
                    }
                }
                if (UPD == UPD_FRESH) {
                    // the grad rows as they are NOW, right before the step: all loads first, then the stores; the weight
                    // rows take their step as a reduction at L2 and are not read again
                    float4 fg[NJ];
#pragma unroll
                    for (int j = 0; j < NJ; ++j)
                        if (j * R + g < K) fg[j] = ldcg4(P.item_g + (size_t)nid[j] * D + 4 * sl);
#pragma unroll
                    for (int j = 0; j < NJ; ++j) {
                        if (j * R + g < K) {
                            float4 gk = fg[j] + dl[j];
                            gk = clamp4(gk, P.clip);
                            red4(P.item_w + (size_t)nid[j] * D + 4 * sl, (-P.lr) * gk);
                            stcg4(P.item_g + (size_t)nid[j] * D + 4 * sl, gk);
                        }
                    }
                }
            } else {
                // Duplicate negative ids in this sample: replay the reference's sequential k loop
                // (grad row re-read per k, last writer wins on the weight row).
#pragma unroll
                for (int j = 0; j < NJ; ++j) {
#pragma unroll
                    for (int gg = 0; gg < R; ++gg) {
                        if (gg == g && j * R + g < K) {
                            const float lg = (sc[j] * inv_esum) * score_mul;
                            const float r_u3_n = (ru * ru) * (ru * rn[j]), r_u_n3 = (ru * rn[j]) * (rn[j] * rn[j]);
                            const float4 gu_n = r_u3_n * (uu * nw[j] - un[j] * ua);
                            const float4 gn_n = r_u_n3 * (nn[j] * ua - un[j] * nw[j]);
                            fma4(gacc, lg, gu_n - gu_p);
                            lgsum += lg;
                            float *grow = P.item_g + (size_t)nid[j] * D + 4 * sl;
                            float4 gk = ldcg4(grow);
                            fma4(gk, lg, gn_n);
                            gk = clamp4(gk, P.clip);
                            float4 nk = nw[j];
                            fma4(nk, -P.lr, gk);
                            stcg4(P.item_w + (size_t)nid[j] * D + 4 * sl, nk);
                            stcg4(grow, gk);
                        }
                        __syncwarp();
                    }
                }
            }
            gacc = cross_sum4<VL>(gacc);
            lgsum = cross_sum<VL>(lgsum);
            float4 gu = gu_old + gacc;
            float4 gp = gp_old;
            fma4(gp, lgsum, gp_p);

            // ---- aggregator backward (behavior_aggregators.cpp:129-153): this sample's (mean, 0.6 g_u) pair goes into the
            //      CTA's ring; the warp that completes a half of 32 applies it (forward's iteration += 1, :124)
            {
                unsigned idx = 0;
                if (lane == 0) {
                    idx = atomicAdd(&rctl[0], 1u);
                    while (idx - *reinterpret_cast<volatile unsigned *>(&rctl[1]) >= 64u) { }   // the half is still being applied (rare)
                }
                idx = __shfl_sync(FULL, idx, 0);
                float *slot_ = ring + (idx & 63u) * 2 * D;
                if (g == 0) {
                    *reinterpret_cast<float4 *>(slot_ + 4 * sl) = mean;
                    *reinterpret_cast<float4 *>(slot_ + D + 4 * sl) = omg * gu;
                }
                __syncwarp();
                const unsigned half = (idx >> 5) & 1u;
                unsigned done = 0;
                if (lane == 0) { __threadfence_block(); done = atomicAdd(&rctl[2 + half], 1u) + 1u; }
                done = __shfl_sync(FULL, done, 0);
                if (done == 32u) {                             // exactly one warp per completed half
                    __threadfence_block();
                    w0_flush<D>(ring + half * 32 * 2 * D, W0s, P.w0, 32, P.lr_w0, lane);
                    __syncwarp();
                    if (lane == 0) { rctl[2 + half] = 0u; __threadfence_block(); atomicAdd(&rctl[1], 32u); }
                    __syncwarp();
                }
            }
            gu = gamma * gu;

            // ---- user / positive step and write-back (matrix_factorization.cpp:166-174)
            gu = clamp4(gu, P.clip);
            fma4(ua, -P.lr, gu);
            float *prow = P.item_w + (size_t)p * D + 4 * sl, *pgrow = P.item_g + (size_t)p * D + 4 * sl;
            if (UPD == UPD_STORE) {
                gp = clamp4(gp, P.clip);
                fma4(pw, -P.lr, gp);
                if (R == 1 || g == 0) stcg4(prow, pw);
                if (R == 1 || g == 1) stcg4(pgrow, gp);
            } else if (UPD == UPD_ATOMIC) {
                gp = clamp4(gp, P.clip);
                if (R == 1 || g == 0) red4(prow, (-P.lr) * gp);
                if (R == 1 || g == 1) red4(pgrow, gp - gp_old);
            } else {
                // UPD_FRESH: every row group reads the grad row as it is now (the run cache needs the applied step on all lanes)
                gp = ldcg4(pgrow);
                fma4(gp, lgsum, gp_p);
                gp = clamp4(gp, P.clip);
                if (R == 1 || g == 0) red4(prow, (-P.lr) * gp);
                if (R == 1 || g == 1) stcg4(pgrow, gp);
            }
            if (R == 1 || g == 0) stcg4(P.user_w + (size_t)u * D + 4 * sl, ua);
            if (R == 1 || g == 1) stcg4(P.user_g + (size_t)u * D + 4 * sl, gu);
            if (CACHE) {
                u_keep = ua; gu_keep = gu;
                age = regather ? REGATHER : age + 1;
                if (pmult != 0 && !regather) fma4(acc, -P.lr * (float)pmult, gp);     // I[p] moved by -lr*gp
            }
            if (P.loss_out && lane == 0) P.loss_out[s - P.begin] = loss;
            loss_acc += (double)loss;
        }
    }
    // ---- the leftover (< 32) pairs of the half that was being filled (the reference drops a partial batch: drop_w0_tail)
    __syncthreads();
    if (warp == 0 && !P.drop_w0_tail) {
        const unsigned claimed = rctl[0], applied = rctl[1];
        if (claimed > applied) w0_flush<D>(ring + ((applied >> 5) & 1u) * 32 * 2 * D, W0s, P.w0, (int)(claimed - applied), P.lr_w0, lane);
    }
    if (lane == 0 && loss_acc != 0.0) atomicAdd(P.loss_sum, loss_acc);
}

}  // namespace heat
