// heat_b200/csrc/train_kernel_bigk.cuh -- training kernel for the large-negative regime (SimpleX's
// num_negs in the hundreds to thousands, BASELINE.json configs[3]: emb_dim 128, num_negs 1000).
//
// One CTA of 512 threads owns one sample; a sample moves (2+K) weight rows and (2+K) persistent grad
// rows in and out (2.1 MB at D=128, K=1000), so the kernel is a row gather/scatter engine and the
// arithmetic has to stay out of its way.  Same per-sample arithmetic as train_kernel.cuh /
// train_kernel_cta.cuh (reference: models/matrix_factorization.cpp:15-181,
// behavior_aggregators/behavior_aggregators.cpp:51-153, optimizers/sgd.cpp:14-26); what differs is where things live:
//
//   * a row is D/4 lanes x float4; the CTA is G = 512/(D/4) row groups (16 at D=128), group g owns the
//     negatives k = g, g+G, ...; UNR rows per group are in flight per loop trip (LDG.128, ld.global.cg).
//   * the K negatives do not fit in registers, so the sample runs in two passes around the softmax.
//     Pass 1 reads every weight row once and leaves (u.n_k, n_k.n_k) per k in shared memory; the first
//     `cache_rows` rows (as many as fit in the 227 KB: ~300 of 1000 at D=128) are fetched with cp.async straight
//     into a shared-memory cache while the registers stream the others, and stay there for pass 2.  The 2*UNR dot
//     products of a trip are reduced with a TRANSPOSING butterfly (2*UNR-1 shuffles instead of 2*UNR*log2(lanes)):
//     afterwards each lane holds one total, so the per-row scalar math (rsqrt, exp) runs once per trip, one row per
//     lane.  Between the passes one thread per k turns (u.n, n.n) into the two step coefficients; pass 2 is then
//     loads, three FMAs and a clamp per slice, and stores; it re-reads only the weight rows that did not fit.
//   * W0 (D x D): at D <= 64 a register slice per thread (column = tid / (512/D)) plus the thread's slice of the W0
//     gradient accumulator (behavior_aggregators.cpp:131-146, applied every 32 samples with atomicAdd; the returned
//     value refreshes the register copy).  At D = 128 those 64 registers are needed for rows in flight: mean.W0
//     reads W0 through L2 (64 KB per 2.1 MB sample) and the (mean, 0.6 g_u) pairs of the last 32 samples sit in
//     shared-memory rings for one rank-32 update.
//   * duplicate negatives (certain at K=1000: K^2/2I expected pairs) follow the reference's sequential k loop
//     (matrix_factorization.cpp:127-150): all copies see the same snapshot and the same lg, the grad row takes the
//     clipped add once per copy, the last copy's weight row wins.  A shared-memory hash set gives every id one
//     owner slot and its multiplicity m; the owner applies g = clamp(g + dg) m times and stores once.  A negative
//     equal to the positive is overwritten by the positive's write-back (:169-174), so it is skipped.
//   * the history is gathered for every sample (with K=1000 a negative hits the history of most samples, so
//     the run cache of the small-K kernels would be invalidated anyway).
//
// With one CTA the samples are applied strictly in file order (deterministic mode, used by the parity tests).
#pragma once
#include "train_kernel_cta.cuh"

namespace heat {

template <int D> struct BigKCfg {
    static_assert(D == 32 || D == 64 || D == 128, "emb_dim 32, 64 or 128");
    static constexpr int NT = 512;
    static constexpr int LPR = D / 4;          // lanes per row
    static constexpr int G = NT / LPR;         // row groups per CTA
    static constexpr int RQ = NT / D;          // threads per W0 column
    static constexpr int IQ = D / RQ;          // W0 rows per thread
    static constexpr int UNR = D == 32 ? 4 : 8;    // rows in flight per group, pass 1 (2*UNR <= LPR)
    static constexpr int UNR2 = D == 128 ? 8 : 4;  // pass 2 (a weight and a grad slice per row)
    static constexpr int UH = 8;               // history rows in flight per group
    static constexpr bool RING = D == 128;     // W0 gradient from shared-memory rings, W0 read through L2
};

// Shared-memory layout (bytes), a function of (D, K) evaluated identically on host and device.
struct BigKLayout {
    int Kp, HS, hshift;
    int off_red, off_sE, off_sB, off_un, off_nn, off_neg, off_hkey, off_hcnt, off_mult, off_misc, off_ring, off_cache;
};
HEAT_HD BigKLayout bigk_layout(int D, int G, int K, bool ring)
{
    BigKLayout L;
    L.Kp = (K + 3) & ~3;
    int hs = 64, sh = 26;
    while (hs < 2 * K) { hs <<= 1; --sh; }
    L.HS = hs; L.hshift = sh;
    int o = 3 * D * 4;                        // mean, f, 0.6*g_u (the latter two unused slots when the rings hold them)
    L.off_red = o; o += G * D * 4;
    L.off_sE = o; o += G * 4;
    L.off_sB = o; o += G * 4;
    L.off_un = o; o += L.Kp * 4;              // u.n_k; before pass 1: hash slot | owner bit
    L.off_nn = o; o += L.Kp * 4;
    L.off_neg = o; o += L.Kp * 4;
    L.off_hkey = o; o += hs * 4;              // hash keys; after the multiplicities are known: int2 {id, m} per k
    L.off_hcnt = o; o += hs * 4;              // hash counts;                                   float2 {c1, c2} per k
    L.off_mult = o; o += (L.Kp + 15) & ~15;
    L.off_misc = o; o += 16;
    L.off_ring = o; o += ring ? 2 * 32 * D * 4 : 0;
    L.off_cache = o;
    return L;
}

// Totals of N per-lane partial values over the LPR lanes of a row group (N a power of two, N <= LPR): log2(N)
// halving exchanges (N-1 shuffles in all), then plain butterflies over what is left.  On return lane sl holds the
// total of value number sl / (LPR/N).
template <int LPR, int N> __device__ __forceinline__ float group_sum_transposed(float (&v)[N], int sl)
{
    static_assert(N <= LPR, "one value per lane at least");
    int o = LPR / 2;
#pragma unroll
    for (int n = N / 2; n >= 1; n >>= 1, o >>= 1) {
        const bool up = (sl & o) != 0;
#pragma unroll
        for (int i = 0; i < n; ++i) {
            const float send = up ? v[i] : v[i + n];
            const float keep = up ? v[i + n] : v[i];
            v[i] = keep + __shfl_xor_sync(FULL, send, o);
        }
    }
    float r = v[0];
#pragma unroll
    for (int q = LPR / N / 2; q >= 1; q >>= 1) r += __shfl_xor_sync(FULL, r, q);
    return r;
}

template <int D>
__global__ void __launch_bounds__(BigKCfg<D>::NT, 1) train_bigk_kernel(const TrainParams P)
{
    using C = BigKCfg<D>;
    constexpr int NT = C::NT, LPR = C::LPR, G = C::G, RQ = C::RQ, IQ = C::IQ, UNR = C::UNR, UNR2 = C::UNR2, UH = C::UH;
    constexpr bool RING = C::RING;
    constexpr int NV = 2 * UNR, PAIR = LPR / NV;             // values per trip; lanes that end up holding one total
    extern __shared__ __align__(16) unsigned char smraw[];
    const int K = P.K;
    const BigKLayout L = bigk_layout(D, G, K, RING);
    float *mean0 = reinterpret_cast<float *>(smraw), *f_s = mean0 + D, *g060 = f_s + D;
    float *red = reinterpret_cast<float *>(smraw + L.off_red);
    float *sE = reinterpret_cast<float *>(smraw + L.off_sE), *sB = reinterpret_cast<float *>(smraw + L.off_sB);
    float *un_s = reinterpret_cast<float *>(smraw + L.off_un), *nn_s = reinterpret_cast<float *>(smraw + L.off_nn);
    int *tmp_s = reinterpret_cast<int *>(smraw + L.off_un);
    int *negid = reinterpret_cast<int *>(smraw + L.off_neg);
    int *hkey = reinterpret_cast<int *>(smraw + L.off_hkey), *hcnt = reinterpret_cast<int *>(smraw + L.off_hcnt);
    int2 *recA = reinterpret_cast<int2 *>(smraw + L.off_hkey);
    float2 *recB = reinterpret_cast<float2 *>(smraw + L.off_hcnt);
    unsigned char *mult = smraw + L.off_mult;
    unsigned long long *sh_chunk = reinterpret_cast<unsigned long long *>(smraw + L.off_misc);
    float *mring = reinterpret_cast<float *>(smraw + L.off_ring), *gring = mring + 32 * D;
    float *cache = reinterpret_cast<float *>(smraw + L.off_cache);
    const int CR = P.cache_rows, HS = L.HS;

    const int tid = threadIdx.x;
    const int g = tid / LPR, sl = tid % LPR;                 // row group, lane within the row
    const int col = tid / RQ, rb = tid % RQ;                 // W0 column, row block
    const float gamma = 0.4f, omg = 1.0f - gamma;
    const float score_mul = (float)(1.0 / 0.07);
    constexpr int TILE = G * UNR;
    const int ntile = (K + TILE - 1) / TILE;

    constexpr int IQR = RING ? 1 : IQ;
    float wreg[IQR], wacc[IQR];
    if (!RING) {
#pragma unroll
        for (int i = 0; i < IQR; ++i) { wreg[i] = __ldcg(P.w0 + (rb * IQ + i) * D + col); wacc[i] = 0.f; }
    }
    // W0 -= lr0 * acc / 32 (behavior_aggregators.cpp:141-146) over the last `ns` samples
    auto w0_step = [&](int ns) {
        if (RING) {
            float acc[IQ];
#pragma unroll
            for (int i = 0; i < IQ; ++i) acc[i] = 0.f;
            for (int q = 0; q < ns; ++q) {
                const float *pm = mring + q * D + rb * IQ;
                const float gq = gring[q * D + col];
#pragma unroll
                for (int i4 = 0; i4 < IQ; i4 += 4) {
                    const float4 m4 = *reinterpret_cast<const float4 *>(pm + i4);
                    acc[i4 + 0] = fmaf(m4.x, gq, acc[i4 + 0]); acc[i4 + 1] = fmaf(m4.y, gq, acc[i4 + 1]);
                    acc[i4 + 2] = fmaf(m4.z, gq, acc[i4 + 2]); acc[i4 + 3] = fmaf(m4.w, gq, acc[i4 + 3]);
                }
            }
#pragma unroll
            for (int i = 0; i < IQ; ++i) atomicAdd(P.w0 + (rb * IQ + i) * D + col, -(P.lr_w0 * (acc[i] / 32.0f)));
        } else {
#pragma unroll
            for (int i = 0; i < IQR; ++i) {
                const float delta = -(P.lr_w0 * (wacc[i] / 32.0f));
                const float old = atomicAdd(P.w0 + (rb * IQ + i) * D + col, delta);   // the value before: picks up other CTAs' steps
                wreg[i] = old + delta;
                wacc[i] = 0.f;
            }
        }
    };

    unsigned it = 0;
    double loss_acc = 0.0;

    for (;;) {
        __syncthreads();
        if (tid == 0) *sh_chunk = atomicAdd(P.chunk_counter, 1ULL);
        __syncthreads();
        const long long ch = (long long)*sh_chunk;
        long long s0, s1;
        if (!chunk_range(P, ch, s0, s1)) break;
        for (long long s = s0; s < s1; ++s) {
            const int2 click = __ldg(&P.clicks[s]);
            const int u = click.x, p = click.y;
            const int H = __ldg(P.masks + u);
            const float rH = (float)(1.0 / (double)H);      // behavior_aggregators.cpp:63
            float *mean = RING ? mring + (it & 31) * D : mean0;
            float *g06 = RING ? gring + (it & 31) * D : g060;

            // ---- A: empty the hash set
            for (int i = tid; i < HS; i += NT) { hkey[i] = -1; hcnt[i] = 0; }
            __syncthreads();                                                              // #1

            // ---- B: negatives (one Philox block = four slots, philox.cuh) into the hash set; history gather
            {
                const unsigned long long gs = P.sample_offset + (unsigned long long)s;
                for (int q = tid; q < L.Kp / 4; q += NT) {
                    const Philox4 r = philox4x32_10((uint32_t)gs, (uint32_t)(gs >> 32), P.epoch, (uint32_t)q, (uint32_t)P.seed,
                                                    (uint32_t)(P.seed >> 32));
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const int k = 4 * q + t;
                        if (k < K) {
                            int d = (int)map_draw(r.v[t], P.seed, P.epoch, gs, (uint32_t)P.num_items, P.tile_size, P.refresh_interval);
                            if (d == p) d = sample_negative(P, s, k, p);
                            negid[k] = d;
                            unsigned h = ((unsigned)d * 2654435761u) >> L.hshift;
                            for (;;) {
                                const int old = atomicCAS(&hkey[h], -1, d);
                                if (old == -1 || old == d) break;
                                h = (h + 1) & (unsigned)(HS - 1);
                            }
                            const int c = atomicAdd(&hcnt[h], 1);
                            tmp_s[k] = (int)h | (c == 0 ? (int)0x80000000 : 0);          // the first arrival owns the id
                        }
                    }
                }
                // history rows of the user (behavior_aggregators.cpp:96-105): partial column sums per group
                const int *hrow = P.his + (size_t)u * P.max_his;
                float4 a4 = f4zero();
                for (int h0 = g; h0 < H; h0 += G * UH) {
                    int hid[UH];
                    float4 hr[UH];
#pragma unroll
                    for (int q = 0; q < UH; ++q) hid[q] = (h0 + q * G < H) ? __ldg(hrow + h0 + q * G) : -1;
#pragma unroll
                    for (int q = 0; q < UH; ++q) hr[q] = hid[q] >= 0 ? ldcg4(P.item_w + (size_t)hid[q] * D + 4 * sl) : f4zero();
#pragma unroll
                    for (int q = 0; q < UH; ++q) a4 = a4 + hr[q];
                }
                *reinterpret_cast<float4 *>(red + g * D + 4 * sl) = a4;
            }
            __syncthreads();                                                              // #2

            // ---- C: the cached rows start moving (every lane copies the slice it reads later: completion needs no
            //      barrier); multiplicities; history mean
            for (int k = g; k < CR; k += G) cp_async16(cache + (size_t)k * D + 4 * sl, P.item_w + (size_t)negid[k] * D + 4 * sl);
            cp_async_commit();
            if (tid < D) {
                float a = 0.f;
#pragma unroll 8
                for (int q = 0; q < G; ++q) a += red[q * D + tid];
                mean[tid] = a * rH;
            }
            for (int k = tid; k < K; k += NT) {
                const int v = tmp_s[k];
                int m = v < 0 ? hcnt[v & 0x7fffffff] : 0;
                if (negid[k] == p) m = 0;                     // overwritten by the positive's write-back (:169-174)
                mult[k] = (unsigned char)min(m, 255);
            }
            __syncthreads();                                                              // #3

            // ---- D: f = mean . W0 (behavior_aggregators.cpp:118)
            {
                const float *pm = mean + rb * IQ;
                float f = 0.f;
                if (RING) {
                    float wv[IQ];
#pragma unroll
                    for (int i = 0; i < IQ; ++i) wv[i] = __ldcg(P.w0 + (rb * IQ + i) * D + col);
#pragma unroll
                    for (int i = 0; i < IQ; ++i) f = fmaf(pm[i], wv[i], f);
                } else {
#pragma unroll
                    for (int i = 0; i < IQR; ++i) f = fmaf(pm[i], wreg[i], f);
                }
#pragma unroll
                for (int o = 1; o < RQ; o <<= 1) f += __shfl_xor_sync(FULL, f, o);
                if (rb == 0) f_s[col] = f;
            }
            __syncthreads();                                                              // #4

            // ---- E: u <- gamma*u + (1-gamma)*f (:122); positive dots; pass 1 over the negatives
            float4 ua = gamma * ldcg4(P.user_w + (size_t)u * D + 4 * sl);
            fma4(ua, omg, *reinterpret_cast<const float4 *>(f_s + 4 * sl));
            const float4 pw = ldcg4(P.item_w + (size_t)p * D + 4 * sl);
            float uu, pp, upd;
            {
                float r[3] = {dot4(ua, ua), dot4(pw, pw), dot4(ua, pw)};
                group_sum_n<LPR, 3>(r);
                uu = r[0]; pp = r[1]; upd = r[2];
            }
            const float rinv_u = rsqrt_eps(uu), rinv_p = rsqrt_eps(pp);
            const float Ep = expf(-((upd * (rinv_u * rinv_p)) * score_mul));
            {
                float es = 0.f, bs = 0.f;
                float4 gpart = f4zero();
                const int vi = sl / PAIR;                    // which total this lane holds after the transposing butterfly
                const int jrow = vi >> 1;                    // ... i.e. row jrow of the trip (even: u.n, odd: n.n)
                const bool primary = (sl % (2 * PAIR)) == 0; // one lane per row accumulates / publishes
                const int gbase = (tid & 31) & ~(LPR - 1);   // first lane of this row group within the warp
                bool waited = false;
                for (int t = ntile - 1; t >= 0; --t) {       // uncached rows first: the cp.async rows land meanwhile
                    const int k0 = t * TILE + g;
                    if (!waited && t * TILE < CR) { cp_async_wait0(); waited = true; }
                    float4 w[UNR];
#pragma unroll
                    for (int j = 0; j < UNR; ++j) {
                        const int k = k0 + j * G;
                        w[j] = k >= K ? f4zero()
                             : k < CR ? *reinterpret_cast<const float4 *>(cache + (size_t)k * D + 4 * sl)
                                      : ldcg4(P.item_w + (size_t)negid[k] * D + 4 * sl);
                    }
                    float r[NV];
#pragma unroll
                    for (int j = 0; j < UNR; ++j) { r[2 * j] = dot4(ua, w[j]); r[2 * j + 1] = dot4(w[j], w[j]); }
                    const float mine = group_sum_transposed<LPR, NV>(r, sl);
                    const float other = __shfl_xor_sync(FULL, mine, PAIR);
                    const float un = (vi & 1) ? other : mine, nn = (vi & 1) ? mine : other;
                    const int k = k0 + jrow * G;
                    const float rinv_n = rsqrt_eps(nn);
                    const float E = k < K ? expf((un * (rinv_u * rinv_n)) * score_mul) : 0.f;
                    const float r3 = (rinv_u * rinv_u) * (rinv_u * rinv_n);             // 1 / (|u|^3 |n|)   (:136)
                    const float a = (E * r3) * uu;
                    if (primary && k < K) {
                        es += E;
                        bs = fmaf(E, r3 * un, bs);
                        un_s[k] = un; nn_s[k] = nn;
                    }
#pragma unroll
                    for (int j = 0; j < UNR; ++j) fma4(gpart, __shfl_sync(FULL, a, gbase + j * 2 * PAIR), w[j]);
                }
                if (!waited) cp_async_wait0();
                es = group_sum<LPR>(es); bs = group_sum<LPR>(bs);
                *reinterpret_cast<float4 *>(red + g * D + 4 * sl) = gpart;
                if (sl == 0) { sE[g] = es; sB[g] = bs; }
            }
            __syncthreads();                                                              // #5

            // ---- F: softmax with the implicit zero logit (matrix_factorization.cpp:55-61, 91-109), as in the CTA kernel:
            //      esum = 1 + Ep * sum_k E_k, lg_k = cE * E_k;  one thread per k: lg*gn_n = c1*u + c2*N[k] (:139-143)
            float SE = 0.f;
#pragma unroll 8
            for (int q = 0; q < G; ++q) SE += sE[q];
            const float esum = fmaf(Ep, SE, 1.0f);
            const float cE = (Ep * score_mul) / esum;
            for (int k = tid; k < K; k += NT) {
                const float un = un_s[k], nn = nn_s[k];
                const float rinv_n = rsqrt_eps(nn);
                const float E = expf((un * (rinv_u * rinv_n)) * score_mul);
                const float lg = cE * E;                                                  // :109
                const float rn3 = (rinv_u * rinv_n) * (rinv_n * rinv_n);                  // 1 / (|u| |n|^3)   (:137)
                recA[k] = make_int2(negid[k], (int)mult[k]);
                recB[k] = make_float2((lg * rn3) * nn, -(lg * rn3) * un);
            }
            __syncthreads();                                                              // #5b
            if (g == 0) {
                // user row: grad (:60-62, :141), aggregator backward, clipped step (:152, :166, :171-172)
                float4 gs4 = f4zero();
                float SB = 0.f;
#pragma unroll 8
                for (int q = 0; q < G; ++q) { gs4 = gs4 + *reinterpret_cast<const float4 *>(red + q * D + 4 * sl); SB += sB[q]; }
                const float q3 = (cE * SE) * ((rinv_u * rinv_u) * (rinv_u * rinv_p));    // lgsum / (|u|^3 |p|)
                float4 gu = ldcg4(P.user_g + (size_t)u * D + 4 * sl);
                fma4(gu, cE, gs4);
                fma4(gu, fmaf(q3, upd, -(cE * SB)), ua);
                fma4(gu, -q3 * uu, pw);
                *reinterpret_cast<float4 *>(g06 + 4 * sl) = omg * gu;                     // behavior_aggregators.cpp:131-139
                gu = gamma * gu;                                                          // :148-152
                gu = clamp4(gu, P.clip);
                float4 un4 = ua;
                fma4(un4, -P.lr, gu);
                stcg4(P.user_w + (size_t)u * D + 4 * sl, un4);
                stcg4(P.user_g + (size_t)u * D + 4 * sl, gu);
            } else if (g == 1) {
                // positive row (:142, :169, :173-174)
                const float q = (cE * SE) * ((rinv_u * rinv_p) * (rinv_p * rinv_p));     // lgsum / (|u| |p|^3)
                float4 gp = ldcg4(P.item_g + (size_t)p * D + 4 * sl);
                fma4(gp, -q * pp, ua); fma4(gp, q * upd, pw);
                gp = clamp4(gp, P.clip);
                float4 wp = pw;
                fma4(wp, -P.lr, gp);
                stcg4(P.item_w + (size_t)p * D + 4 * sl, wp);
                stcg4(P.item_g + (size_t)p * D + 4 * sl, gp);
            }
            // ---- pass 2: clipped step and write-back of the negative rows (:127-150; sgd.cpp:14-26)
            for (int k0 = g; k0 < K; k0 += G * UNR2) {
                int2 ra[UNR2];
                float4 w[UNR2], gr[UNR2];
#pragma unroll
                for (int j = 0; j < UNR2; ++j) {
                    const int k = k0 + j * G;
                    ra[j] = k < K ? recA[k] : make_int2(0, 0);
                }
#pragma unroll
                for (int j = 0; j < UNR2; ++j) {
                    const int k = k0 + j * G;
                    if (ra[j].y) {
                        gr[j] = ldcg4(P.item_g + (size_t)ra[j].x * D + 4 * sl);
                        w[j] = k < CR ? *reinterpret_cast<const float4 *>(cache + (size_t)k * D + 4 * sl)
                                      : ldcg4(P.item_w + (size_t)ra[j].x * D + 4 * sl);
                    }
                }
#pragma unroll
                for (int j = 0; j < UNR2; ++j) {
                    if (ra[j].y) {
                        const float2 c = recB[k0 + j * G];
                        float4 gk = gr[j];
                        for (int rep = 0; rep < ra[j].y; ++rep) {
                            fma4(gk, c.x, ua); fma4(gk, c.y, w[j]);
                            gk = clamp4(gk, P.clip);
                        }
                        float4 wk = w[j];
                        fma4(wk, -P.lr, gk);
                        stcg4(P.item_w + (size_t)ra[j].x * D + 4 * sl, wk);               // :148
                        stcg4(P.item_g + (size_t)ra[j].x * D + 4 * sl, gk);               // :149
                    }
                }
            }
            __syncthreads();                                                              // #6
            // ---- W0 gradient: acc += mean^T (0.6 g_u) (behavior_aggregators.cpp:131-139), stepped every 32 samples
            if (!RING) {
                const float gq = g06[col];
                const float *pm = mean + rb * IQ;
#pragma unroll
                for (int i = 0; i < IQR; ++i) wacc[i] = fmaf(pm[i], gq, wacc[i]);
            }
            if (tid == 0) {
                const float loss = logf(esum);                                            // :108
                loss_acc += (double)loss;
                if (P.loss_out) P.loss_out[s - P.begin] = loss;
            }
            it += 1;
            if ((it & 31) == 0) w0_step(32);
        }
    }
    // leftover (< 32) W0 gradients: the reference drops them with the per-thread aggregator (drop_w0_tail)
    if ((it & 31) != 0 && !P.drop_w0_tail) w0_step((int)(it & 31));
    if (tid == 0 && loss_acc != 0.0) atomicAdd(P.loss_sum, loss_acc);
}

}  // namespace heat
