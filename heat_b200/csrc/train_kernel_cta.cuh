// heat_b200/csrc/train_kernel_cta.cuh -- latency-oriented training kernel: one CTA per sample.
//
// Why: the reference's update is a clipped RUNNING SUM applied at every touch, so its dynamics depend
// on how many samples are in flight: the loss / Recall of the reference at 1-128
// threads is reproduced only while a few hundred samples at most are being processed concurrently.
// Throughput at a FIXED number of samples in flight is (samples in flight) / (latency of one sample),
// so this kernel spends a whole CTA on one sample to make that latency small (~1 us instead of the
// ~5 us a single warp needs for the same arithmetic):
//
//   slots 0..15   (D/4 lanes each)  own the negative rows k = slot, slot+16, ...   (NPS per slot)
//   slot 16                          owns the positive row
//   slot 17                          owns the user row (kept in registers across a user's run)
//   threads 0..4D-1                  also compute mean.W0 (4 partial GEMVs) and own the D*D/(4D)
//                                    = D/4 elements each of the W0 gradient accumulator (registers)
//   control warp                     prepares the NEXT sample while the rows of this one are in
//                                    flight: click, Philox negatives, duplicate detection, history
//                                    membership; patches the cached history column sums
//
// Per sample: issue every row load, overlap the L2 round trip with mean.W0 and the user aggregation,
// then dots -> [barrier] -> softmax (recomputed per slot from 2K+3 shared scalars) -> row updates ->
// [barrier]; the user step of a sample overlaps the next sample's loads.  Three block barriers, one L2
// round trip on the critical path.
// Same arithmetic as train_kernel.cuh (reference: models/matrix_factorization.cpp:15-181,
// behavior_aggregators/behavior_aggregators.cpp:51-153, optimizers/sgd.cpp:14-26); with one CTA the
// samples are applied strictly in file order (deterministic mode).
#pragma once
#include "train_kernel.cuh"

namespace heat {

struct CtaMeta {            // one sample, prepared by the control warp
    int u, p, fresh, dup, H, done, pmult, regather;
    long long s;
    int negid[64];
};

template <int D, int NPS> struct CtaCfg {
    static constexpr int LPR = D / 4;                       // lanes per row
    static constexpr int NSLOT = 18;
    static constexpr int ROWT = NSLOT * LPR;
    static constexpr int CTRL0 = (ROWT + 31) / 32 * 32;      // first thread of the control warp
    static constexpr int NT = CTRL0 + 32;
    static constexpr int KMAX = 16 * NPS;
    static constexpr int GT = 4 * D;                         // GEMV / W0-accumulator threads
    static constexpr int IQ = D / 4;                         // W0 rows per GEMV thread
    static_assert(GT <= ROWT, "GEMV threads are row threads");
    // shared memory, in floats
    static constexpr int OFF_W0 = 0;
    static constexpr int OFF_MEAN = OFF_W0 + D * D;          // [2][D]
    static constexpr int OFF_ACC = OFF_MEAN + 2 * D;         // [D] history column sums
    static constexpr int OFF_FPART = OFF_ACC + D;            // [4][D]
    static constexpr int OFF_UA = OFF_FPART + 4 * D;         // [D] the user's stored row (input of the aggregation)
    static constexpr int OFF_PW = OFF_UA + D;                // [D]
    static constexpr int OFF_GP = OFF_PW + D;                // [D] (spare)
    static constexpr int OFF_GU06 = OFF_GP + D;              // [D] 0.6 * user grad (W0 outer product)
    static constexpr int OFF_PART = OFF_GU06 + D;            // [16][D] user-grad partial sums of the negative slots
    static constexpr int OFF_HPART = OFF_PART + 16 * D;      // [NSLOT][D] history-gather partial sums
    static constexpr int OFF_SC = OFF_HPART + NSLOT * D;     // un[KMAX], nn[KMAX], pp, up
    static constexpr int OFF_META = OFF_SC + 2 * KMAX + 8;   // 2 x CtaMeta
    static constexpr int FLOATS = OFF_META + 2 * (int)(sizeof(CtaMeta) / 4) + 8;
};

// Barrier over the row threads only (named barrier 1): the control warp is not on the critical path of
// the sample being processed, it meets the row threads at the two full barriers [B] and [C].
template <int NROW> __device__ __forceinline__ void row_bar(bool is_ctrl)
{
    if (!is_ctrl) asm volatile("bar.sync 1, %0;" ::"n"(NROW) : "memory");
}

template <int VL> __device__ __forceinline__ float group_max(float v, unsigned mask)
{
#pragma unroll
    for (int o = VL / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(mask, v, o));
    return v;
}

// Control warp: fill `m` with the sample after (s_prev, chunk_end); fetches a new chunk when needed.
// Carries the run state (cur_u, H, age, regather) in registers and returns the history multiplicity of
// the sample's positive (for the patch of the cached column sums).
template <int D, int NPS>
__device__ __forceinline__ void cta_prepare(const TrainParams &P, CtaMeta *m, int lane, long long &s, long long &s_end,
                                            long long nchunks, int &cur_u, int &H, int &age, bool &regather, int &pmult,
                                            bool no_cache)
{
    constexpr int KMAX = 16 * NPS;
    if (s >= s_end) {                                      // next chunk (warp-uniform)
        unsigned long long c = 0;
        if (lane == 0) c = atomicAdd(P.chunk_counter, 1ULL);
        c = __shfl_sync(FULL, c, 0);
        if ((long long)c >= nchunks) {
            if (lane == 0) m->done = 1;
            return;
        }
        s = P.begin + (long long)c * P.chunk;
        s_end = min(P.end, s + (long long)P.chunk);
        cur_u = -1;
    }
    const int2 up = __ldg(&P.clicks[s]);
    const int u = up.x, p = up.y;
    const int K = P.K;
    bool dup = false, nhit = false;
    int myneg[NPS];
#pragma unroll
    for (int n = 0; n < NPS; ++n) {
        const int k = n * 16 + (lane & 15);
        int id = -1 - (lane & 15) - 16 * n;
        if (lane < 16 && k < K) id = sample_negative(P, s, k, p);
        id = __shfl_sync(FULL, id, lane & 15);             // both half-warps hold the 16 ids of this pass
        myneg[n] = id;
        if (lane < 16) m->negid[k] = id;
    }
    // duplicates among the negatives, or a negative equal to the positive: the row updates of such a
    // sample are replayed in the reference's sequential k order
#pragma unroll
    for (int n = 0; n < NPS; ++n) {
#pragma unroll
        for (int n2 = 0; n2 <= n; ++n2) {
            // compare pass n (lanes 0-15) against pass n2 (lanes 16-31)
            const int v = lane < 16 ? myneg[n] : myneg[n2];
            const unsigned mm = __match_any_sync(FULL, v);
            const unsigned other = (n == n2) ? (mm & 0xffffu & ~(1u << (lane & 15))) : (mm >> 16);
            dup |= (lane < 16) && myneg[n] >= 0 && other != 0;
        }
        dup |= (lane < 16) && myneg[n] == p;
    }
    dup = __any_sync(FULL, dup);
    const bool fresh = no_cache || u != cur_u || age >= 64 || regather;
    if (fresh) { H = __ldg(P.masks + u); cur_u = u; age = 0; }
    // history membership of the rows this sample rewrites (see train_kernel.cuh, CACHE)
    pmult = 0;
    const int *hrow = P.his + (size_t)u * P.max_his;
    for (int c0 = 0; c0 < H; c0 += 16) {
        const int hi = c0 + lane - 16;
        const int hv = (lane >= 16 && hi < H) ? __ldg(hrow + hi) : (-1000 - lane);
        pmult += __popc(__ballot_sync(FULL, hv == p));
#pragma unroll
        for (int n = 0; n < NPS; ++n) {
            const unsigned mm = __match_any_sync(FULL, lane < 16 ? myneg[n] : hv);
            nhit |= (lane < 16) && (mm >> 16) != 0;
        }
    }
    regather = __any_sync(FULL, nhit) || (pmult != 0 && dup);
    age += 1;
    if (lane == 0) { m->u = u; m->p = p; m->fresh = fresh; m->dup = dup; m->H = H; m->done = 0; m->s = s; m->pmult = pmult; m->regather = regather; }
    (void)KMAX;
    s += 1;
}


template <int D, int NPS>
__global__ void __launch_bounds__(CtaCfg<D, NPS>::NT, 1) train_cta_kernel(const TrainParams P)
{
    using C = CtaCfg<D, NPS>;
    constexpr int LPR = C::LPR, NSLOT = C::NSLOT, KMAX = C::KMAX, IQ = C::IQ;
    extern __shared__ __align__(16) float smem[];
    float *W0s = smem + C::OFF_W0, *meanb = smem + C::OFF_MEAN, *accs = smem + C::OFF_ACC, *fpart = smem + C::OFF_FPART;
    float *urow_s = smem + C::OFF_UA, *pw_s = smem + C::OFF_PW, *gu06_s = smem + C::OFF_GU06;
    float *part = smem + C::OFF_PART, *hpart = smem + C::OFF_HPART;
    float *sc_un = smem + C::OFF_SC, *sc_nn = sc_un + KMAX, *sc3 = sc_nn + KMAX;
    CtaMeta *meta = reinterpret_cast<CtaMeta *>(smem + C::OFF_META);

    const int tid = threadIdx.x, lane = tid & 31;
    const bool is_ctrl = tid >= C::CTRL0;
    const bool is_row = tid < C::ROWT;
    const int slot = tid / LPR, sl = tid % LPR;             // valid when is_row
    const bool is_neg = is_row && slot < 16, is_pos = is_row && slot == 16, is_usr = is_row && slot == 17;
    const bool is_gemv = tid < C::GT;
    const int gj = tid % D, giq = tid / D;                   // GEMV column / row block (valid when is_gemv)
    // lanes of this thread's row slot: slots of one warp take different branches, so every shuffle names its own group
    const unsigned gmask = LPR >= 32 ? FULL : (((1u << (LPR & 31)) - 1u) << ((lane / LPR) * LPR));
    const int gbase = (lane / LPR) * LPR;

    for (int i = tid * 4; i < D * D; i += C::NT * 4) *reinterpret_cast<float4 *>(W0s + i) = ldcg4(P.w0 + i);

    const float gamma = 0.4f, omg = 1.0f - gamma, eps = 1e-8f;
    const float score_mul = (float)(1.0 / 0.07);
    const int K = P.K;
    const long long nchunks = (P.end - P.begin + P.chunk - 1) / P.chunk;
    const bool no_cache = P.no_cache != 0;

    long long c_s = 0, c_end = 0;                            // control-warp state
    int c_u = -1, c_H = 0, c_age = 0, c_pmult = 0;
    bool c_regather = false;
    float4 u_keep = f4zero(), gu_keep = f4zero();            // user slot: the run's user row / grad row
    float w0acc[IQ];                                         // GEMV threads: W0 gradient accumulator
#pragma unroll
    for (int i = 0; i < IQ; ++i) w0acc[i] = 0.f;
    unsigned iter = 0;
    double loss_acc = 0.0;
    bool have_tail = false;                                  // W0 outer product of the previous sample pending
    int cur = 0;

    // W0 gradient of the sample whose mean is in `pm_all` (behavior_aggregators.cpp:131-146)
    auto w0_tail = [&](const float *pm_all, bool flush, bool refresh_smem) {
        const float *pm = pm_all + giq * IQ;
        const float gq = gu06_s[gj];
#pragma unroll
        for (int i4 = 0; i4 < IQ; i4 += 4) {
            const float4 m4 = *reinterpret_cast<const float4 *>(pm + i4);
            w0acc[i4 + 0] = fmaf(m4.x, gq, w0acc[i4 + 0]); w0acc[i4 + 1] = fmaf(m4.y, gq, w0acc[i4 + 1]);
            w0acc[i4 + 2] = fmaf(m4.z, gq, w0acc[i4 + 2]); w0acc[i4 + 3] = fmaf(m4.w, gq, w0acc[i4 + 3]);
        }
        if (flush) {
#pragma unroll
            for (int i = 0; i < IQ; ++i) {
                const int e = (giq * IQ + i) * D + gj;
                const float delta = -(P.lr_w0 * (w0acc[i] / 32.0f));
                const float old = atomicAdd(P.w0 + e, delta);
                if (refresh_smem) W0s[e] = old + delta;
                w0acc[i] = 0.f;
            }
        }
    };

    if (is_ctrl) {
        if (lane == 0) meta[0].done = 0;
        cta_prepare<D, NPS>(P, &meta[0], lane, c_s, c_end, nchunks, c_u, c_H, c_age, c_regather, c_pmult, no_cache);
    }
    __syncthreads();

    for (;;) {
        const CtaMeta *m = &meta[cur];
        if (m->done) break;
        const int u = m->u, p = m->p, H = m->H;
        const bool fresh = m->fresh != 0, dup = m->dup != 0;
        const float rH = (float)(1.0 / (double)H);          // behavior_aggregators.cpp:63
        float *mean = meanb + cur * D;

        // ---- issue the row loads of this sample
        int nid[NPS];
        float4 nw[NPS], gn[NPS];
        float4 rw = f4zero(), rg = f4zero();                  // positive / user weight and grad slices
        if (is_neg) {
#pragma unroll
            for (int n = 0; n < NPS; ++n) {
                const int k = slot + 16 * n;
                nid[n] = k < K ? m->negid[k] : 0;
                nw[n] = k < K ? ldcg4(P.item_w + (size_t)nid[n] * D + 4 * sl) : f4zero();
                gn[n] = k < K ? ldcg4(P.item_g + (size_t)nid[n] * D + 4 * sl) : f4zero();
            }
        } else if (is_pos) {
            rw = ldcg4(P.item_w + (size_t)p * D + 4 * sl);
            rg = ldcg4(P.item_g + (size_t)p * D + 4 * sl);
        } else if (is_usr) {
            if (fresh) {
                rw = ldcg4(P.user_w + (size_t)u * D + 4 * sl);
                rg = ldcg4(P.user_g + (size_t)u * D + 4 * sl);
            } else {
                rw = u_keep; rg = gu_keep;
            }
            *reinterpret_cast<float4 *>(urow_s + 4 * sl) = rw;
        }
        // ---- first sample of a run: gather the history rows (behavior_aggregators.cpp:96-105)
        if (fresh) {
            if (is_row) {
                const int *hrow = P.his + (size_t)u * P.max_his;
                float4 a4 = f4zero();
                for (int h = slot; h < H; h += NSLOT) a4 = a4 + ldcg4(P.item_w + (size_t)__ldg(hrow + h) * D + 4 * sl);
                *reinterpret_cast<float4 *>(hpart + slot * D + 4 * sl) = a4;
            }
            row_bar<C::CTRL0>(is_ctrl);
            if (tid < D) {
                float a = 0.f;
#pragma unroll
                for (int q = 0; q < NSLOT; ++q) a += hpart[q * D + tid];
                accs[tid] = a;
                mean[tid] = a * rH;
            }
            row_bar<C::CTRL0>(is_ctrl);
        }
        // ---- a W0 step is due (every 32 samples): it must land before this sample's mean.W0
        const bool flush_due = have_tail && (iter & 31) == 0;
        if (flush_due) {
            row_bar<C::CTRL0>(is_ctrl);                       // the user slot has published 0.6*g_u of the previous sample
            if (is_gemv) w0_tail(meanb + (cur ^ 1) * D, true, true);
        }
        // ---- f = mean . W0 as four partial GEMVs (behavior_aggregators.cpp:118)
        if (is_gemv) {
            const float *pm = mean + giq * IQ;
            float f0 = 0.f, f1 = 0.f, f2 = 0.f, f3 = 0.f;
#pragma unroll
            for (int i4 = 0; i4 < IQ; i4 += 4) {
                const float4 m4 = *reinterpret_cast<const float4 *>(pm + i4);
                const float *w = W0s + (giq * IQ + i4) * D + gj;
                f0 = fmaf(m4.x, w[0], f0); f1 = fmaf(m4.y, w[D], f1); f2 = fmaf(m4.z, w[2 * D], f2); f3 = fmaf(m4.w, w[3 * D], f3);
            }
            fpart[giq * D + gj] = (f0 + f1) + (f2 + f3);
        }
        // ---- control warp: prepare the next sample while this one is processed (meets the others at [B])
        if (is_ctrl) {
            if (lane == 0) meta[cur ^ 1].done = 0;
            cta_prepare<D, NPS>(P, &meta[cur ^ 1], lane, c_s, c_end, nchunks, c_u, c_H, c_age, c_regather, c_pmult, no_cache);
        }
        row_bar<C::CTRL0>(is_ctrl);                                                   // [G1]
        if (have_tail && !flush_due && is_gemv) w0_tail(meanb + (cur ^ 1) * D, false, false);
        // ---- every row slot: u <- gamma*u + (1-gamma)*f (:122), uu
        float4 ua = f4zero();
        float uu = 0.f;
        if (is_row) {
            const float4 f0 = *reinterpret_cast<const float4 *>(fpart + 4 * sl), f1 = *reinterpret_cast<const float4 *>(fpart + D + 4 * sl);
            const float4 f2 = *reinterpret_cast<const float4 *>(fpart + 2 * D + 4 * sl), f3 = *reinterpret_cast<const float4 *>(fpart + 3 * D + 4 * sl);
            const float4 f = (f0 + f1) + (f2 + f3);
            ua = gamma * *reinterpret_cast<const float4 *>(urow_s + 4 * sl);
            fma4(ua, omg, f);
            uu = group_sum<LPR>(dot4(ua, ua), gmask);
        }
        // ---- dots (matrix_factorization.cpp:45-47, 82-85)
        if (is_neg) {
#pragma unroll
            for (int n = 0; n < NPS; ++n) {
                const int k = slot + 16 * n;
                const float un_ = group_sum<LPR>(dot4(ua, nw[n]), gmask), nn_ = group_sum<LPR>(dot4(nw[n], nw[n]), gmask);
                if (sl == 0 && k < K) { sc_un[k] = un_; sc_nn[k] = nn_; }
            }
        } else if (is_pos) {
            const float pp_ = group_sum<LPR>(dot4(rw, rw), gmask), up_ = group_sum<LPR>(dot4(ua, rw), gmask);
            if (sl == 0) { sc3[1] = pp_; sc3[2] = up_; }
            *reinterpret_cast<float4 *>(pw_s + 4 * sl) = rw;
        }
        row_bar<C::CTRL0>(is_ctrl);                                                   // [A]
        // ---- scores and softmax, recomputed by every row slot from the shared scalars
        //      (matrix_factorization.cpp:55-61, 91-109).  log(1 + sum_k exp(score_k)) is evaluated without
        //      the reference's max shift: |score| <= 2/0.07, far inside fp32 range, same value to rounding.
        float lgsum = 0.f, bsum = 0.f, esum = 0.f, unorm = 0.f, unorm3 = 0.f, pnorm = 0.f, pp = 0.f, upd = 0.f;
        constexpr int SPL = (KMAX + LPR - 1) / LPR;
        float k1[SPL], k2[SPL], k3[SPL];                       // per-k update coefficients, lane sl holds k = sl + LPR*q
        if (is_row) {
            pp = sc3[1]; upd = sc3[2];
            unorm = sqrtf(fmaxf(uu, eps)); pnorm = sqrtf(fmaxf(pp, eps));
            unorm3 = unorm * unorm * unorm;
            const float cos_p = upd / (unorm * pnorm);
            float ev[SPL], r3[SPL], rn3[SPL], unv[SPL], nnv[SPL];
#pragma unroll
            for (int q = 0; q < SPL; ++q) {
                const int k = sl + LPR * q;
                ev[q] = 0.f; r3[q] = 0.f; rn3[q] = 0.f; unv[q] = 0.f; nnv[q] = 0.f;
                if (k < K) {
                    nnv[q] = sc_nn[k]; unv[q] = sc_un[k];
                    const float nnorm = sqrtf(nnv[q] < eps ? eps : nnv[q]);
                    ev[q] = expf((unv[q] / (unorm * nnorm) - cos_p) * score_mul);
                    r3[q] = 1.f / (unorm3 * nnorm);                        // :136
                    rn3[q] = 1.f / (unorm * (nnorm * nnorm * nnorm));      // :137
                }
                esum += ev[q];
            }
            esum = group_sum<LPR>(esum, gmask) + 1.0f;         // the implicit zero logit (:107)
            const float inv_esum = 1.f / esum;
#pragma unroll
            for (int q = 0; q < SPL; ++q) {
                const float lg = (ev[q] * inv_esum) * score_mul;   // :109
                lgsum += lg; bsum = fmaf(lg, -(r3[q] * unv[q]), bsum);
                k1[q] = lg * rn3[q] * nnv[q];                  // lg*gn_n = k1*u + k2*N[k] (:139,:143)
                k2[q] = -(lg * rn3[q]) * unv[q];
                k3[q] = lg * r3[q] * uu;                       // the N[k] part of lg*gu_n (:138,:141)
            }
            lgsum = group_sum<LPR>(lgsum, gmask);
            bsum = group_sum<LPR>(bsum, gmask);
        }
        // ---- negative rows: grads, clipped step, write-back (:127-150; sgd.cpp:14-26)
        if (is_neg) {
            float4 gpart = f4zero();
            float c1[NPS], c2[NPS];
#pragma unroll
            for (int n = 0; n < NPS; ++n) {
                const int k = slot + 16 * n;                   // its coefficients live in lane k % LPR, entry k / LPR
                float c3 = 0.f;
                c1[n] = 0.f; c2[n] = 0.f;
#pragma unroll
                for (int q = 0; q < SPL; ++q) {
                    const float a1 = __shfl_sync(gmask, k1[q], gbase + (k % LPR));
                    const float a2 = __shfl_sync(gmask, k2[q], gbase + (k % LPR));
                    const float a3 = __shfl_sync(gmask, k3[q], gbase + (k % LPR));
                    if (q == k / LPR && k < K) { c1[n] = a1; c2[n] = a2; c3 = a3; }
                }
                fma4(gpart, c3, nw[n]);
            }
            *reinterpret_cast<float4 *>(part + slot * D + 4 * sl) = gpart;
            if (!dup) {
#pragma unroll
                for (int n = 0; n < NPS; ++n) {
                    if (slot + 16 * n < K) {
                        float4 gk = gn[n];
                        fma4(gk, c1[n], ua); fma4(gk, c2[n], nw[n]);
                        gk = clamp4(gk, P.clip);
                        float4 wk = nw[n];
                        fma4(wk, -P.lr, gk);
                        stcg4(P.item_w + (size_t)nid[n] * D + 4 * sl, wk);   // :148
                        stcg4(P.item_g + (size_t)nid[n] * D + 4 * sl, gk);   // :149
                    }
                }
            } else {                                         // keep the coefficients for the replay below
#pragma unroll
                for (int n = 0; n < NPS; ++n) gn[n] = make_float4(c1[n], c2[n], 0.f, 0.f);
            }
        }
        if (dup) {
            // duplicate ids: replay in the reference's k order (grad row re-read per k, last writer
            // wins on the weight row); the positive row is written after all negatives (:169-174)
            for (int k = 0; k < K; ++k) {
                if (is_neg && slot == (k & 15)) {
#pragma unroll
                    for (int n = 0; n < NPS; ++n) {
                        if (n == (k >> 4)) {
                            float *grow = P.item_g + (size_t)nid[n] * D + 4 * sl;
                            float4 gk = ldcg4(grow);
                            fma4(gk, gn[n].x, ua); fma4(gk, gn[n].y, nw[n]);
                            gk = clamp4(gk, P.clip);
                            float4 wk = nw[n];
                            fma4(wk, -P.lr, gk);
                            stcg4(P.item_w + (size_t)nid[n] * D + 4 * sl, wk);
                            stcg4(grow, gk);
                        }
                    }
                }
                row_bar<C::CTRL0>(is_ctrl);
            }
        }
        // ---- positive row (:142, :169, :173-174); patch of the cached history sums
        if (is_pos) {
            const float r_u_p3 = 1.f / (unorm * (pnorm * pnorm * pnorm));
            float4 gp = rg;
            fma4(gp, -(lgsum * r_u_p3) * pp, ua); fma4(gp, (lgsum * r_u_p3) * upd, rw);   // lgsum * gp_p, gp_p = -(pp*u - up*p)*r
            gp = clamp4(gp, P.clip);
            float4 wp = rw;
            fma4(wp, -P.lr, gp);
            stcg4(P.item_w + (size_t)p * D + 4 * sl, wp);
            stcg4(P.item_g + (size_t)p * D + 4 * sl, gp);
            // I[p] moved by -lr*gp: carry that into the column sums and publish the mean of the run's next
            // sample (if the next sample starts a new run or re-gathers, the gather overwrites both)
            float4 a4 = *reinterpret_cast<const float4 *>(accs + 4 * sl);
            fma4(a4, m->regather ? 0.f : -P.lr * (float)m->pmult, gp);
            *reinterpret_cast<float4 *>(accs + 4 * sl) = a4;
            *reinterpret_cast<float4 *>(meanb + (cur ^ 1) * D + 4 * sl) = rH * a4;
            if (sl == 0) {
                const float loss = logf(esum);                   // :108
                loss_acc += (double)loss;
                if (P.loss_out) P.loss_out[m->s - P.begin] = loss;
            }
        }
        __syncthreads();                                                              // [B]
        // ---- user row: grad, aggregator backward, step (:141, :152, :166, :171-172); overlaps the next sample's loads
        if (is_usr) {
            float4 gacc = f4zero();
#pragma unroll
            for (int q = 0; q < 16; ++q) gacc = gacc + *reinterpret_cast<const float4 *>(part + q * D + 4 * sl);
            const float4 pw = *reinterpret_cast<const float4 *>(pw_s + 4 * sl);
            const float r_u3_p = 1.f / (unorm3 * pnorm);
            // + sum_k lg_k*(-r*un_k)*u  - lgsum*gu_p,  gu_p = (uu*p - up*u)*r_u3_p  (:60-62)
            fma4(gacc, bsum + lgsum * r_u3_p * upd, ua);
            fma4(gacc, -(lgsum * r_u3_p) * uu, pw);
            float4 gu = rg + gacc;
            *reinterpret_cast<float4 *>(gu06_s + 4 * sl) = omg * gu;      // behavior_aggregators.cpp:131-139
            gu = gamma * gu;                                               // :148-152
            gu = clamp4(gu, P.clip);
            fma4(ua, -P.lr, gu);
            stcg4(P.user_w + (size_t)u * D + 4 * sl, ua);
            stcg4(P.user_g + (size_t)u * D + 4 * sl, gu);
            u_keep = ua; gu_keep = gu;
        }
        iter += 1;
        have_tail = true;
        cur ^= 1;
    }
    // ---- W0 gradient of the last sample and the leftover (< 32 samples) accumulators
    __syncthreads();
    if (have_tail && is_gemv) w0_tail(meanb + (cur ^ 1) * D, ((iter & 31) == 0) || !P.drop_w0_tail, false);
    if (is_pos && sl == 0 && loss_acc != 0.0) atomicAdd(P.loss_sum, loss_acc);
}

}  // namespace heat
