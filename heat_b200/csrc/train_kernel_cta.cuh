// heat_b200/csrc/train_kernel_cta.cuh -- latency-oriented training kernel: one CTA per sample.
//
// Why: the reference's update is a clipped RUNNING SUM applied at every touch, so its dynamics depend
// on how many samples are in flight: the loss / Recall of the reference at 1-128
// threads is reproduced only while a few hundred samples at most are being processed concurrently.
// Throughput at a FIXED number of samples in flight is (samples in flight) / (latency of one sample),
// so this kernel spends a whole CTA on one sample to make that latency small:
//
//   neg slots 0..15 (D/4 lanes each, whole warps)  own the negative rows k = slot, slot+16, ... (NPS per slot)
//   pos warp (D/4 lanes used)                       owns the positive row, the loss and the history-sum patch
//   usr warp (D/4 lanes used)                       owns the user row (kept in registers across a user's run)
//   threads 0..4D-1 (= the neg slots)               also compute mean.W0 as 4 partial GEMVs against a
//                                                   REGISTER-resident slice of W0 (D <= 64) and own the
//                                                   matching D/4 elements of the W0 gradient accumulator
//   control warp                                    prepares the NEXT sample while this one is processed:
//                                                   click, Philox negatives, duplicate detection, history
//                                                   membership against the run's history ids (registers)
//
// Per sample, three block barriers and one L2 round trip on the critical path:
//   issue all row loads | mean.W0  -[G1]-  u <- 0.4u + 0.6f, the 3 dots of every slot in one interleaved
//   butterfly, E_k = exp(cos_k/T)  -[A]-  esum = 1 + exp(-cos_p/T)*sum_k E_k from 16 shared partials (no
//   cross-slot shuffles), coefficients, clipped row steps, stores, W0 outer product of the PREVIOUS
//   sample  -[B]-  user step (overlaps the next sample's loads and GEMV).
// Same arithmetic as train_kernel.cuh (reference: models/matrix_factorization.cpp:15-181,
// behavior_aggregators/behavior_aggregators.cpp:51-153, optimizers/sgd.cpp:14-26) up to fp32 rounding:
// 1/sqrt via MUFU.RSQ (2 ulp), softmax as exp(a)*exp(-b) instead of exp(a-b-max).  With one CTA the
// samples are applied strictly in file order (deterministic mode).
#pragma once
#include "train_kernel.cuh"

namespace heat {

struct CtaMeta {            // one sample, prepared by the control warp
    int u, p, fresh, dup, H, done, pmult, regather;
    long long s;
    float rH;
    int late;               // shares an item row with the previous sample: its rows may not be prefetched
    int negid[64];
};

constexpr int CTA_HIS_REGS = 4;                              // history ids per control-warp lane
constexpr int CTA_HIS_MAX = 32 * CTA_HIS_REGS;              // max_his served from registers / shared memory

template <int D, int NPS> struct CtaCfg {
    static constexpr int LPR = D / 4;                       // lanes per row
    static constexpr int NEGT = 16 * LPR;                    // threads of the negative slots (whole warps)
    static constexpr int SLOTW = LPR < 32 ? 32 : LPR;        // threads reserved for the pos / usr slot (one warp)
    static constexpr int POS0 = NEGT, USR0 = POS0 + SLOTW, CTRL0 = USR0 + SLOTW;
    static constexpr int NT = CTRL0 + 32;
    static constexpr int NSLOT = 18;
    static constexpr int KMAX = 16 * NPS;
    static constexpr int GT = 4 * D;                         // GEMV / W0-accumulator threads
    static constexpr int IQ = D / 4;                         // W0 rows per GEMV thread
    static constexpr bool WREG = D <= 64;                    // W0 slice in registers (else shared memory)
    static constexpr bool PF = D <= 64;                      // rows of the next sample prefetched (cp.async) during this one
    static_assert(GT == NEGT, "GEMV threads are exactly the negative-slot threads");
    // shared memory, in floats
    static constexpr int OFF_W0 = 0;
    static constexpr int OFF_MEAN = OFF_W0 + (WREG ? 0 : D * D);   // [3][D]
    static constexpr int OFF_ACC = OFF_MEAN + 3 * D;         // [D] history column sums
    static constexpr int OFF_FPART = OFF_ACC + D;            // [4][D]
    static constexpr int OFF_UROW = OFF_FPART + 4 * D;       // [D] the user's stored row (input of the aggregation)
    static constexpr int OFF_PW = OFF_UROW + D;              // [D] positive row snapshot (user step)
    static constexpr int OFF_GU06 = OFF_PW + D;              // [D] 0.6 * user grad (W0 outer product)
    static constexpr int OFF_PART = OFF_GU06 + D;            // [16][D] user-grad partial sums of the negative slots
    static constexpr int OFF_HPART = OFF_PART + 16 * D;      // [NSLOT][D] history-gather partial sums
    static constexpr int OFF_SC = OFF_HPART + NSLOT * D;     // sE[16], sB[16], pp, up, Ep
    static constexpr int OFF_HID = OFF_SC + 40;              // [3][CTA_HIS_MAX] history ids of the run (ints)
    static constexpr int OFF_META = OFF_HID + 3 * CTA_HIS_MAX;   // 3 x CtaMeta (this sample, next, the one being prepared)
    static constexpr int OFF_STAGE = OFF_META + 3 * (int)(sizeof(CtaMeta) / 4) + 8;   // [2*NPS][CTRL0] float4: prefetched rows
    static constexpr int FLOATS = OFF_STAGE + (PF ? 2 * NPS * CTRL0 * 4 : 0);
};

// Barrier over the row threads only (named barrier 1): the control warp is not on the critical path of
// the sample being processed, it meets the row threads at the full barrier [B].
template <int NROW> __device__ __forceinline__ void row_bar(bool is_ctrl)
{
    if (!is_ctrl) asm volatile("bar.sync 1, %0;" ::"n"(NROW) : "memory");
}

// N independent sums over the VL lanes of a row group in ONE butterfly: the N shuffles of a step are
// issued back to back, so the latency of a step is paid once, not N times.
template <int VL, int N> __device__ __forceinline__ void group_sum_n(float (&v)[N], unsigned mask)
{
#pragma unroll
    for (int o = VL / 2; o > 0; o >>= 1) {
        float t[N];
#pragma unroll
        for (int i = 0; i < N; ++i) t[i] = __shfl_xor_sync(mask, v[i], o);
#pragma unroll
        for (int i = 0; i < N; ++i) v[i] += t[i];
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

// 16-byte asynchronous copy global -> shared (LDGSTS, L2-coherent): the row prefetch.  Every thread copies
// exactly the slice it consumes later, so completion needs no cross-thread synchronisation.
__device__ __forceinline__ void cp_async16(float *smem_dst, const float *gsrc)
{
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Control-warp state carried across samples.
template <int NPS> struct CtaCtrl {
    long long s = 0, s_end = 0;
    int cur_u = -1, ids_u = -1, H = 0, age = 0;
    bool regather = false;
    float rH = 0.f;
    long long blk0 = 0;                                      // lane holds click blk0 + lane (clk) and blk0 + 32 + lane (clkn)
    int2 clk = make_int2(0, 0), clkn = make_int2(0, 0);
    int hreg[CTA_HIS_REGS];                                  // history ids lane, lane+32, ... of user ids_u
    int prevneg[NPS];                                        // ids of the previously prepared sample
    int prev_p = -5000;
};

// Control warp: fill `m` with the next sample of this CTA (fetches a new chunk when needed) and, when the
// sample starts a run or re-gathers, publish the run's history ids in `hid_dst`.
template <int D, int NPS>
__device__ __forceinline__ void cta_prepare(const TrainParams &P, CtaMeta *m, int *hid_dst, int lane, CtaCtrl<NPS> &c,
                                            long long nchunks, bool no_cache)
{
    if (c.s >= c.s_end) {                                  // next chunk (warp-uniform)
        unsigned long long ch = 0;
        if (lane == 0) ch = atomicAdd(P.chunk_counter, 1ULL);
        ch = __shfl_sync(FULL, ch, 0);
        if ((long long)ch >= nchunks) {
            if (lane == 0) m->done = 1;
            return;
        }
        c.s = P.begin + (long long)ch * P.chunk;
        c.s_end = min(P.end, c.s + (long long)P.chunk);
        c.cur_u = -1;
        c.blk0 = c.s;
        c.clk = (c.s + lane < c.s_end) ? __ldg(&P.clicks[c.s + lane]) : make_int2(0, 0);
        c.clkn = (c.s + 32 + lane < c.s_end) ? __ldg(&P.clicks[c.s + 32 + lane]) : make_int2(0, 0);
    } else if (c.s - c.blk0 == 32) {                       // next block of 32 clicks of this chunk
        c.blk0 = c.s;
        c.clk = c.clkn;
        c.clkn = (c.s + 32 + lane < c.s_end) ? __ldg(&P.clicks[c.s + 32 + lane]) : make_int2(0, 0);
    }
    const long long s = c.s;
    const int u = __shfl_sync(FULL, c.clk.x, (int)(s - c.blk0)), p = __shfl_sync(FULL, c.clk.y, (int)(s - c.blk0));
    const int K = P.K;
    const bool big_his = P.max_his > CTA_HIS_MAX;
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
    // an item row shared with the previously prepared sample (which rewrites it): no prefetch
    bool conf = p == c.prev_p;
#pragma unroll
    for (int n = 0; n < NPS; ++n) {
#pragma unroll
        for (int n2 = 0; n2 < NPS; ++n2) {
            const unsigned mm = __match_any_sync(FULL, lane < 16 ? myneg[n] : c.prevneg[n2]);
            conf |= (lane < 16) && myneg[n] >= 0 && (mm >> 16) != 0;
        }
        conf |= myneg[n] == c.prev_p || c.prevneg[n] == p;
    }
    const bool late = __any_sync(FULL, conf);
#pragma unroll
    for (int n = 0; n < NPS; ++n) c.prevneg[n] = myneg[n];
    c.prev_p = p;
    const bool fresh = no_cache || u != c.cur_u || c.age >= 64 || c.regather;
    if (fresh) {
        if (u != c.ids_u) {                                // a new user: history length and ids
            c.H = __ldg(P.masks + u);
            c.rH = (float)(1.0 / (double)c.H);             // behavior_aggregators.cpp:63
            c.ids_u = u;
            if (!big_his) {
                const int *hrow = P.his + (size_t)u * P.max_his;
#pragma unroll
                for (int j = 0; j < CTA_HIS_REGS; ++j) {
                    const int idx = lane + 32 * j;
                    c.hreg[j] = idx < c.H ? __ldg(hrow + idx) : (-1000 - idx);
                }
            }
        }
        c.cur_u = u; c.age = 0;
        if (!big_his) {
#pragma unroll
            for (int j = 0; j < CTA_HIS_REGS; ++j) hid_dst[lane + 32 * j] = c.hreg[j];
        }
    }
    // history membership of the rows this sample rewrites (see train_kernel.cuh, CACHE)
    int pmult = 0;
    if (!big_his) {
#pragma unroll
        for (int j = 0; j < CTA_HIS_REGS; ++j) pmult += __popc(__ballot_sync(FULL, c.hreg[j] == p));
#pragma unroll
        for (int n = 0; n < NPS; ++n) {
#pragma unroll
            for (int t = 0; t < 16; ++t) {
                const int v = __shfl_sync(FULL, myneg[n], t);
#pragma unroll
                for (int j = 0; j < CTA_HIS_REGS; ++j) nhit |= c.hreg[j] == v;
            }
        }
    } else {
        const int *hrow = P.his + (size_t)u * P.max_his;
        for (int c0 = 0; c0 < c.H; c0 += 16) {
            const int hi = c0 + lane - 16;
            const int hv = (lane >= 16 && hi < c.H) ? __ldg(hrow + hi) : (-1000 - lane);
            pmult += __popc(__ballot_sync(FULL, hv == p));
#pragma unroll
            for (int n = 0; n < NPS; ++n) {
                const unsigned mm = __match_any_sync(FULL, lane < 16 ? myneg[n] : hv);
                nhit |= (lane < 16) && (mm >> 16) != 0;
            }
        }
    }
    c.regather = __any_sync(FULL, nhit) || (pmult != 0 && dup);
    c.age += 1;
    if (lane == 0) {
        m->u = u; m->p = p; m->fresh = fresh; m->dup = dup; m->H = c.H; m->done = 0; m->s = s;
        m->pmult = pmult; m->regather = c.regather; m->rH = c.rH; m->late = late;
    }
    c.s = s + 1;
}

// Phase timing for kernel development (-DHEAT_CTA_PROFILE through HEAT_NVCC_DEFS, build.py): one thread of
// neg slot 0, the pos slot, the usr slot and the control warp print clock64() at the marked points for a
// few iterations of block 0.  (clock reads placed right after a barrier do not include the wait: the
// barrier blocks the warp at its next dependent instruction; compare absolute times across warps.)
#ifdef HEAT_CTA_PROFILE
#define HEAT_PT(i) do { long long t_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_) :: "memory"); if (prof_on) prof_abs[i] = t_; } while (0)
#else
#define HEAT_PT(i) do { } while (0)
#endif

template <int D, int NPS>
__global__ void __launch_bounds__(CtaCfg<D, NPS>::NT, 1) train_cta_kernel(const TrainParams P)
{
    using C = CtaCfg<D, NPS>;
    constexpr int LPR = C::LPR, NSLOT = C::NSLOT, IQ = C::IQ;
    constexpr bool WREG = C::WREG, PF = C::PF;
    extern __shared__ __align__(16) float smem[];
    float *W0s = smem + C::OFF_W0, *meanb = smem + C::OFF_MEAN, *accs = smem + C::OFF_ACC, *fpart = smem + C::OFF_FPART;
    float *urow_s = smem + C::OFF_UROW, *pw_s = smem + C::OFF_PW, *gu06_s = smem + C::OFF_GU06;
    float *part = smem + C::OFF_PART, *hpart = smem + C::OFF_HPART;
    float *sE = smem + C::OFF_SC, *sB = sE + 16, *sc3 = sB + 16;      // sc3: pp, up, Ep
    int *hid_s = reinterpret_cast<int *>(smem + C::OFF_HID);
    CtaMeta *meta = reinterpret_cast<CtaMeta *>(smem + C::OFF_META);
    float *stage = smem + C::OFF_STAGE + 4 * threadIdx.x;    // this thread's prefetch slots: [2*n] weight, [2*n+1] grad slice
    constexpr int STG = C::CTRL0 * 4;                        // floats between two slots

    const int tid = threadIdx.x, lane = tid & 31;
    const bool is_ctrl = tid >= C::CTRL0;
    const bool is_neg = tid < C::NEGT;                       // warp-uniform
    const bool is_pos = tid >= C::POS0 && tid < C::POS0 + LPR;
    const bool is_usr = tid >= C::USR0 && tid < C::USR0 + LPR;
    const bool is_row = is_neg || is_pos || is_usr;
    const int slot = is_neg ? tid / LPR : (is_pos ? 16 : 17), sl = tid % LPR;
    const int gj = tid % D, giq = tid / D;                   // GEMV column / row block (neg threads)
    const unsigned smask = LPR >= 32 ? FULL : ((1u << (LPR & 31)) - 1u);   // the pos / usr slot's lanes of its warp

    const float gamma = 0.4f, omg = 1.0f - gamma;
    const float score_mul = (float)(1.0 / 0.07);
    const int K = P.K;
    const long long nchunks = (P.end - P.begin + P.chunk - 1) / P.chunk;
    const bool no_cache = P.no_cache != 0;
    const bool big_his = P.max_his > CTA_HIS_MAX;

    // this thread's slice of W0 (rows giq*IQ .. +IQ, column gj) and of its gradient accumulator
    float wreg[WREG ? IQ : 1], w0acc[IQ];
#pragma unroll
    for (int i = 0; i < IQ; ++i) w0acc[i] = 0.f;
    if (WREG) {
        if (is_neg) {
#pragma unroll
            for (int i = 0; i < IQ; ++i) wreg[WREG ? i : 0] = __ldcg(P.w0 + (giq * IQ + i) * D + gj);
        }
    } else {
        for (int i = tid * 4; i < D * D; i += C::NT * 4) *reinterpret_cast<float4 *>(W0s + i) = ldcg4(P.w0 + i);
    }

    CtaCtrl<NPS> ctl;
#pragma unroll
    for (int n = 0; n < NPS; ++n) ctl.prevneg[n] = -2000 - n;
    float4 u_keep = f4zero(), gu_keep = f4zero();            // usr slot: the row / grad row of user keep_u after its last step
    int keep_u = -1, pf_u = -1;                              // usr slot: user of u_keep; user whose row is in the prefetch slot
    bool have_pf = false;                                    // the prefetch slots hold this sample's item rows
    int nidn[NPS];
    unsigned iter = 0;
    double loss_acc = 0.0;
    float loss_f = 0.f;
    bool have_tail = false;                                  // W0 outer product of the previous sample pending
    int cur = 0, mb = 0;                                     // meta / history-id buffer and mean buffer of this sample (rings of 3)

    // W0 gradient of the sample whose mean is `pm_all` (behavior_aggregators.cpp:131-146); every 32 samples
    // the accumulated step is applied (W0 -= lr0 * acc / 32) and the thread's copy of W0 refreshed
    auto w0_tail = [&](const float *pm_all, bool flush) {
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
                const float old = atomicAdd(P.w0 + e, delta);    // returns the value before: picks up other CTAs' steps
                if (WREG) wreg[WREG ? i : 0] = old + delta; else W0s[e] = old + delta;
                w0acc[i] = 0.f;
            }
        }
    };

#ifdef HEAT_CTA_PROFILE
    const bool prof_on = blockIdx.x == 0 && lane == 0;
    long long prof_abs[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
#endif
    if (is_ctrl) {
        if (lane == 0) meta[0].done = 0;
        cta_prepare<D, NPS>(P, &meta[0], hid_s, lane, ctl, nchunks, no_cache);
        if (lane == 0) meta[1].done = 0;
        cta_prepare<D, NPS>(P, &meta[1], hid_s + CTA_HIS_MAX, lane, ctl, nchunks, no_cache);
    }
    __syncthreads();

    for (;;) {
        const CtaMeta *m = &meta[cur];
        if (m->done) break;
        const int u = m->u, p = m->p, H = m->H;
        const bool fresh = m->fresh != 0, dup = m->dup != 0;
        const float rH = m->rH;
        float *mean = meanb + mb * D;
        const float *mean_prev = meanb + (mb == 0 ? 2 : mb - 1) * D;
        float *mean_next = meanb + (mb == 2 ? 0 : mb + 1) * D;
        HEAT_PT(0);

        // ---- the rows of this sample: prefetched during the previous sample, or loaded now
        const int nx = cur == 2 ? 0 : cur + 1, n2 = nx == 2 ? 0 : nx + 1;
        const CtaMeta *mn = &meta[nx];
        const bool pf = PF && !mn->done && !mn->late;          // block-uniform
        int nid[NPS];
        float4 nw[NPS], gn[NPS];
        float4 rw = f4zero(), rg = f4zero();                  // positive / user weight and grad slices
        if (is_neg) {
            if (have_pf) cp_async_wait0();
#pragma unroll
            for (int n = 0; n < NPS; ++n) {
                const int k = slot + 16 * n;
                if (have_pf) {
                    nid[n] = nidn[n];
                    nw[n] = k < K ? *reinterpret_cast<const float4 *>(stage + (2 * n) * STG) : f4zero();
                    gn[n] = k < K ? *reinterpret_cast<const float4 *>(stage + (2 * n + 1) * STG) : f4zero();
                } else {
                    nid[n] = k < K ? m->negid[k] : 0;
                    nw[n] = k < K ? ldcg4(P.item_w + (size_t)nid[n] * D + 4 * sl) : f4zero();
                    gn[n] = k < K ? ldcg4(P.item_g + (size_t)nid[n] * D + 4 * sl) : f4zero();
                }
            }
            if (pf) {
#pragma unroll
                for (int n = 0; n < NPS; ++n) {
                    const int k = slot + 16 * n;
                    nidn[n] = k < K ? mn->negid[k] : 0;
                    if (k < K) {
                        cp_async16(stage + (2 * n) * STG, P.item_w + (size_t)nidn[n] * D + 4 * sl);
                        cp_async16(stage + (2 * n + 1) * STG, P.item_g + (size_t)nidn[n] * D + 4 * sl);
                    }
                }
                cp_async_commit();
            }
        } else if (is_pos) {
            if (have_pf) {
                cp_async_wait0();
                rw = *reinterpret_cast<const float4 *>(stage); rg = *reinterpret_cast<const float4 *>(stage + STG);
            } else {
                rw = ldcg4(P.item_w + (size_t)p * D + 4 * sl);
                rg = ldcg4(P.item_g + (size_t)p * D + 4 * sl);
            }
            if (pf) {
                cp_async16(stage, P.item_w + (size_t)mn->p * D + 4 * sl);
                cp_async16(stage + STG, P.item_g + (size_t)mn->p * D + 4 * sl);
                cp_async_commit();
            }
        } else if (is_usr) {
            if (!fresh || u == keep_u) {                      // the run continues: the row is in registers
                rw = u_keep; rg = gu_keep;
            } else if (pf_u == u) {
                cp_async_wait0();
                rw = *reinterpret_cast<const float4 *>(stage); rg = *reinterpret_cast<const float4 *>(stage + STG);
            } else {
                rw = ldcg4(P.user_w + (size_t)u * D + 4 * sl);
                rg = ldcg4(P.user_g + (size_t)u * D + 4 * sl);
            }
            *reinterpret_cast<float4 *>(urow_s + 4 * sl) = rw;
            pf_u = -1;
            if (pf && mn->fresh && mn->u != u) {              // the next sample starts another user's run
                pf_u = mn->u;
                cp_async16(stage, P.user_w + (size_t)pf_u * D + 4 * sl);
                cp_async16(stage + STG, P.user_g + (size_t)pf_u * D + 4 * sl);
                cp_async_commit();
            }
        }
        have_pf = pf;
        HEAT_PT(1);
        // ---- first sample of a run: gather the history rows (behavior_aggregators.cpp:96-105); the ids come
        //      from the control warp's copy in shared memory
        if (fresh) {
            if (is_row) {
                const int *hg = P.his + (size_t)u * P.max_his;
                const int *hs = hid_s + cur * CTA_HIS_MAX;
                float4 a4 = f4zero();
                for (int h = slot; h < H; h += NSLOT) {
                    const int id = big_his ? __ldg(hg + h) : hs[h];
                    a4 = a4 + ldcg4(P.item_w + (size_t)id * D + 4 * sl);
                }
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
            row_bar<C::CTRL0>(is_ctrl);                       // the usr slot has published 0.6*g_u of the previous sample
            if (is_neg) w0_tail(mean_prev, true);
        }
        HEAT_PT(8);
        // ---- f = mean . W0 as four partial GEMVs (behavior_aggregators.cpp:118)
        if (is_neg) {
            const float *pm = mean + giq * IQ;
            float f0 = 0.f, f1 = 0.f, f2 = 0.f, f3 = 0.f;
#pragma unroll
            for (int i4 = 0; i4 < IQ; i4 += 4) {
                const float4 m4 = *reinterpret_cast<const float4 *>(pm + i4);
                if (WREG) {
                    f0 = fmaf(m4.x, wreg[WREG ? i4 : 0], f0); f1 = fmaf(m4.y, wreg[WREG ? i4 + 1 : 0], f1);
                    f2 = fmaf(m4.z, wreg[WREG ? i4 + 2 : 0], f2); f3 = fmaf(m4.w, wreg[WREG ? i4 + 3 : 0], f3);
                } else {
                    const float *w = W0s + (giq * IQ + i4) * D + gj;
                    f0 = fmaf(m4.x, w[0], f0); f1 = fmaf(m4.y, w[D], f1); f2 = fmaf(m4.z, w[2 * D], f2); f3 = fmaf(m4.w, w[3 * D], f3);
                }
            }
            fpart[giq * D + gj] = (f0 + f1) + (f2 + f3);
        }
        HEAT_PT(2);
        // ---- control warp: prepare the sample after the next while this one is processed (meets the others at [B])
        if (is_ctrl) {
            if (lane == 0) meta[n2].done = 0;
            cta_prepare<D, NPS>(P, &meta[n2], hid_s + n2 * CTA_HIS_MAX, lane, ctl, nchunks, no_cache);
        }
        HEAT_PT(3);
        row_bar<C::CTRL0>(is_ctrl);                                                   // [G1]
        // ---- every row slot: u <- gamma*u + (1-gamma)*f (:122); its dots in one butterfly
        //      (matrix_factorization.cpp:45-53, 82-85); E_k = exp(cos_k / T), published as per-slot sums
        float4 ua = f4zero();
        float uu = 0.f, rinv_u = 0.f, pp = 0.f, upd = 0.f;
        float un[NPS], nn[NPS], Ek[NPS], r3[NPS], rn3[NPS];
        if (is_row) {
            const float4 f0 = *reinterpret_cast<const float4 *>(fpart + 4 * sl), f1 = *reinterpret_cast<const float4 *>(fpart + D + 4 * sl);
            const float4 f2 = *reinterpret_cast<const float4 *>(fpart + 2 * D + 4 * sl), f3 = *reinterpret_cast<const float4 *>(fpart + 3 * D + 4 * sl);
            const float4 f = (f0 + f1) + (f2 + f3);
            ua = gamma * *reinterpret_cast<const float4 *>(urow_s + 4 * sl);
            fma4(ua, omg, f);
        }
#ifdef HEAT_CTA_PROFILE
        if (ua.x == 123.456f) ua.y = 0.f;                     // force ua to be materialised here
#endif
        HEAT_PT(9);
        if (is_neg) {
            float r[1 + 2 * NPS];
            r[0] = dot4(ua, ua);
#pragma unroll
            for (int n = 0; n < NPS; ++n) { r[1 + 2 * n] = dot4(ua, nw[n]); r[2 + 2 * n] = dot4(nw[n], nw[n]); }
            group_sum_n<LPR, 1 + 2 * NPS>(r, FULL);
            uu = r[0];
            rinv_u = rsqrt_eps(uu);
            float es = 0.f, bs = 0.f;
#pragma unroll
            for (int n = 0; n < NPS; ++n) {
                un[n] = r[1 + 2 * n]; nn[n] = r[2 + 2 * n];
                const float rinv_n = rsqrt_eps(nn[n]);
                Ek[n] = (slot + 16 * n < K) ? expf((un[n] * (rinv_u * rinv_n)) * score_mul) : 0.f;
                r3[n] = (rinv_u * rinv_u) * (rinv_u * rinv_n);             // 1 / (|u|^3 |n|)   (:136)
                rn3[n] = (rinv_u * rinv_n) * (rinv_n * rinv_n);            // 1 / (|u| |n|^3)   (:137)
                es += Ek[n];
                bs = fmaf(Ek[n], r3[n] * un[n], bs);
            }
            if (sl == 0) { sE[slot] = es; sB[slot] = bs; }
        } else if (is_pos) {
            float r[3] = {dot4(ua, ua), dot4(rw, rw), dot4(ua, rw)};
            group_sum_n<LPR, 3>(r, smask);
            uu = r[0]; pp = r[1]; upd = r[2];
            rinv_u = rsqrt_eps(uu);
            const float cos_p = upd * (rinv_u * rsqrt_eps(pp));
            if (sl == 0) { sc3[0] = pp; sc3[1] = upd; sc3[2] = expf(-(cos_p * score_mul)); }
            *reinterpret_cast<float4 *>(pw_s + 4 * sl) = rw;
        } else if (is_usr) {
            float r[1] = {dot4(ua, ua)};
#ifdef HEAT_CTA_PROFILE
            if (r[0] == 123.456f) r[0] = 0.f;
#endif
            HEAT_PT(10);
            group_sum_n<LPR, 1>(r, smask);
            uu = r[0];
#ifdef HEAT_CTA_PROFILE
            if (uu == 123.456f) uu = 0.f;
#endif
            HEAT_PT(11);
            rinv_u = rsqrt_eps(uu);
        }
        HEAT_PT(4);
        row_bar<C::CTRL0>(is_ctrl);                                                   // [A]
        // ---- softmax with the implicit zero logit (matrix_factorization.cpp:55-61, 91-109):
        //      esum = 1 + sum_k exp((cos_k - cos_p)/T) = 1 + Ep * sum_k E_k;  lg_k = (E_k Ep / esum) / T.
        //      No max shift: |cos/T| <= 14.3, far inside fp32 range.
        float cE = 0.f, SE = 0.f, SB = 0.f, esum = 1.f;
        if (is_row) {
            const float4 e0 = *reinterpret_cast<const float4 *>(sE), e1 = *reinterpret_cast<const float4 *>(sE + 4);
            const float4 e2 = *reinterpret_cast<const float4 *>(sE + 8), e3 = *reinterpret_cast<const float4 *>(sE + 12);
            const float4 e = (e0 + e1) + (e2 + e3);
            SE = (e.x + e.y) + (e.z + e.w);
            const float Ep = sc3[2];
            esum = fmaf(Ep, SE, 1.0f);
            cE = (Ep * score_mul) / esum;
        }
        HEAT_PT(5);
        // ---- negative rows: grads, clipped step, write-back (:127-150; sgd.cpp:14-26)
        //      lg*gn_n = c1*u + c2*N[k];  the N[k] part of lg*gu_n is c3*N[k] (:138-143)
        float c1[NPS], c2[NPS];
        if (is_neg) {
            float4 gpart = f4zero();
#pragma unroll
            for (int n = 0; n < NPS; ++n) {
                const float lg = cE * Ek[n];                   // :109
                c1[n] = (lg * rn3[n]) * nn[n];
                c2[n] = -(lg * rn3[n]) * un[n];
                fma4(gpart, (lg * r3[n]) * uu, nw[n]);
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
                            fma4(gk, c1[n], ua); fma4(gk, c2[n], nw[n]);
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
        // ---- positive row (:142, :169, :173-174); patch of the cached history sums; loss (:108)
        if (is_pos) {
            const float rinv_p = rsqrt_eps(pp);
            const float lgsum = cE * SE;
            const float q = lgsum * ((rinv_u * rinv_p) * (rinv_p * rinv_p));   // lgsum / (|u| |p|^3)
            float4 gp = rg;
            fma4(gp, -q * pp, ua); fma4(gp, q * upd, rw);          // lgsum * gp_p, gp_p = -(pp*u - up*p) / (|u||p|^3)
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
            *reinterpret_cast<float4 *>(mean_next + 4 * sl) = rH * a4;
            if (sl == 0) {
                const float loss = logf(esum);
                loss_f += loss;
                if ((iter & 31) == 31) { loss_acc += (double)loss_f; loss_f = 0.f; }
                if (P.loss_out) P.loss_out[m->s - P.begin] = loss;
            }
        }
        // ---- W0 outer product of the PREVIOUS sample (its 0.6*g_u was published before [G1])
        if (have_tail && !flush_due && is_neg) w0_tail(mean_prev, false);
        HEAT_PT(6);
        __syncthreads();                                                              // [B]
        // ---- user row: grad, aggregator backward, step (:141, :152, :166, :171-172); overlaps the next sample's loads
        if (is_usr) {
            const float4 b0 = *reinterpret_cast<const float4 *>(sB), b1 = *reinterpret_cast<const float4 *>(sB + 4);
            const float4 b2 = *reinterpret_cast<const float4 *>(sB + 8), b3 = *reinterpret_cast<const float4 *>(sB + 12);
            const float4 b = (b0 + b1) + (b2 + b3);
            SB = (b.x + b.y) + (b.z + b.w);
            pp = sc3[0]; upd = sc3[1];
            float4 gacc = f4zero();
#pragma unroll
            for (int q = 0; q < 16; ++q) gacc = gacc + *reinterpret_cast<const float4 *>(part + q * D + 4 * sl);
            const float4 pw = *reinterpret_cast<const float4 *>(pw_s + 4 * sl);
            const float lgsum = cE * SE;
            const float q3 = lgsum * ((rinv_u * rinv_u) * (rinv_u * rsqrt_eps(pp)));   // lgsum / (|u|^3 |p|)
            // sum_k lg_k*(-un_k/(|u|^3|n_k|))*u  - lgsum*gu_p,  gu_p = (uu*p - up*u)/(|u|^3|p|)  (:60-62)
            fma4(gacc, fmaf(q3, upd, -(cE * SB)), ua);
            fma4(gacc, -q3 * uu, pw);
            float4 gu = rg + gacc;
            *reinterpret_cast<float4 *>(gu06_s + 4 * sl) = omg * gu;      // behavior_aggregators.cpp:131-139
            gu = gamma * gu;                                               // :148-152
            gu = clamp4(gu, P.clip);
            fma4(ua, -P.lr, gu);
            stcg4(P.user_w + (size_t)u * D + 4 * sl, ua);
            stcg4(P.user_g + (size_t)u * D + 4 * sl, gu);
            u_keep = ua; gu_keep = gu; keep_u = u;
        }
        HEAT_PT(7);
#ifdef HEAT_CTA_PROFILE
        if (prof_on && iter == 1000)
            printf("abs tid %d iter %u fresh %d dup %d: top %lld issue %lld gemv %lld prep %lld dots %lld softmax %lld update %lld user %lld pregemv %lld ua %lld dot %lld bfly %lld\n", tid, iter, (int)fresh, (int)dup,
                   prof_abs[0] % 1000000, prof_abs[1] % 1000000, prof_abs[2] % 1000000, prof_abs[3] % 1000000, prof_abs[4] % 1000000, prof_abs[5] % 1000000,
                   prof_abs[6] % 1000000, prof_abs[7] % 1000000, prof_abs[8] % 1000000, prof_abs[9] % 1000000, prof_abs[10] % 1000000, prof_abs[11] % 1000000);
#endif
        iter += 1;
        have_tail = true;
        cur = nx;
        mb = (mb == 2) ? 0 : mb + 1;
    }
    // ---- W0 gradient of the last sample and the leftover (< 32 samples) accumulators
    __syncthreads();
    if (have_tail && is_neg) w0_tail(meanb + (mb == 0 ? 2 : mb - 1) * D, ((iter & 31) == 0) || !P.drop_w0_tail);
    if (is_pos && sl == 0) {
        loss_acc += (double)loss_f;
        if (loss_acc != 0.0) atomicAdd(P.loss_sum, loss_acc);
    }
}

}  // namespace heat
