// heat_b200/csrc/train_kernel_cta.cuh -- latency-oriented training kernel: one CTA per sample.
//
// Why: the reference's update is a clipped RUNNING SUM applied at every touch, so its dynamics depend
// on how many samples are in flight: the loss / Recall of the reference at 1-128
// threads is reproduced only while a few hundred samples at most are being processed concurrently.
// Throughput at a FIXED number of samples in flight is (samples in flight) / (latency of one sample),
// so this kernel spends a whole CTA on one sample and keeps only true dependencies on the sample's
// critical path.  Warp roles (every warp is role-uniform):
//
//   neg warps   16 slots of D/4 lanes; slot q owns the negative rows k = q, q+16, ... (NPS per slot).
//               The same 4D threads apply the batched W0 step every 32 samples from shared-memory rings of
//               (mean, 0.6 g_u), D/4 rows of one column each, and leave the stepped W0 in shared memory.
//   gemv warp   D/32 columns of W0 per lane in REGISTERS: f = mean . W0 of the next sample while the other
//               warps are still storing rows (it used to sit on the neg warps, the last to reach [G1]); also
//               the loss (log esum and its running sums).  It runs its own copy of the loop skeleton (same
//               barriers, no row work), so that its D^2/32 registers are not live in the row warps' code;
//               one warp keeps the CTA at 12 warps = 3 per scheduler (168 registers a thread).
//   pos warp    the positive row, the patch of the cached history column sums -> next history mean.
//   usr warp    the user row, kept in registers across a user's run.
//   ctl warp    runs ahead on its own loop: one LANE per sample, 32 samples per batch -- click, Philox
//               negatives, duplicate / conflict detection, history membership -- into a ring of three
//               batches; it meets the row warps once per batch (named barrier 2).
//
// Per sample the row warps meet at two full barriers; between them sits only what the next stage needs:
//   [G1] all: issue the cp.async prefetch of the NEXT sample's rows; u <- 0.4u + 0.6f; the slot's dots in one
//        interleaved butterfly; E_k = exp(cos_k/T) and the UNNORMALISED partial user grads, published per slot
//   [A]  all: esum = 1 + exp(-cos_p/T) * sum_k E_k from 16 shared partials (no cross-slot shuffles);
//        usr: user grad, aggregator backward, clipped step, next sample's input row     (the critical chain)
//        pos: clipped step, patched column sums, next history mean -> named barrier 3 -> gemv: mean.W0 of the next sample
//        neg: clipped row steps, stores
//   Rows come from the prefetch slots; a sample sharing an item row with its predecessor ("late") or starting
//   a run ("fresh": history gather) loads after an extra barrier, behind the predecessor's stores.
// Same arithmetic as train_kernel.cuh (reference: models/matrix_factorization.cpp:15-181,
// behavior_aggregators/behavior_aggregators.cpp:51-153, optimizers/sgd.cpp:14-26) up to fp32 rounding:
// 1/sqrt as MUFU.RSQ + one Newton step, softmax as exp(a)*exp(-b) instead of exp(a-b-max).  With one CTA
// the samples are applied strictly in file order (deterministic mode).
#pragma once
#include "train_kernel.cuh"

namespace heat {

template <int NPS> struct CtaMeta {      // one sample, prepared by the control warp
    int u, p, H, pmult;
    int flags, pad;                      // CTA_FRESH | CTA_DUP | CTA_LATE | CTA_REGATHER
    float rH, pad2;
    long long s;
    int negid[16 * NPS];
};
enum { CTA_FRESH = 1, CTA_DUP = 2, CTA_LATE = 4, CTA_REGATHER = 8 };

template <int NPS> struct CtaBatch {
    int cnt, pad[3];                     // samples in this batch; 0 = no more work
    CtaMeta<NPS> meta[32];
};

template <int D, int NPS> struct CtaCfg {
    static_assert(D == 32 || D == 64, "the CTA engine covers emb_dim 32 and 64");
    static constexpr int LPR = D / 4;                       // lanes per row
    static constexpr int NEGT = 16 * LPR;                    // threads of the negative slots (whole warps)
    static constexpr int POS0 = NEGT, USR0 = POS0 + 32, GMV0 = USR0 + 32, CTRL0 = GMV0 + 32;
    static constexpr int NLD = GMV0;                         // threads that load rows (negative slots, pos, usr)
    static constexpr int NT = CTRL0 + 32;
    static constexpr int NSLOT = 18;
    static constexpr int KMAX = 16 * NPS;
    static constexpr int IQ = D / 4;                         // W0 rows per neg thread (the batched W0 step)
    static constexpr int MRING = 64, GRING = 32;             // history means of the last 64 samples, 0.6*g_u of the last 32
    // shared memory, in floats
    static constexpr int OFF_MEAN = 0;                       // [MRING][D]
    static constexpr int OFF_GU06 = OFF_MEAN + MRING * D;    // [GRING][D] 0.6 * user grad (W0 outer products)
    static constexpr int OFF_ACC = OFF_GU06 + GRING * D;     // [D] history column sums
    static constexpr int OFF_F = OFF_ACC + D;                // [D] f = mean . W0
    static constexpr int OFF_UROW = OFF_F + D;               // [D] the user's stored row (input of the aggregation)
    static constexpr int OFF_PW = OFF_UROW + D;              // [D] positive row snapshot (user step)
    static constexpr int OFF_PART = OFF_PW + D;              // [16][D] unnormalised user-grad partial sums of the negative slots
    static constexpr int OFF_HPART = OFF_PART + 16 * D;      // [NSLOT][D] history-gather partial sums
    static constexpr int OFF_SC = OFF_HPART + NSLOT * D;     // sE[16], sB[16], pp, up, Ep
    static constexpr int OFF_W0S = OFF_SC + 40;              // [D][D] W0 as of this CTA's last step (neg warps -> gemv warp)
    static constexpr int OFF_RING = OFF_W0S + D * D;         // 3 x CtaBatch
    static constexpr int OFF_STAGE = OFF_RING + 3 * (int)(sizeof(CtaBatch<NPS>) / 4);   // [2*NPS][NLD] float4: prefetched rows
    static constexpr int FLOATS = OFF_STAGE + 2 * NPS * NLD * 4 + 8;
};

// named barriers: 1 = the row warps, 2 = row warps + control warp (once per batch),
// 3 = pos warp (arrives: the next history mean is published) + gemv warp (waits for it)
template <int N> __device__ __forceinline__ void row_bar() { asm volatile("bar.sync 1, %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void batch_bar() { asm volatile("bar.sync 2, %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void mean_arrive() { asm volatile("bar.arrive 3, %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void mean_wait() { asm volatile("bar.sync 3, %0;" ::"n"(N) : "memory"); }

// N independent sums over the VL lanes of a row group in ONE butterfly: the N shuffles of a step are
// issued back to back, so the latency of a step is paid once, not N times.
template <int VL, int N> __device__ __forceinline__ void group_sum_n(float (&v)[N])
{
#pragma unroll
    for (int o = VL / 2; o > 0; o >>= 1) {
        float t[N];
#pragma unroll
        for (int i = 0; i < N; ++i) t[i] = __shfl_xor_sync(FULL, v[i], o);
#pragma unroll
        for (int i = 0; i < N; ++i) v[i] += t[i];
    }
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

// Control-warp state carried across batches (warp-uniform).
template <int NPS> struct CtaCtrl {
    long long s = 0, s_end = 0, chunk_begin = 0;
    int carry_u = -1, carry_regather = 0;
    const CtaMeta<NPS> *prev_last = nullptr;                 // last sample of the previous batch (still in the ring)
};

// Control warp: prepare up to 32 consecutive samples of this CTA's current chunk, one LANE per sample
// (fetches the next chunk when the current one is used up; cnt = 0 when there is none).
template <int D, int NPS, bool TILE>
__device__ __forceinline__ void cta_prepare_batch(const TrainParams &P, CtaBatch<NPS> *B, int lane, CtaCtrl<NPS> &c,
                                                  bool no_cache)
{
    constexpr int KM = 16 * NPS;
    if (c.s >= c.s_end) {                                  // next chunk (warp-uniform)
        unsigned long long ch = 0;
        if (lane == 0) ch = atomicAdd(P.chunk_counter, 1ULL);
        ch = __shfl_sync(FULL, ch, 0);
        if (!chunk_range(P, (long long)ch, c.s, c.s_end)) {
            c.s = c.s_end = 0;
            if (lane == 0) B->cnt = 0;
            return;
        }
        c.chunk_begin = c.s;
        c.carry_u = -1;
    }
    const int cnt = (int)min((long long)32, c.s_end - c.s);
    const bool act = lane < cnt;
    const long long sj = c.s + lane;
    const int K = P.K;
    const int2 click = act ? __ldg(&P.clicks[sj]) : make_int2(0, -7000 - lane);
    const int u = click.x, p = click.y;
    const int H = act ? __ldg(P.masks + u) : 0;
    CtaMeta<NPS> *my = &B->meta[lane];

    // ---- negatives: one Philox block gives four slots (philox.cuh); the rare "draw == positive" case
    //      takes the stateless look-back path
    int neg[KM];
    const unsigned long long gs = P.sample_offset + (unsigned long long)sj;
#pragma unroll
    for (int q = 0; q < KM / 4; ++q) {
        Philox4 r;
        if (act && 4 * q < K)
            r = philox4x32_10((uint32_t)gs, (uint32_t)(gs >> 32), P.epoch, (uint32_t)q, (uint32_t)P.seed, (uint32_t)(P.seed >> 32));
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int k = 4 * q + t;
            int d = -1 - k;                                // unused slot: a unique negative sentinel
            if (act && k < K) {
                d = (int)map_draw(r.v[t], P.seed, P.epoch, gs, (uint32_t)P.num_items, TILE ? P.tile_size : 0u, P.refresh_interval);
                if (d == p) d = sample_negative(P, sj, k, p);
            }
            neg[k] = d;
            if (act) my->negid[k] = d;
        }
    }
    __syncwarp();
    // ---- duplicates among the negatives or a negative equal to the positive (such a sample replays its row
    //      steps in the reference's sequential k order); item rows shared with the previous sample (which
    //      rewrites them: no prefetch).  The other side of each comparison is read from the ring.
    int self = 0;
    bool conf = false, dup = false;
    const CtaMeta<NPS> *prev = lane > 0 ? &B->meta[lane - 1] : c.prev_last;
    const bool has_prev = act && prev != nullptr;
    int prev_p = __shfl_up_sync(FULL, p, 1);               // lane-1's positive (its meta fields are written at the end)
    if (lane == 0) prev_p = c.prev_last ? c.prev_last->p : -8000;
    if (!has_prev) prev_p = -8000;
#pragma unroll 4
    for (int b = 0; b < K; ++b) {
        const int vm = act ? my->negid[b] : -8100, vp = has_prev ? prev->negid[b] : -8200;
#pragma unroll
        for (int a = 0; a < KM; ++a) { self += neg[a] == vm; conf |= neg[a] == vp; }
        conf |= p == vp;
    }
#pragma unroll
    for (int a = 0; a < KM; ++a) { dup |= neg[a] == p; conf |= neg[a] == prev_p; }
    dup |= self != K;                                      // every valid id matches itself exactly once
    conf |= p == prev_p;
    // ---- history membership of the rows this sample rewrites (see train_kernel.cuh, CACHE): the positive's
    //      multiplicity patches the cached column sums, a negative in the history forces a re-gather
    int pmult = 0;
    bool nhit = false;
    const int *hrow = P.his + (size_t)(act ? u : 0) * P.max_his;
    const int Hmax = __reduce_max_sync(FULL, H);
#pragma unroll 4
    for (int h = 0; h < Hmax; ++h) {
        const int id = h < H ? __ldg(hrow + h) : -9000;
        pmult += id == p;
#pragma unroll
        for (int a = 0; a < KM; ++a) nhit |= neg[a] == id;
    }
    const bool regather = nhit || (pmult != 0 && dup);
    // ---- start of a run: first sample of a chunk, a new user, after a re-gather, every 64 samples of a chunk
    int pu = __shfl_up_sync(FULL, u, 1), pr = __shfl_up_sync(FULL, (int)regather, 1);
    if (lane == 0) { pu = c.carry_u; pr = c.carry_regather; }
    const bool fresh = no_cache || u != pu || pr != 0 || ((sj - c.chunk_begin) & 63) == 0;
    if (act) {
        my->u = u; my->p = p; my->H = H; my->pmult = pmult;
        my->flags = (fresh ? CTA_FRESH : 0) | (dup ? CTA_DUP : 0) | (conf ? CTA_LATE : 0) | (regather ? CTA_REGATHER : 0);
        my->rH = (float)(1.0 / (double)H);                  // behavior_aggregators.cpp:63
        my->s = sj;
    }
    if (lane == 0) B->cnt = cnt;
    c.carry_u = __shfl_sync(FULL, u, cnt - 1);
    c.carry_regather = __shfl_sync(FULL, (int)regather, cnt - 1);
    c.prev_last = &B->meta[cnt - 1];
    c.s += cnt;
}


// Phase timing for kernel development (-DHEAT_CTA_PROFILE through HEAT_NVCC_DEFS, build.py): lane 0 of every
// row warp of block 0 prints clock64() at the marked points of one iteration.  (A clock read placed right
// after a barrier does not include the wait; compare absolute times across warps.)
#ifdef HEAT_CTA_PROFILE
#define HEAT_PT(i) do { long long t_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_) :: "memory"); if (prof_on) prof_abs[i] = t_; } while (0)
#else
#define HEAT_PT(i) do { } while (0)
#endif

// TILE: the hot-tile sampler (philox.cuh) is compiled in; the default instantiation carries none of its code
template <int D, int NPS, bool TILE = false>
__global__ void __launch_bounds__(CtaCfg<D, NPS>::NT, 1) train_cta_kernel(const TrainParams P)
{
    using C = CtaCfg<D, NPS>;
    constexpr int LPR = C::LPR, NSLOT = C::NSLOT, IQ = C::IQ, NROW = C::CTRL0, NT = C::NT;
    constexpr int MRING = C::MRING, GRING = C::GRING;
    extern __shared__ __align__(16) float smem[];
    float *mean_ring = smem + C::OFF_MEAN, *gu06_ring = smem + C::OFF_GU06, *accs = smem + C::OFF_ACC, *f_s = smem + C::OFF_F;
    float *urow_s = smem + C::OFF_UROW, *pw_s = smem + C::OFF_PW;
    float *part = smem + C::OFF_PART, *hpart = smem + C::OFF_HPART;
    float *sE = smem + C::OFF_SC, *sB = sE + 16, *sc3 = sB + 16;      // sc3: pp, up, Ep, 1/|p|
    float *w0s = smem + C::OFF_W0S;
    CtaBatch<NPS> *ring = reinterpret_cast<CtaBatch<NPS> *>(smem + C::OFF_RING);

    const int tid = threadIdx.x, lane = tid & 31;
    const int role = tid < C::NEGT ? 0 : tid < C::USR0 ? 1 : tid < C::GMV0 ? 2 : tid < C::CTRL0 ? 4 : 3;   // warp-uniform
    const bool no_cache = P.no_cache != 0;

    // ======== control warp: its own loop, two batches ahead of the row warps ========
    if (role == 3) {
        CtaCtrl<NPS> ctl;
        auto prepare = [&](int slot_, bool more) -> bool {
            if (!more) {
                if (lane == 0) ring[slot_].cnt = 0;
                return false;
            }
            cta_prepare_batch<D, NPS, TILE>(P, &ring[slot_], lane, ctl, no_cache);
            __syncwarp();
            return ring[slot_].cnt != 0;
        };
        const bool c0 = prepare(0, true);
        bool cprev = prepare(1, c0);
        batch_bar<NT>();                                     // #1: batches 0 and 1 are visible
        if (!c0) return;
        for (int t = 2;; ++t) {
            const bool ccur = prepare(t % 3, cprev);
            batch_bar<NT>();                                 // #t: batch t is visible; the rows finished batch t-2
            if (!cprev) return;
            cprev = ccur;
        }
    }

    // ======== gemv warp: the row warps' loop skeleton (same barriers, same block-uniform conditions), no row work ========
    if (role == 4) {
        constexpr int CPL = D / 32;                          // adjacent columns of W0 per lane
        float wcol[D][CPL];
#pragma unroll
        for (int k = 0; k < D; ++k)
#pragma unroll
            for (int c = 0; c < CPL; ++c) wcol[k][c] = __ldcg(P.w0 + k * D + CPL * lane + c);
        // f = mean . W0 (behavior_aggregators.cpp:118): per column four row quarters of four interleaved chains each
        auto gemv = [&](const float *mean_) {
            float fr[4][CPL];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float f[4][CPL];
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int c = 0; c < CPL; ++c) f[j][c] = 0.f;
#pragma unroll
                for (int i4 = 0; i4 < IQ; i4 += 4) {
                    const float4 m4 = *reinterpret_cast<const float4 *>(mean_ + q * IQ + i4);
                    const float mj[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j)
#pragma unroll
                        for (int c = 0; c < CPL; ++c) f[j][c] = fmaf(mj[j], wcol[q * IQ + i4 + j][c], f[j][c]);
                }
#pragma unroll
                for (int c = 0; c < CPL; ++c) fr[q][c] = (f[0][c] + f[1][c]) + (f[2][c] + f[3][c]);
            }
#pragma unroll
            for (int c = 0; c < CPL; ++c) f_s[CPL * lane + c] = (fr[0][c] + fr[1][c]) + (fr[2][c] + fr[3][c]);
        };
        bool have_f = false;
        unsigned it = 0;
        double loss_acc = 0.0;
        float loss_f = 0.f;
        batch_bar<NT>();                                     // #1
        for (int rbuf = 0;; rbuf = rbuf == 2 ? 0 : rbuf + 1) {
            const CtaBatch<NPS> *Bt = &ring[rbuf], *Bn = &ring[rbuf == 2 ? 0 : rbuf + 1];
            const int cnt = Bt->cnt;
            if (cnt == 0) break;
            for (int j = 0; j < cnt; ++j) {
                const CtaMeta<NPS> *mn = (j + 1 < cnt) ? &Bt->meta[j + 1] : (Bn->cnt > 0 ? &Bn->meta[0] : nullptr);
                const int fl = Bt->meta[j].flags, nfl = mn ? mn->flags : 0;
                if (fl & CTA_LATE) row_bar<NROW>();
                if (fl & CTA_FRESH) { row_bar<NROW>(); row_bar<NROW>(); row_bar<NROW>(); }   // the history gather
                if (it > 0 && (it & 31) == 0) {               // a W0 step: the neg warps leave the stepped W0 in w0s
                    row_bar<NROW>();
                    row_bar<NROW>();
#pragma unroll
                    for (int k = 0; k < D; ++k) {
                        if constexpr (CPL == 2) {
                            const float2 w2 = *reinterpret_cast<const float2 *>(w0s + k * D + 2 * lane);
                            wcol[k][0] = w2.x; wcol[k][1] = w2.y;
                        } else {
                            wcol[k][0] = w0s[k * D + lane];
                        }
                    }
                }
                if (!have_f) gemv(mean_ring + (it & (MRING - 1)) * D);
                row_bar<NROW>();                              // [G1]
                row_bar<NROW>();                              // [A]: f_s has been consumed
                have_f = mn != nullptr && !(nfl & CTA_FRESH) && ((it + 1) & 31) != 0;
                {
                    // loss = log(esum) (matrix_factorization.cpp:108), esum exactly as the row warps form it below
                    const float4 e0 = *reinterpret_cast<const float4 *>(sE), e1 = *reinterpret_cast<const float4 *>(sE + 4);
                    const float4 e2 = *reinterpret_cast<const float4 *>(sE + 8), e3 = *reinterpret_cast<const float4 *>(sE + 12);
                    const float4 e = (e0 + e1) + (e2 + e3);
                    const float esum = fmaf(sc3[2], (e.x + e.y) + (e.z + e.w), 1.0f);
                    if (lane == 0) {
                        const float loss = logf(esum);
                        loss_f += loss;
                        if ((it & 31) == 31) { loss_acc += (double)loss_f; loss_f = 0.f; }
                        if (P.loss_out) P.loss_out[Bt->meta[j].s - P.begin] = loss;
                    }
                }
                mean_wait<64>();                              // the pos warp has published the next history mean
                if (have_f) gemv(mean_ring + ((it + 1) & (MRING - 1)) * D);   // off everyone's critical path
                if (fl & CTA_DUP)
                    for (int k = 0; k < P.K; ++k) row_bar<NROW>();
                it += 1;
            }
            batch_bar<NT>();
        }
        row_bar<NROW>();                                     // the leftover W0 step
        if (lane == 0) {
            loss_acc += (double)loss_f;
            if (loss_acc != 0.0) atomicAdd(P.loss_sum, loss_acc);
        }
        return;
    }

    // ======== row warps ========
    const int sl = tid % LPR;                                // lane within the row slot (the pos / usr warps mirror their slot)
    const int slot = role == 0 ? tid / LPR : 15 + role;      // 0..15 negatives, 16 positive, 17 user
    float *stage = smem + C::OFF_STAGE + 4 * tid;            // this thread's prefetch slots: [2n] weight, [2n+1] grad slice
    constexpr int STG = C::NLD * 4;                          // floats between two slots

    const float gamma = 0.4f, omg = 1.0f - gamma;
    const float score_mul = (float)(1.0 / 0.07);
    const int K = P.K;

    // neg threads: rows rb*IQ .. +IQ of column col of W0 in the batched step; 4 lanes share a column
    const int col = tid >> 2, rb = tid & 3;
    // neg threads: the W0 step of the last `ns` samples (behavior_aggregators.cpp:131-146) as one batched
    // rank-ns update from the rings: W0 -= lr0 * (sum_s mean_s^T (0.6 g_s)) / 32; the stepped slice goes to w0s
    auto w0_step = [&](unsigned upto, int ns) {
        float acc[IQ];
#pragma unroll
        for (int i = 0; i < IQ; ++i) acc[i] = 0.f;
        for (int q = ns; q > 0; --q) {
            const unsigned sidx = upto - (unsigned)q;
            const float *pm = mean_ring + (sidx & (MRING - 1)) * D + rb * IQ;
            const float gq = gu06_ring[(sidx & (GRING - 1)) * D + col];
#pragma unroll
            for (int i4 = 0; i4 < IQ; i4 += 4) {
                const float4 m4 = *reinterpret_cast<const float4 *>(pm + i4);
                acc[i4 + 0] = fmaf(m4.x, gq, acc[i4 + 0]); acc[i4 + 1] = fmaf(m4.y, gq, acc[i4 + 1]);
                acc[i4 + 2] = fmaf(m4.z, gq, acc[i4 + 2]); acc[i4 + 3] = fmaf(m4.w, gq, acc[i4 + 3]);
            }
        }
#pragma unroll
        for (int i = 0; i < IQ; ++i) {
            const float delta = -(P.lr_w0 * (acc[i] / 32.0f));
            const float old = atomicAdd(P.w0 + (rb * IQ + i) * D + col, delta);   // returns the value before: picks up other CTAs' steps
            w0s[(rb * IQ + i) * D + col] = old + delta;
        }
    };

    float4 u_keep = f4zero(), gu_keep = f4zero();            // usr: the row / grad row of user keep_u after its last step
    int keep_u = -1, pf_u = -1;                              // usr: user of u_keep; user whose row is in the prefetch slot
    bool urow_done = false;                                  // usr: urow_s already holds this sample's user row
    bool have_pf = false;                                    // the prefetch slots hold this sample's item rows
    int nidn[NPS];                                           // rows this thread prefetches for the next sample
    // the prefetch after [G1] is the same two LDGSTS per row for every role: only table and row id differ
    const float *pf_w = (role == 2 ? P.user_w : P.item_w) + 4 * sl, *pf_g = (role == 2 ? P.user_g : P.item_g) + 4 * sl;
    unsigned pf_rows = 1u;                                   // bit n: prefetch slot n is in use
    if (role == 0) {
        pf_rows = 0u;
#pragma unroll
        for (int n = 0; n < NPS; ++n) pf_rows |= (slot + 16 * n < K ? 1u : 0u) << n;
    }
    unsigned it = 0;                                         // samples processed by this CTA
#ifdef HEAT_CTA_PROFILE
    const bool prof_on = blockIdx.x == 0 && lane == 0;
    long long prof_abs[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
#endif
    batch_bar<NT>();                                         // #1

    for (int rbuf = 0;; rbuf = rbuf == 2 ? 0 : rbuf + 1) {
        const CtaBatch<NPS> *Bt = &ring[rbuf], *Bn = &ring[rbuf == 2 ? 0 : rbuf + 1];
        const int cnt = Bt->cnt;
        if (cnt == 0) break;
        for (int j = 0; j < cnt; ++j) {
            const CtaMeta<NPS> *m = &Bt->meta[j];
            // the next sample's record: always an address inside the ring, valid only if has_next (no 64-bit selects)
            const bool has_next = j + 1 < cnt || Bn->cnt > 0;
            const CtaMeta<NPS> *mn = (j + 1 < cnt) ? m + 1 : &Bn->meta[0];
            const int u = m->u, p = m->p, H = m->H, fl = m->flags;
            const bool fresh = (fl & CTA_FRESH) != 0, dup = (fl & CTA_DUP) != 0;
            const int nfl = has_next ? mn->flags : 0, nu = has_next ? mn->u : -1;
            const bool pf = has_next && !(nfl & CTA_LATE);                       // block-uniform
            float *mean = mean_ring + (it & (MRING - 1)) * D;
            float *mean_next = mean_ring + ((it + 1) & (MRING - 1)) * D;
            HEAT_PT(0);

            // ---- the rows of this sample: prefetched during the previous sample, or loaded now (first sample,
            //      or a sample sharing an item row with its predecessor: after that one's stores)
            if (fl & CTA_LATE) row_bar<NROW>();
            int nid[NPS];
            bool pf_me = pf;                                   // this warp prefetches rows of the next sample (warp-uniform)
            float4 nw[NPS], gn[NPS];
            float4 rw = f4zero(), rg = f4zero();              // positive / user weight and grad slices
            if (role == 0) {
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
                if (pf) {                                      // ids of the rows prefetched after [G1]: read them off its path
#pragma unroll
                    for (int n = 0; n < NPS; ++n) nidn[n] = slot + 16 * n < K ? mn->negid[slot + 16 * n] : 0;
                }
            } else if (role == 1) {
                if (pf) nidn[0] = mn->p;
                if (have_pf) {
                    cp_async_wait0();
                    rw = *reinterpret_cast<const float4 *>(stage); rg = *reinterpret_cast<const float4 *>(stage + STG);
                } else {
                    rw = ldcg4(P.item_w + (size_t)p * D + 4 * sl);
                    rg = ldcg4(P.item_g + (size_t)p * D + 4 * sl);
                }
            } else if (role == 2) {
                if (u == keep_u) {                             // the run continues: the row is in registers
                    rw = u_keep; rg = gu_keep;
                } else if (pf_u == u) {
                    cp_async_wait0();
                    rw = *reinterpret_cast<const float4 *>(stage); rg = *reinterpret_cast<const float4 *>(stage + STG);
                } else {
                    rw = ldcg4(P.user_w + (size_t)u * D + 4 * sl);
                    rg = ldcg4(P.user_g + (size_t)u * D + 4 * sl);
                }
                if (!urow_done) *reinterpret_cast<float4 *>(urow_s + 4 * sl) = rw;
                pf_me = pf && nu != u;                         // the next sample starts another user's run
                pf_u = pf_me ? nu : -1;
                nidn[0] = nu;
            }
            HEAT_PT(1);
            // ---- first sample of a run: gather the history rows (behavior_aggregators.cpp:96-105)
            if (fresh) {
                row_bar<NROW>();                              // the previous sample's row stores are done
                const int *hrow = P.his + (size_t)u * P.max_his;
                float4 a4 = f4zero();
                for (int h0 = slot; h0 < H; h0 += 8 * NSLOT) {   // 8 ids, then 8 rows in flight per thread
                    int hid[8];
                    float4 hr[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q) hid[q] = (h0 + q * NSLOT < H) ? __ldg(hrow + h0 + q * NSLOT) : -1;
#pragma unroll
                    for (int q = 0; q < 8; ++q) hr[q] = hid[q] >= 0 ? ldcg4(P.item_w + (size_t)hid[q] * D + 4 * sl) : f4zero();
#pragma unroll
                    for (int q = 0; q < 8; ++q) a4 = a4 + hr[q];
                }
                *reinterpret_cast<float4 *>(hpart + slot * D + 4 * sl) = a4;
                row_bar<NROW>();
                if (tid < D) {
                    float a = 0.f;
#pragma unroll
                    for (int q = 0; q < NSLOT; ++q) a += hpart[q * D + tid];
                    accs[tid] = a;
                    mean[tid] = a * m->rH;
                }
                row_bar<NROW>();
            }
            // ---- a W0 step is due (every 32 samples): it must land before this sample's mean.W0
            if (it > 0 && (it & 31) == 0) {
                row_bar<NROW>();                              // the rings hold the last 32 samples
                if (role == 0) w0_step(it, 32);
                row_bar<NROW>();                              // the stepped W0 is in w0s
            }
            // (gemv warp: f = mean . W0 now, unless it was computed during the previous sample)
            HEAT_PT(2);
            row_bar<NROW>();                                                          // [G1]
            // ---- prefetch of the next sample's rows (ordered after every store of the previous sample by [G1];
            //      a row shared with THIS sample makes the next one "late" instead)
            if (pf_me) {
#pragma unroll
                for (int n = 0; n < NPS; ++n) {
                    if (pf_rows >> n & 1u) {
                        cp_async16(stage + (2 * n) * STG, pf_w + (size_t)nidn[n] * D);
                        cp_async16(stage + (2 * n + 1) * STG, pf_g + (size_t)nidn[n] * D);
                    }
                }
                cp_async_commit();
            }
            have_pf = pf;
            HEAT_PT(9);
            // ---- u <- gamma*u + (1-gamma)*f (:122); the slot's dots in one butterfly
            //      (matrix_factorization.cpp:45-53, 82-85); E_k = exp(cos_k / T), published as per-slot sums
            float4 ua = gamma * *reinterpret_cast<const float4 *>(urow_s + 4 * sl);
            fma4(ua, omg, *reinterpret_cast<const float4 *>(f_s + 4 * sl));
            float uu, rinv_u, pp = 0.f, upd = 0.f;
            float un[NPS], nn[NPS], Ek[NPS], rn3[NPS];
            if (role == 0) {
                float r[1 + 2 * NPS];
                r[0] = dot4(ua, ua);
#pragma unroll
                for (int n = 0; n < NPS; ++n) { r[1 + 2 * n] = dot4(ua, nw[n]); r[2 + 2 * n] = dot4(nw[n], nw[n]); }
                group_sum_n<LPR, 1 + 2 * NPS>(r);
                uu = r[0];
                HEAT_PT(10);
                rinv_u = rsqrt_eps(uu);
                float es = 0.f, bs = 0.f;
                float4 gpart = f4zero();
#pragma unroll
                for (int n = 0; n < NPS; ++n) {
                    un[n] = r[1 + 2 * n]; nn[n] = r[2 + 2 * n];
                    const float rinv_n = rsqrt_eps(nn[n]);
                    Ek[n] = (slot + 16 * n < K) ? expf((un[n] * (rinv_u * rinv_n)) * score_mul) : 0.f;
                    const float r3 = (rinv_u * rinv_u) * (rinv_u * rinv_n);    // 1 / (|u|^3 |n|)   (:136)
                    rn3[n] = (rinv_u * rinv_n) * (rinv_n * rinv_n);            // 1 / (|u| |n|^3)   (:137)
                    es += Ek[n];
                    bs = fmaf(Ek[n], r3 * un[n], bs);
                    fma4(gpart, (Ek[n] * r3) * uu, nw[n]);     // the N[k] part of gu_n (:138), still without lg's common factor
                }
                if (sl == 0) { sE[slot] = es; sB[slot] = bs; }
                *reinterpret_cast<float4 *>(part + slot * D + 4 * sl) = gpart;
            } else if (role == 1) {
                float r[3] = {dot4(ua, ua), dot4(rw, rw), dot4(ua, rw)};
                group_sum_n<LPR, 3>(r);
                uu = r[0]; pp = r[1]; upd = r[2];
                rinv_u = rsqrt_eps(uu);
                const float rinv_p = rsqrt_eps(pp);
                const float cos_p = upd * (rinv_u * rinv_p);
                if (sl == 0) { sc3[0] = pp; sc3[1] = upd; sc3[2] = expf(-(cos_p * score_mul)); sc3[3] = rinv_p; }
                *reinterpret_cast<float4 *>(pw_s + 4 * sl) = rw;
            } else if (role == 2) {
                float r[1] = {dot4(ua, ua)};
                group_sum_n<LPR, 1>(r);
                uu = r[0];
                rinv_u = rsqrt_eps(uu);
            }
            HEAT_PT(3);
            row_bar<NROW>();                                                          // [A]
            // ---- softmax with the implicit zero logit (matrix_factorization.cpp:55-61, 91-109):
            //      esum = 1 + sum_k exp((cos_k - cos_p)/T) = 1 + Ep * sum_k E_k;  lg_k = cE * E_k, cE = (Ep / esum) / T.
            //      No max shift: |cos/T| <= 14.3, far inside fp32 range.
            float SE;
            {
                const float4 e0 = *reinterpret_cast<const float4 *>(sE), e1 = *reinterpret_cast<const float4 *>(sE + 4);
                const float4 e2 = *reinterpret_cast<const float4 *>(sE + 8), e3 = *reinterpret_cast<const float4 *>(sE + 12);
                const float4 e = (e0 + e1) + (e2 + e3);
                SE = (e.x + e.y) + (e.z + e.w);
            }
            const float Ep = sc3[2];
            const float esum = fmaf(Ep, SE, 1.0f);
            const float cE = (Ep * score_mul) / esum;
            float c1[NPS], c2[NPS];
            float4 gp = f4zero();
            HEAT_PT(6);
            if (role == 2) {
                // ---- user row: grad (:60-62, :141), aggregator backward, step (:152, :166, :171-172).  This is the
                //      only work between [A] and the next sample's [G1] that the next sample waits for.
                float4 g0 = f4zero(), g1 = f4zero(), g2 = f4zero(), g3 = f4zero();
#pragma unroll
                for (int q = 0; q < 16; q += 4) {
                    g0 = g0 + *reinterpret_cast<const float4 *>(part + q * D + 4 * sl);
                    g1 = g1 + *reinterpret_cast<const float4 *>(part + (q + 1) * D + 4 * sl);
                    g2 = g2 + *reinterpret_cast<const float4 *>(part + (q + 2) * D + 4 * sl);
                    g3 = g3 + *reinterpret_cast<const float4 *>(part + (q + 3) * D + 4 * sl);
                }
                const float4 b0 = *reinterpret_cast<const float4 *>(sB), b1 = *reinterpret_cast<const float4 *>(sB + 4);
                const float4 b2 = *reinterpret_cast<const float4 *>(sB + 8), b3 = *reinterpret_cast<const float4 *>(sB + 12);
                const float4 b = (b0 + b1) + (b2 + b3);
                const float SB = (b.x + b.y) + (b.z + b.w);
                pp = sc3[0]; upd = sc3[1];
                const float4 pw = *reinterpret_cast<const float4 *>(pw_s + 4 * sl);
                const float q3 = (cE * SE) * ((rinv_u * rinv_u) * (rinv_u * sc3[3]));   // lgsum / (|u|^3 |p|)
                // sum_k lg_k*gu_n_k - lgsum*gu_p;  gu_n = (uu*n - un*u)/(|u|^3|n|),  gu_p = (uu*p - up*u)/(|u|^3|p|)
                float4 gu = rg;
                fma4(gu, cE, (g0 + g1) + (g2 + g3));
                fma4(gu, fmaf(q3, upd, -(cE * SB)), ua);
                fma4(gu, -q3 * uu, pw);
                *reinterpret_cast<float4 *>(gu06_ring + (it & (GRING - 1)) * D + 4 * sl) = omg * gu;   // behavior_aggregators.cpp:131-139
                gu = gamma * gu;                                                                       // :148-152
                gu = clamp4(gu, P.clip);
                fma4(ua, -P.lr, gu);
                urow_done = nu == u;
                if (urow_done) *reinterpret_cast<float4 *>(urow_s + 4 * sl) = ua;                      // the next sample's input row
                stcg4(P.user_w + (size_t)u * D + 4 * sl, ua);
                stcg4(P.user_g + (size_t)u * D + 4 * sl, gu);
                u_keep = ua; gu_keep = gu; keep_u = u;
            } else if (role == 1) {
                // ---- positive row grad (:142), clipped (:169); I[p] moves by -lr*gp: carry that into the cached column
                //      sums and publish the mean of the run's next sample (a gather overwrites both)
                const float rinv_p = rsqrt_eps(pp);
                const float q = (cE * SE) * ((rinv_u * rinv_p) * (rinv_p * rinv_p));   // lgsum / (|u| |p|^3)
                gp = rg;
                fma4(gp, -q * pp, ua); fma4(gp, q * upd, rw);          // lgsum * gp_p, gp_p = -(pp*u - up*p) / (|u||p|^3)
                gp = clamp4(gp, P.clip);
                float4 a4 = *reinterpret_cast<const float4 *>(accs + 4 * sl);
                fma4(a4, (fl & CTA_REGATHER) ? 0.f : -P.lr * (float)m->pmult, gp);
                *reinterpret_cast<float4 *>(accs + 4 * sl) = a4;
                *reinterpret_cast<float4 *>(mean_next + 4 * sl) = m->rH * a4;
                __syncwarp();
                mean_arrive<64>();
            } else {
                // ---- negative rows: lg*gn_n = c1*u + c2*N[k] (:139-143), clipped step, write-back (:127-150; sgd.cpp:14-26)
#pragma unroll
                for (int n = 0; n < NPS; ++n) {
                    const float lg = cE * Ek[n];               // :109
                    c1[n] = (lg * rn3[n]) * nn[n];
                    c2[n] = -(lg * rn3[n]) * un[n];
                }
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
                HEAT_PT(7);
            }
            HEAT_PT(4);
            if (dup) {
                // duplicate ids: replay in the reference's k order (grad row re-read per k, last writer
                // wins on the weight row); the positive row is written after all negatives (:169-174)
                for (int k = 0; k < K; ++k) {
                    if (role == 0 && slot == (k & 15)) {
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
                    row_bar<NROW>();
                }
            }
            if (role == 1) {
                // positive row write-back (:173-174); the loss is the gemv warp's
                float4 wp = rw;
                fma4(wp, -P.lr, gp);
                stcg4(P.item_w + (size_t)p * D + 4 * sl, wp);
                stcg4(P.item_g + (size_t)p * D + 4 * sl, gp);
            }
            HEAT_PT(5);
#ifdef HEAT_CTA_PROFILE
            if (prof_on && it == 1000)
                printf("abs tid %3d fresh %d: top %lld rows %lld G1 %lld pfi %lld bfly %lld A %lld cE %lld st %lld mw %lld S %lld end %lld\n", tid, (int)fresh,
                       prof_abs[0] % 1000000, prof_abs[1] % 1000000, prof_abs[2] % 1000000, prof_abs[9] % 1000000, prof_abs[10] % 1000000, prof_abs[3] % 1000000,
                       prof_abs[6] % 1000000, prof_abs[7] % 1000000, prof_abs[8] % 1000000, prof_abs[4] % 1000000, prof_abs[5] % 1000000);
#endif
            it += 1;
        }
        batch_bar<NT>();                                     // this batch is consumed; the batch after the next is visible
    }
    // ---- W0 step of the leftover (<= 32) samples (the reference drops a partial batch of W0 grads: drop_w0_tail)
    row_bar<NROW>();
    {
        const int left = (it & 31) == 0 ? (it > 0 ? 32 : 0) : (P.drop_w0_tail ? 0 : (int)(it & 31));
        if (left > 0 && role == 0) w0_step(it, left);
    }
}

}  // namespace heat
