// heat_b200/csrc/eval_kernel.cu -- evaluation: user x item score GEMM on tcgen05/TMEM fed by TMA,
// fused with a per-user masked top-k; Recall / NDCG / HitRate from the top-k lists.
//
// Replaces Engine::evaluate0 (train/engine.cpp:388-400: Eigen fp32 U.I^T into a dense
// [num_users, num_items] matrix, copied to numpy by init_modules.cpp:122-129) and
// metrics.evaluate_metrics (cf/metrics.py:5-36: mask train items, argpartition + argsort top-k,
// per-user Recall/NDCG/HitRate).  The score matrix is never stored unless heat_eval_dense asks.
//
// Numerics: the reference contracts in fp32.  tcgen05 has no fp32 input kind, so every fp32 value
// x is split once per call into hi = tf32_round(x) and lo = tf32_round(x - hi) and the tile
// product is accumulated in fp32 TMEM as  A_lo.B_hi + A_hi.B_lo + A_hi.B_hi  ("3xTF32"): the
// dropped lo.lo term and the rounding of lo are ~2^-22 relative per product, i.e. fp32-level.
//
// One CTA owns a block of 128 users (TMEM lanes) and streams every item tile past it:
//   warp 0      TMA producer   A (users, hi+lo) once per block; B (items, hi+lo) ring of stages
//   warp 1      MMA issuer     one elected thread, 3 x D/8 tcgen05.mma.kind::tf32 per item tile,
//                              accumulators double/quad-buffered in TMEM; owns TMEM alloc/dealloc
//   warps 2..5  epilogue       tcgen05.ld: thread = user row, NT consecutive item scores; a max-of-16
//                              threshold vote per group, survivors appended to the row's candidate
//                              buffer in shared memory; a full buffer is folded into the row's sorted
//                              list by the whole warp (train-mask binary search + bitonic sort)
#include <algorithm>
#include <cmath>
#include <vector>
#include "heat_ctx.cuh"
#include "sm100_ptx.cuh"

namespace heat {
namespace {

constexpr int EV_M = 128;          // users per CTA block = TMEM lanes
constexpr int EV_THREADS = 192;    // 6 warps
constexpr int ACC_STAGES = 4;      // TMEM accumulator ring
constexpr int MAX_B_STAGES = 4;

struct EvalParams {
    int U, I, k;                   // users in this launch, items, top-k (0 = dense output)
    int terms;                     // 3 = 3xTF32 (fp32-level), 1 = single TF32 pass
    int b_stages;
    int n_row_blocks, n_item_tiles;
    const long long *train_indptr; // [U+1] or null; per-user item lists sorted ascending
    const int *train_items;
    int *topk_ids;                 // [U, k]
    float *topk_scores;            // [U, k]
    float *dense_out;              // [U, I] when k == 0
    unsigned long long *glist;     // [grid, 128, 32*sort_e - CAP] sorted lists of the user block in flight (L2-resident scratch)
    int sort_e;                    // keys per lane in the compaction sort: 4, 8 or 16
};

// x -> hi = tf32(x) (round to nearest, low 13 mantissa bits zero), lo = tf32(x - hi)
__global__ void split_tf32_kernel(const float *__restrict__ src, float *__restrict__ hi, float *__restrict__ lo, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float x = src[i];
        uint32_t h, l;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
        const float hf = __uint_as_float(h);
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(x - hf));
        hi[i] = hf;
        lo[i] = __uint_as_float(l);
    }
}

template <int D, int NT>
struct EvalSmem {
    static constexpr int KA = D / 32;                       // 128-byte swizzle atoms along K
    static constexpr int A_ATOM = EV_M * 128;               // bytes of one [128 rows x 32 floats] atom
    static constexpr int B_ATOM = NT * 128;
    static constexpr int A_BYTES = 2 * KA * A_ATOM;         // hi + lo
    static constexpr int B_STAGE = 2 * KA * B_ATOM;         // hi + lo
    static constexpr int BAR_BYTES = 256;
    static constexpr int CAP = NT >= 64 ? 64 : 32;          // candidate-buffer entries per user row
    static size_t bytes(int stages, int k) { return 1024 + A_BYTES + (size_t)stages * B_STAGE + (k ? (size_t)CAP * EV_M * 8 : 0) + BAR_BYTES; }
};

// Top-k entries are 64-bit keys: order-preserving bits of the score in the high word, (0x7fffffff - item id) in the
// low word, so ONE unsigned compare orders by score and, among equal scores, puts the smaller id first.
__device__ __forceinline__ unsigned long long topk_key(float s, int id)
{
    unsigned u = __float_as_uint(s);
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    return ((unsigned long long)u << 32) | (unsigned)(0x7fffffff - id);
}
__device__ __forceinline__ float topk_key_score(unsigned long long key)
{
    unsigned u = (unsigned)(key >> 32);
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    return __uint_as_float(u);
}
__device__ __forceinline__ int topk_key_id(unsigned long long key) { return 0x7fffffff - (int)(unsigned)key; }

// Bitonic sort, descending, of 32*E keys held E per lane in blocked order (element index = lane*E + i).
template <int E>
__device__ __forceinline__ void bitonic_sort_desc(unsigned long long (&key)[E], int lane)
{
#pragma unroll
    for (int size = 2; size <= 32 * E; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            if (stride >= E) {
                const int lmask = stride / E;
                const bool keep_max = (((lane * E) & size) == 0) == ((lane & lmask) == 0);
#pragma unroll
                for (int i = 0; i < E; ++i) {
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, key[i], lmask);
                    const bool take = keep_max ? (o > key[i]) : (o < key[i]);
                    key[i] = take ? o : key[i];
                }
            } else {
#pragma unroll
                for (int i = 0; i < E; ++i) {
                    if ((i & stride) == 0) {
                        const int j = i | stride;
                        const bool desc = (((lane * E + i) & size) == 0);
                        const unsigned long long a = key[i], b = key[j];
                        const bool sw = desc ? (b > a) : (b < a);
                        key[i] = sw ? b : a;
                        key[j] = sw ? a : b;
                    }
                }
            }
        }
    }
}

// Per-lane (= per user row) state of the streaming top-k.  While the item tiles stream by, a lane only APPENDS the
// scores that beat its row's threshold to the row's candidate buffer (one shared-memory store); the threshold is
// the k-th best score as of the row's last compaction, so it lags -- that only admits a few extra candidates.
struct RowState {
    float thr;          // k-th best unmasked score at the last compaction (valid when full)
    bool full;          // the row's list holds k unmasked items
    bool has_list;      // the row's list in global scratch has been written for this user block
    int cnt;            // candidates waiting in the buffer
    long long tb, te;   // the user's train items [tb, te), sorted ascending (the -inf mask of metrics.py:21-24)
};

// Fold the candidate buffers of the rows in `rows` (bit r = row of lane r) into their sorted lists; the whole warp
// works on one row at a time: the candidates that are train items of the user are dropped (binary search in the
// user's sorted list -- deferred to here so that the streaming path never touches the CSR), list + candidates are
// sorted as 32*E keys, the best LCAP = 32*E - CAP go back to the row's list (global scratch, L2-resident) and the
// k-th one becomes the row's new threshold.  With `final` the first k entries are written out instead.
template <int E, int CAP>
__device__ __noinline__ void topk_compact(unsigned rows, RowState &st, int lane, int q, const unsigned long long *buf,
                                          unsigned long long *glist, const int *__restrict__ train_items, int k, bool final,
                                          int u_base, int *__restrict__ out_ids, float *__restrict__ out_scores)
{
    constexpr int LCAP = 32 * E - CAP, LE = LCAP / 32, CE = CAP / 32;
    const unsigned has = __ballot_sync(0xffffffffu, st.has_list);
    while (rows) {
        const int r = __ffs(rows) - 1;
        rows &= rows - 1;
        const int row = q * 32 + r;
        const int n = __shfl_sync(0xffffffffu, st.cnt, r);
        const long long tb = __shfl_sync(0xffffffffu, st.tb, r), te = __shfl_sync(0xffffffffu, st.te, r);
        unsigned long long key[E];
        unsigned long long *gl = glist + (size_t)row * LCAP;
#pragma unroll
        for (int i = 0; i < LE; ++i) key[i] = ((has >> r) & 1u) ? __ldcg(gl + lane + 32 * i) : 0ull;
#pragma unroll
        for (int i = 0; i < CE; ++i) {
            const int j = lane + 32 * i;
            unsigned long long c = j < n ? buf[j * EV_M + row] : 0ull;
            if (c != 0ull && tb < te) {
                const int col = 0x7fffffff - (int)(unsigned)c;
                long long lo = tb, hi = te;
                while (lo < hi) {
                    const long long mid = (lo + hi) >> 1;
                    if (__ldg(train_items + mid) < col) lo = mid + 1; else hi = mid;
                }
                if (lo < te && __ldg(train_items + lo) == col) c = 0ull;
            }
            key[LE + i] = c;
        }
        bitonic_sort_desc<E>(key, lane);
        // k-th best: element index k-1 = lane (k-1)/E, register (k-1)%E
        unsigned long long kth = 0ull;
#pragma unroll
        for (int i = 0; i < E; ++i) kth = ((k - 1) % E == i) ? key[i] : kth;
        kth = __shfl_sync(0xffffffffu, kth, (k - 1) / E);
        if (final) {
            const int u = u_base + row;
#pragma unroll
            for (int i = 0; i < E; ++i) {
                const int j = lane * E + i;
                if (j < k) {
                    out_ids[(size_t)u * k + j] = key[i] ? topk_key_id(key[i]) : -1;
                    if (out_scores) out_scores[(size_t)u * k + j] = key[i] ? topk_key_score(key[i]) : -INFINITY;
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < E; ++i) {
                const int j = lane * E + i;
                if (j < LCAP) __stcg(gl + j, key[i]);
            }
        }
        if (lane == r) {
            st.cnt = 0; st.has_list = true;
            if (kth != 0ull) { st.thr = topk_key_score(kth); st.full = true; }
        }
    }
    __syncwarp();
}

template <int CAP>
__device__ __forceinline__ void topk_compact_any(int E, unsigned rows, RowState &st, int lane, int q, const unsigned long long *buf,
                                                 unsigned long long *glist, const int *train_items, int k, bool final, int u_base,
                                                 int *out_ids, float *out_scores)
{
    if (E == 4) topk_compact<4, CAP>(rows, st, lane, q, buf, glist, train_items, k, final, u_base, out_ids, out_scores);
    else if (E == 8) topk_compact<8, CAP>(rows, st, lane, q, buf, glist, train_items, k, final, u_base, out_ids, out_scores);
    else topk_compact<16, CAP>(rows, st, lane, q, buf, glist, train_items, k, final, u_base, out_ids, out_scores);
}

template <int D, int NT>
__global__ void __launch_bounds__(EV_THREADS, 1)
eval_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
            const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo, const EvalParams P)
{
    using S = EvalSmem<D, NT>;
    constexpr int KA = S::KA;
    constexpr uint32_t TMEM_COLS = ACC_STAGES * NT;          // 256 (NT=64) or 128 (NT=32)
    constexpr uint32_t IDESC = ptx::idesc_tf32(EV_M, NT);

    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *sA = smem;                                       // [hi|lo][KA][128 x 128B]
    uint8_t *sB = sA + S::A_BYTES;                            // [stage][hi|lo][KA][NT x 128B]
    unsigned long long *buf = reinterpret_cast<unsigned long long *>(sB + (size_t)P.b_stages * S::B_STAGE);   // [CAP][128] candidate keys
    uint64_t *bars = reinterpret_cast<uint64_t *>(buf + (P.k ? (size_t)S::CAP * EV_M : 0));
    uint64_t *full = bars, *empty = bars + MAX_B_STAGES;
    uint64_t *tfull = bars + 2 * MAX_B_STAGES, *tempty = tfull + ACC_STAGES;
    uint64_t *a_full = tempty + ACC_STAGES, *a_empty = a_full + 1;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(a_empty + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int terms = P.terms;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA_hi); ptx::prefetch_tmap(&tmA_lo); ptx::prefetch_tmap(&tmB_hi); ptx::prefetch_tmap(&tmB_lo);
        for (int s = 0; s < MAX_B_STAGES; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
        for (int s = 0; s < ACC_STAGES; ++s) { ptx::mbar_init(&tfull[s], 1); ptx::mbar_init(&tempty[s], 4); }
        ptx::mbar_init(a_full, 1); ptx::mbar_init(a_empty, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            uint32_t it_b = 0, it_a = 0;
            for (int rb = blockIdx.x; rb < P.n_row_blocks; rb += gridDim.x, ++it_a) {
                ptx::mbar_wait(a_empty, (it_a & 1) ^ 1);
                ptx::mbar_expect_tx(a_full, (terms == 3 ? 2 : 1) * KA * S::A_ATOM);
                for (int a = 0; a < KA; ++a) {
                    ptx::tma_load_2d(sA + a * S::A_ATOM, &tmA_hi, a_full, a * 32, rb * EV_M);
                    if (terms == 3) ptx::tma_load_2d(sA + (KA + a) * S::A_ATOM, &tmA_lo, a_full, a * 32, rb * EV_M);
                }
                for (int t = 0; t < P.n_item_tiles; ++t, ++it_b) {
                    const int s = it_b % P.b_stages;
                    ptx::mbar_wait(&empty[s], ((it_b / P.b_stages) & 1) ^ 1);
                    ptx::mbar_expect_tx(&full[s], (terms == 3 ? 2 : 1) * KA * S::B_ATOM);
                    uint8_t *dst = sB + (size_t)s * S::B_STAGE;
                    for (int a = 0; a < KA; ++a) {
                        ptx::tma_load_2d(dst + a * S::B_ATOM, &tmB_hi, &full[s], a * 32, t * NT);
                        if (terms == 3) ptx::tma_load_2d(dst + (KA + a) * S::B_ATOM, &tmB_lo, &full[s], a * 32, t * NT);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            uint32_t it_b = 0, it_acc = 0, it_a = 0;
            const uint32_t a_addr = ptx::smem_u32(sA);
            for (int rb = blockIdx.x; rb < P.n_row_blocks; rb += gridDim.x, ++it_a) {
                ptx::mbar_wait(a_full, it_a & 1);
                for (int t = 0; t < P.n_item_tiles; ++t, ++it_b, ++it_acc) {
                    const int s = it_b % P.b_stages, acc = it_acc % ACC_STAGES;
                    ptx::mbar_wait(&tempty[acc], ((it_acc / ACC_STAGES) & 1) ^ 1);
                    ptx::mbar_wait(&full[s], (it_b / P.b_stages) & 1);
                    ptx::tc_fence_after();
                    const uint32_t b_addr = ptx::smem_u32(sB + (size_t)s * S::B_STAGE);
                    const uint32_t d_tmem = tmem_base + acc * NT;
                    uint32_t accumulate = 0;
                    // small terms first: A_lo.B_hi, A_hi.B_lo, then A_hi.B_hi
                    for (int term = (terms == 3 ? 0 : 2); term < 3; ++term) {
                        const uint32_t a_off = (term == 0 ? KA : 0) * S::A_ATOM;
                        const uint32_t b_off = (term == 1 ? KA : 0) * S::B_ATOM;
#pragma unroll
                        for (int a = 0; a < KA; ++a) {
#pragma unroll
                            for (int kk = 0; kk < 4; ++kk) {      // 4 x (K = 8 tf32 = 32 bytes) per 128-byte atom
                                const uint64_t da = ptx::smem_desc_k_sw128(a_addr + a_off + a * S::A_ATOM + kk * 32);
                                const uint64_t db = ptx::smem_desc_k_sw128(b_addr + b_off + a * S::B_ATOM + kk * 32);
                                ptx::mma_tf32(d_tmem, da, db, IDESC, accumulate);
                                accumulate = 1;
                            }
                        }
                    }
                    ptx::mma_commit(&empty[s]);       // B stage free once these MMAs have read it
                    ptx::mma_commit(&tfull[acc]);     // accumulator ready for the epilogue
                }
                ptx::mma_commit(a_empty);             // A block free for the next user block
            }
        }
    } else {
        // ================= epilogue: warps 2..5, TMEM lane quadrant = warp % 4 =================
        constexpr int CAP = S::CAP;
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int k = P.k, E = P.sort_e;
        unsigned long long *glist = P.glist + (size_t)blockIdx.x * EV_M * (32 * E - CAP);
        unsigned long long *my = buf + row;
        uint32_t it_acc = 0;
        for (int rb = blockIdx.x; rb < P.n_row_blocks; rb += gridDim.x) {
            const int u = rb * EV_M + row;
            const bool row_ok = u < P.U;
            RowState st;
            st.thr = row_ok ? -INFINITY : INFINITY; st.full = !row_ok; st.has_list = false; st.cnt = 0; st.tb = 0; st.te = 0;
            if (row_ok && P.train_indptr) { st.tb = P.train_indptr[u]; st.te = P.train_indptr[u + 1]; }
            for (int t = 0; t < P.n_item_tiles; ++t, ++it_acc) {
                const int acc = it_acc % ACC_STAGES;
                ptx::mbar_wait(&tfull[acc], (it_acc / ACC_STAGES) & 1);
                ptx::tc_fence_after();
                float v[NT];
#pragma unroll
                for (int c0 = 0; c0 < NT; c0 += 32)
                    ptx::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * NT + c0, *reinterpret_cast<float(*)[32]>(&v[c0]));
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&tempty[acc]);     // accumulator drained into registers
                const int col0 = t * NT;
                if (k == 0) {
                    if (!row_ok) continue;
                    float *out = P.dense_out + (size_t)u * P.I + col0;
#pragma unroll
                    for (int i = 0; i < NT; ++i)
                        if (col0 + i < P.I) out[i] = v[i];
                    continue;
                }
                if (col0 + NT > P.I) {                              // ragged last tile: the padding columns never qualify
#pragma unroll
                    for (int i = 0; i < NT; ++i)
                        if (col0 + i >= P.I) v[i] = __int_as_float(0x7fc00000);
                }
                // Streaming path: ONE warp vote per 16 scores (max-of-16 against the row thresholds).  Only when some
                // row of the warp has a survivor do the lanes test their 16 scores and append the survivors.
#pragma unroll
                for (int g = 0; g < NT; g += 16) {
                    float m = v[g];
#pragma unroll
                    for (int e = 1; e < 16; ++e) m = fmaxf(m, v[g + e]);
                    if (!__any_sync(0xffffffffu, !st.full || m > st.thr)) continue;
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        const float sc = v[g + e];
                        const bool cand = st.full ? (sc > st.thr) : (sc == sc);
                        if (cand) { my[st.cnt * EV_M] = topk_key(sc, col0 + g + e); ++st.cnt; }
                    }
                    const unsigned over = __ballot_sync(0xffffffffu, st.cnt > CAP - 16);
                    if (over) {
                        __syncwarp();
                        topk_compact_any<CAP>(E, over, st, lane, q, buf, glist, P.train_items, k, false, 0, nullptr, nullptr);
                    }
                }
            }
            if (k > 0) {
                const unsigned rows = __ballot_sync(0xffffffffu, row_ok);
                __syncwarp();
                topk_compact_any<CAP>(E, rows, st, lane, q, buf, glist, P.train_items, k, true, rb * EV_M, P.topk_ids, P.topk_scores);
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<TMEM_COLS>(tmem_base);
}

// Recall / NDCG / HitRate at 20 and 50 from the sorted top-k ids (cf/metrics.py:39-48, 87-113,
// 131-139; natural-log DCG, Recall denominator len(true)+1e-12), one warp per user, summed over
// users that have test items.  sums[0..5] = R@20, R@50, N@20, N@50, H@20, H@50; sums[6] = users.
__global__ void metrics_kernel(const int *__restrict__ topk_ids, int k, int U, const long long *__restrict__ test_indptr,
                               const int *__restrict__ test_items, double *sums)
{
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    double acc[7] = {0, 0, 0, 0, 0, 0, 0};
    for (int u = w; u < U; u += nw) {
        const long long tb = test_indptr[u], te = test_indptr[u + 1];
        const long long nt = te - tb;
        if (nt <= 0) continue;
        int hits20 = 0, hits50 = 0;
        double dcg20 = 0.0, dcg50 = 0.0;
        for (int r = 0; r < k && r < 50; ++r) {
            const int id = topk_ids[(size_t)u * k + r];
            bool hit = false;
            for (long long j = tb + lane; j < te; j += 32) hit |= (test_items[j] == id);
            hit = __any_sync(0xffffffffu, hit) && id >= 0;
            if (hit) {
                const double g = 1.0 / log(2.0 + (double)r);
                hits50 += 1; dcg50 += g;
                if (r < 20) { hits20 += 1; dcg20 += g; }
            }
        }
        if (lane == 0) {
            double idcg20 = 0.0, idcg50 = 0.0;
            for (int r = 0; r < 50 && r < nt; ++r) {
                const double g = 1.0 / log(2.0 + (double)r);
                idcg50 += g;
                if (r < 20) idcg20 += g;
            }
            acc[0] += hits20 / ((double)nt + 1e-12); acc[1] += hits50 / ((double)nt + 1e-12);
            acc[2] += dcg20 / (idcg20 + 1e-12);      acc[3] += dcg50 / (idcg50 + 1e-12);
            acc[4] += hits20 > 0 ? 1.0 : 0.0;        acc[5] += hits50 > 0 ? 1.0 : 0.0;
            acc[6] += 1.0;
        }
    }
    if (lane == 0 && acc[6] != 0.0)
        for (int i = 0; i < 7; ++i) atomicAdd(&sums[i], acc[i]);
}

// ---- host side ---------------------------------------------------------------------------------

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// fp32 row-major [rows, D] table; box = 32 floats (one 128-byte swizzle atom) x box_rows
int make_table_map(CUtensorMap *m, const float *base, uint64_t rows, int D, int box_rows)
{
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return fail("cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)D, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)D * sizeof(float)};
    cuuint32_t box[2] = {32, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return 0;
}

struct EvalScratch {
    float *user_hi = nullptr, *user_lo = nullptr, *item_hi = nullptr, *item_lo = nullptr;
    size_t user_elems = 0, item_elems = 0;
    int *topk_ids = nullptr; float *topk_scores = nullptr; size_t topk_elems = 0;
    long long *train_indptr = nullptr; int *train_items = nullptr; size_t train_n = 0, indptr_n = 0;
    long long *test_indptr = nullptr; int *test_items = nullptr; size_t test_n = 0;
    double *sums = nullptr;
    float *dense = nullptr; size_t dense_elems = 0;
    unsigned long long *glist = nullptr; size_t glist_elems = 0;
};

template <class T> int ensure(T **p, size_t *have, size_t need)
{
    if (*have >= need && *p) return 0;
    if (*p) cudaFree(*p);
    *p = nullptr; *have = 0;
    HEAT_CUDA(cudaMalloc(p, std::max<size_t>(need, 1) * sizeof(T)));
    *have = need;
    return 0;
}

EvalScratch *scratch(Ctx *c)
{
    if (!c->eval_scratch) c->eval_scratch = new EvalScratch();
    return static_cast<EvalScratch *>(c->eval_scratch);
}

int split_tables(Ctx *c, EvalScratch *s)
{
    const size_t D = c->cfg.emb_dim, ue = c->cfg.num_users * D, ie = c->cfg.num_items * D;
    size_t h;
    h = s->user_elems; if (int rc = ensure(&s->user_hi, &h, ue)) return rc;
    h = s->user_elems; if (int rc = ensure(&s->user_lo, &h, ue)) return rc;
    s->user_elems = ue;
    h = s->item_elems; if (int rc = ensure(&s->item_hi, &h, ie)) return rc;
    h = s->item_elems; if (int rc = ensure(&s->item_lo, &h, ie)) return rc;
    s->item_elems = ie;
    split_tf32_kernel<<<c->num_sms * 8, 256, 0, c->stream>>>(c->user_w, s->user_hi, s->user_lo, ue);
    split_tf32_kernel<<<c->num_sms * 8, 256, 0, c->stream>>>(c->item_w, s->item_hi, s->item_lo, ie);
    c->launches += 2;
    HEAT_CUDA(cudaGetLastError());
    return 0;
}

template <int D, int NT>
int launch_eval(Ctx *c, EvalScratch *s, uint64_t u_begin, uint64_t u_count, EvalParams P)
{
    using S = EvalSmem<D, NT>;
    CUtensorMap mA_hi, mA_lo, mB_hi, mB_lo;
    if (int rc = make_table_map(&mA_hi, s->user_hi + u_begin * D, u_count, D, EV_M)) return rc;
    if (int rc = make_table_map(&mA_lo, s->user_lo + u_begin * D, u_count, D, EV_M)) return rc;
    if (int rc = make_table_map(&mB_hi, s->item_hi, c->cfg.num_items, D, NT)) return rc;
    if (int rc = make_table_map(&mB_lo, s->item_lo, c->cfg.num_items, D, NT)) return rc;
    int dev_smem = 0;
    HEAT_CUDA(cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->device));
    int stages = MAX_B_STAGES;
    while (stages > 0 && S::bytes(stages, P.k) > (size_t)dev_smem) --stages;
    if (stages == 0) return fail("evaluation kernel: top-k list and operand tiles do not fit shared memory (k too large)");
    P.b_stages = stages;
    P.U = (int)u_count; P.I = (int)c->cfg.num_items;
    P.n_row_blocks = (int)((u_count + EV_M - 1) / EV_M);
    P.n_item_tiles = (int)((c->cfg.num_items + NT - 1) / NT);
    auto kern = eval_kernel<D, NT>;
    const size_t smem = S::bytes(stages, P.k);
    HEAT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = std::min(P.n_row_blocks, c->num_sms);
    if (P.k > 0) {
        // compaction sort width: 32*E keys >= k + CAP, E in {4, 8, 16}; the lists of the user blocks in flight
        int E = 4;
        while (32 * E < P.k + S::CAP) E *= 2;
        if (E > 16) return fail("evaluation kernel: top-k size too large");
        P.sort_e = E;
        size_t have = s->glist_elems;
        if (int rc = ensure(&s->glist, &have, (size_t)grid * EV_M * (32 * E - S::CAP))) return rc;
        s->glist_elems = have;
        P.glist = s->glist;
    }
    kern<<<grid, EV_THREADS, smem, c->stream>>>(mA_hi, mA_lo, mB_hi, mB_lo, P);
    c->launches += 1;
    HEAT_CUDA(cudaGetLastError());
    return 0;
}

int dispatch_eval(Ctx *c, EvalScratch *s, uint64_t u_begin, uint64_t u_count, const EvalParams &P)
{
    switch (c->cfg.emb_dim) {
    case 32: return launch_eval<32, 64>(c, s, u_begin, u_count, P);
    case 64: return launch_eval<64, 64>(c, s, u_begin, u_count, P);
    case 128: return launch_eval<128, 32>(c, s, u_begin, u_count, P);
    default: return fail("evaluation kernel supports emb_dim 32, 64 and 128");
    }
}

// CSR of items per user -> device, each user's list sorted ascending (the kernel's mask cursor needs it)
int upload_csr(Ctx *c, const int64_t *indptr, const int32_t *items, uint64_t U, long long **d_indptr, size_t *indptr_have,
               int **d_items, size_t *items_have, bool sort_rows)
{
    const size_t nnz = (size_t)indptr[U];
    if (indptr[0] != 0) return fail("CSR indptr must start at 0");
    for (uint64_t u = 0; u < U; ++u)
        if (indptr[u + 1] < indptr[u]) return fail("CSR indptr must be non-decreasing");
    std::vector<int32_t> sorted;
    const int32_t *src = items;
    if (sort_rows) {
        bool need = false;
        for (uint64_t u = 0; u < U && !need; ++u)
            need = !std::is_sorted(items + indptr[u], items + indptr[u + 1]);
        if (need) {
            sorted.assign(items, items + nnz);
            for (uint64_t u = 0; u < U; ++u) std::sort(sorted.begin() + indptr[u], sorted.begin() + indptr[u + 1]);
            src = sorted.data();
        }
    }
    if (int rc = ensure(d_indptr, indptr_have, (size_t)U + 1)) return rc;
    if (int rc = ensure(d_items, items_have, nnz)) return rc;
    HEAT_CUDA(cudaMemcpyAsync(*d_indptr, indptr, (U + 1) * sizeof(long long), cudaMemcpyHostToDevice, c->stream));
    if (nnz) HEAT_CUDA(cudaMemcpyAsync(*d_items, src, nnz * sizeof(int), cudaMemcpyHostToDevice, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));      // `sorted` is a host temporary
    return 0;
}

int topk_device(Ctx *c, EvalScratch *s, uint32_t k, const int64_t *train_indptr, const int32_t *train_items)
{
    if (!c->user_w) return fail("heat_bind_tables has not been called");
    if (k == 0 || k > 256) return fail("top-k size must be in 1..256");
    const uint64_t U = c->cfg.num_users;
    if (train_indptr) {
        if (!train_items) return fail("train_items is null");
        if (int rc = upload_csr(c, train_indptr, train_items, U, &s->train_indptr, &s->indptr_n, &s->train_items, &s->train_n, true)) return rc;
    }
    if (int rc = ensure(&s->topk_ids, &s->topk_elems, (size_t)U * k)) return rc;
    size_t h = 0;
    if (s->topk_scores) { cudaFree(s->topk_scores); s->topk_scores = nullptr; }
    if (int rc = ensure(&s->topk_scores, &h, (size_t)U * k)) return rc;
    HEAT_CUDA(cudaEventRecord(c->ev0, c->stream));
    if (int rc = split_tables(c, s)) return rc;
    EvalParams P{};
    P.k = (int)k; P.terms = 3;
    P.train_indptr = train_indptr ? s->train_indptr : nullptr;
    P.train_items = train_indptr ? s->train_items : nullptr;
    P.topk_ids = s->topk_ids; P.topk_scores = s->topk_scores;
    if (int rc = dispatch_eval(c, s, 0, U, P)) return rc;
    HEAT_CUDA(cudaEventRecord(c->ev1, c->stream));
    return 0;
}

}  // namespace

int eval_topk(Ctx *c, uint32_t k, const int64_t *train_indptr, const int32_t *train_items, int32_t *ids_host, float *scores_host)
{
    EvalScratch *s = scratch(c);
    if (int rc = topk_device(c, s, k, train_indptr, train_items)) return rc;
    const size_t n = (size_t)c->cfg.num_users * k;
    HEAT_CUDA(cudaMemcpyAsync(ids_host, s->topk_ids, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    if (scores_host) HEAT_CUDA(cudaMemcpyAsync(scores_host, s->topk_scores, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&c->last_eval_ms, c->ev0, c->ev1);
    return 0;
}

int eval_metrics(Ctx *c, const int64_t *train_indptr, const int32_t *train_items, const int64_t *test_indptr,
                 const int32_t *test_items, double out[6], double sums_out[7])
{
    EvalScratch *s = scratch(c);
    const uint64_t U = c->cfg.num_users;
    size_t tip = 0;
    if (s->test_indptr) { cudaFree(s->test_indptr); s->test_indptr = nullptr; }
    if (int rc = upload_csr(c, test_indptr, test_items, U, &s->test_indptr, &tip, &s->test_items, &s->test_n, false)) return rc;
    if (!s->sums) HEAT_CUDA(cudaMalloc(&s->sums, 7 * sizeof(double)));
    const uint32_t k = 50;
    if (int rc = topk_device(c, s, k, train_indptr, train_items)) return rc;
    HEAT_CUDA(cudaMemsetAsync(s->sums, 0, 7 * sizeof(double), c->stream));
    metrics_kernel<<<c->num_sms * 4, 256, 0, c->stream>>>(s->topk_ids, (int)k, (int)U, s->test_indptr, s->test_items, s->sums);
    c->launches += 1;
    HEAT_CUDA(cudaGetLastError());
    HEAT_CUDA(cudaEventRecord(c->ev1, c->stream));
    double sums[7];
    HEAT_CUDA(cudaMemcpyAsync(sums, s->sums, sizeof(sums), cudaMemcpyDeviceToHost, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&c->last_eval_ms, c->ev0, c->ev1);
    for (int i = 0; i < 6; ++i) out[i] = sums[6] > 0 ? sums[i] / sums[6] : 0.0;
    if (sums_out) std::copy(sums, sums + 7, sums_out);
    return 0;
}

// Dense compatibility path: the full [U, I] fp32 matrix, produced in row slabs of <= 1 GiB.
int eval_dense(Ctx *c, float *out_host)
{
    if (!c->user_w) return fail("heat_bind_tables has not been called");
    EvalScratch *s = scratch(c);
    const uint64_t U = c->cfg.num_users, I = c->cfg.num_items;
    uint64_t slab = std::max<uint64_t>(EV_M, ((1ull << 30) / (I * sizeof(float))) / EV_M * EV_M);
    slab = std::min<uint64_t>(slab, (U + EV_M - 1) / EV_M * EV_M);
    if (int rc = ensure(&s->dense, &s->dense_elems, (size_t)slab * I)) return rc;
    HEAT_CUDA(cudaEventRecord(c->ev0, c->stream));
    if (int rc = split_tables(c, s)) return rc;
    for (uint64_t u0 = 0; u0 < U; u0 += slab) {
        const uint64_t n = std::min(slab, U - u0);
        EvalParams P{};
        P.k = 0; P.terms = 3; P.dense_out = s->dense;
        if (int rc = dispatch_eval(c, s, u0, n, P)) return rc;
        HEAT_CUDA(cudaMemcpyAsync(out_host + u0 * I, s->dense, n * I * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    }
    HEAT_CUDA(cudaEventRecord(c->ev1, c->stream));
    HEAT_CUDA(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&c->last_eval_ms, c->ev0, c->ev1);
    return 0;
}

void eval_free(Ctx *c)
{
    EvalScratch *s = static_cast<EvalScratch *>(c->eval_scratch);
    if (!s) return;
    cudaFree(s->user_hi); cudaFree(s->user_lo); cudaFree(s->item_hi); cudaFree(s->item_lo);
    cudaFree(s->topk_ids); cudaFree(s->topk_scores); cudaFree(s->train_indptr); cudaFree(s->train_items);
    cudaFree(s->test_indptr); cudaFree(s->test_items); cudaFree(s->sums); cudaFree(s->dense); cudaFree(s->glist);
    delete s;
    c->eval_scratch = nullptr;
}

}  // namespace heat
