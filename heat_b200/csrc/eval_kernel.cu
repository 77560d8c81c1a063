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
//   warps 2..9  epilogue       two sets of four warps take alternate item tiles (two warps per scheduler
//                              hide each other's latencies).  tcgen05.ld: thread = user row, NT
//                              consecutive item scores; a max-of-16 threshold vote per group, survivors
//                              appended to the (set, row) candidate buffer in shared memory; a full
//                              buffer is folded into the (set, row) sorted list by the whole warp
//                              (train-mask binary search, bitonic sort of the candidates, bitonic merge);
//                              the sets exchange thresholds and set 0 merges the two lists at the end
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "heat_ctx.cuh"
#include "sm100_ptx.cuh"

namespace heat {
namespace {

#ifdef HEAT_EVAL_STATS
// development counters (-DHEAT_EVAL_STATS through HEAT_NVCC_DEFS): [0] candidates appended, [1] row folds, [2] score groups with a
// survivor somewhere in the warp, [3] score groups
__device__ unsigned long long g_eval_stats[4];
#define EV_STAT(i, n) atomicAdd(&g_eval_stats[i], (unsigned long long)(n))
#else
#define EV_STAT(i, n) do { } while (0)
#endif

constexpr int EV_M = 128;          // users per CTA block = TMEM lanes
constexpr int EV_CAP = 30;         // candidate-buffer entries per (epilogue set, user row): <= 32 (one key per lane in the fold);
                                   // 30 leaves room for a third B stage at D = 64
constexpr int ACC_STAGES = 4;      // TMEM accumulator ring: at most 4 tiles (fewer where 4 x NT + the 2*D columns of the A operand exceed 512)
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
    const float *user_w;           // [U, D] fp32 rows of this launch's users (the A operand is split and written to TMEM in the kernel)
    const int *perm;               // top-k mode: column c of the (norm-ordered) item operand is item perm[c]; null = identity
    unsigned long long *glist;     // [grid, 3, 128, 32*(sort_e-1)] sorted lists of the user block in flight (L2-resident scratch)
    int sort_e;                    // keys per lane in the fold: 4, 8 or 16
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

// ---- item order of the top-k mode ------------------------------------------------------------------------------
// The streaming top-k admits every score above the row's current k-th best, so the number of candidates (and of
// folds, which are what the epilogue's instructions go to) depends on how fast the thresholds rise: ~k ln(I/k) per row
// when the items come in id order.  Popular items are near the top of most users' lists, so the items are streamed in
// order of DESCENDING TRAIN-SET DEGREE: the thresholds are close to final after a small part of the table and the
// candidate count falls ~3x (measured on trained tables; ordering by the items' current norms does as well on trained
// tables but is the WORST order for barely trained ones, whose relevant items have the smaller norms).  The order is a
// property of the train CSR: computed on the host when a CSR is uploaded, resident with it.  The result does not
// depend on the order (keys carry the original item id); the TF32 split writes the item operand in that order.
// the TF32 split of the item table, row c of the output = item perm[c]
__global__ void split_tf32_perm_kernel(const float *__restrict__ src, const int *__restrict__ perm, float *__restrict__ hi,
                                       float *__restrict__ lo, int rows, int D)
{
    const int vpr = D / 4;                                   // float4 per row
    const size_t n4 = (size_t)rows * vpr;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(i / vpr), j = (int)(i % vpr);
        const float4 x = *reinterpret_cast<const float4 *>(src + (size_t)perm[c] * D + 4 * j);
        float4 h, l;
        const float xs[4] = {x.x, x.y, x.z, x.w};
        float hs[4], ls[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            uint32_t hb, lb;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(xs[e]));
            hs[e] = __uint_as_float(hb);
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lb) : "f"(xs[e] - hs[e]));
            ls[e] = __uint_as_float(lb);
        }
        h = make_float4(hs[0], hs[1], hs[2], hs[3]); l = make_float4(ls[0], ls[1], ls[2], ls[3]);
        *reinterpret_cast<float4 *>(hi + i * 4) = h;
        *reinterpret_cast<float4 *>(lo + i * 4) = l;
    }
}

template <int D, int NT>
struct EvalSmem {
    static constexpr int KA = D / 32;                       // 128-byte swizzle atoms along K
    static constexpr int ACC = (512 - 2 * D) / NT < ACC_STAGES ? (512 - 2 * D) / NT : ACC_STAGES;   // accumulator tiles in TMEM
    static constexpr int B_ATOM = NT * 128;
    static constexpr int A_BYTES = 0;                       // the A operand (users) lives in tensor memory
    static constexpr int B_STAGE = 2 * KA * B_ATOM;         // hi + lo
    static constexpr int BAR_BYTES = 256;
    static constexpr int NSETS = D >= 128 ? 1 : 2;          // epilogue warp sets (D = 128: shared memory holds one buffer set)
    static constexpr int THREADS = 64 + NSETS * 128;
    static constexpr int BUF_BYTES = NSETS * EV_CAP * EV_M * 8 + 2 * NSETS * EV_M * 8;   // candidate pairs + published thresholds
    static size_t bytes(int stages, int k) { return 1024 + A_BYTES + (size_t)stages * B_STAGE + (k ? (size_t)BUF_BYTES : 0) + BAR_BYTES; }
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
// the largest float below x (x finite or +inf; -inf stays -inf)
__device__ __forceinline__ float float_below(float x)
{
    const unsigned u = __float_as_uint(x);
    if (x > 0.f) return __uint_as_float(u - 1);
    if (x == 0.f) return __uint_as_float(0x80000001u);
    return x == -INFINITY ? x : __uint_as_float(u + 1);
}

// Warp-wide sorting networks over keys held in STRIDED order: element index = i*32 + lane for register i.
// Strides >= 32 are register-to-register, strides < 32 are shfl_xor exchanges.  All loops are counted so that they
// unroll completely (the key array must stay in registers) and every compare-exchange is branch-free.
template <int N> struct Log2 { static constexpr int v = 1 + Log2<N / 2>::v; };
template <> struct Log2<1> { static constexpr int v = 0; };

// keep the larger (want_max) or the smaller of (mine, other); equal keys are interchangeable
__device__ __forceinline__ unsigned long long cx_pick(unsigned long long mine, unsigned long long other, bool want_max)
{
    const bool gt = other > mine;
    return (gt == want_max) ? other : mine;
}

// Ascending bitonic sort of registers [R0, R0+C) (32*C elements).
template <int E, int R0, int C>
__device__ __forceinline__ void bitonic_sort_asc(unsigned long long (&key)[E], int lane)
{
    constexpr int LN = Log2<32 * C>::v;
#pragma unroll
    for (int ls = 1; ls <= LN; ++ls) {
        const int size = 1 << ls;
#pragma unroll
        for (int lt = ls - 1; lt >= 0; --lt) {
            const int stride = 1 << lt;
            if (stride >= 32) {
                const int rs = stride / 32;
#pragma unroll
                for (int i = 0; i < C; ++i) {
                    if ((i & rs) == 0) {
                        const bool asc = ((i * 32) & size) == 0;
                        const unsigned long long a = key[R0 + i], b = key[R0 + (i | rs)];
                        const bool sw = (b < a) == asc;
                        key[R0 + i] = sw ? b : a;
                        key[R0 + (i | rs)] = sw ? a : b;
                    }
                }
            } else {
                const bool lower = (lane & stride) == 0;
#pragma unroll
                for (int i = 0; i < C; ++i) {
                    const bool asc = (((i * 32) | lane) & size) == 0;
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, key[R0 + i], stride);
                    key[R0 + i] = cx_pick(key[R0 + i], o, asc != lower);
                }
            }
        }
    }
}
// Descending bitonic MERGE of all 32*E elements; the input must be bitonic (here: a descending list followed by
// ascending candidates).
template <int E>
__device__ __forceinline__ void bitonic_merge_desc(unsigned long long (&key)[E], int lane)
{
    constexpr int LE2 = Log2<E>::v;
#pragma unroll
    for (int l = LE2 - 1; l >= 0; --l) {
        const int rs = 1 << l;
#pragma unroll
        for (int i = 0; i < E; ++i) {
            if ((i & rs) == 0) {
                const unsigned long long a = key[i], b = key[i | rs];
                const bool sw = b > a;
                key[i] = sw ? b : a;
                key[i | rs] = sw ? a : b;
            }
        }
    }
#pragma unroll
    for (int lt = 4; lt >= 0; --lt) {
        const int stride = 1 << lt;
        const bool lower = (lane & stride) == 0;
#pragma unroll
        for (int i = 0; i < E; ++i) {
            const unsigned long long o = __shfl_xor_sync(0xffffffffu, key[i], stride);
            key[i] = cx_pick(key[i], o, lower);
        }
    }
}

// Per-lane (= per user row, per epilogue set) state of the streaming top-k lives in registers of the epilogue loop:
//   thr   lower bound of the row's k-th best unmasked score (-inf until k items are known; +inf for padding rows)
//   wptr  shared-memory address of the row's next free candidate slot (slot j of row r at buf + (j*128 + r)*8)
//   cur   how many of the user's train items lie below the columns folded so far (forward-only cursor)
// While the item tiles stream by, a lane only APPENDS the raw (score, column) pairs that beat its row's threshold
// (two predicated shared-memory stores); the threshold is the k-th best score as of the row's last fold (or the
// other set's), so it lags -- that only admits a few extra candidates.

enum { FOLD_STREAM = 0, FOLD_LAST = 1, FOLD_OUTPUT = 2 };

struct FoldArgs {                               // the rarely read arguments (stack); the hot ones travel in registers
    const unsigned long long *other_list;       // FOLD_OUTPUT: the other set's finished lists (or null)
    int u_base;                                 // first user of the block (FOLD_OUTPUT)
    int *out_ids; float *out_scores;
    unsigned long long *half_pub;               // this set's published ceil(k/2)-th scores [128]
};
struct FoldHot {
    int k, lane, row0;                          // top-k size, lane, first row of the warp's quadrant
    uint32_t buf_s;                             // shared-space address of the set's candidate buffers [EV_CAP][128] x 8 B
    unsigned long long *glist;                  // the set's lists of the user block in flight [128][32*(E-1)]
    const int *train_items;
    const int *perm;                            // column -> item id (null = identity)
    unsigned long long *thr_pub;                // this set's published k-th scores [128] (or null)
    unsigned tag;                               // user block tag of the published thresholds
};

__device__ __forceinline__ unsigned long long lds64(uint32_t addr)
{
    unsigned long long v;
    asm volatile("ld.shared.b64 %0, [%1];" : "=l"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }

// Fold the candidate buffers of the rows in `rows` (bit r = row of lane r) into their sorted lists; the whole warp
// works on one row at a time.  Candidates that are train items of the user are dropped: candidates arrive in column
// order and the user's list is sorted, so the warp reads the list 32 items at a time from the row's cursor (one
// coalesced load) and every lane looks its column up by a 5-step shuffle binary search -- the streaming path never
// touches the CSR.  The <= 32 survivors become keys, are sorted ascending and bitonic-merged with the row's
// descending list of LCAP = 32*(E-1) keys (global scratch, L2-resident); the k-th key becomes the row's new
// threshold.  FOLD_OUTPUT (end of the user block) also merges the other set's finished list, 32 keys at a time, and
// writes the first k entries out instead of storing the list.
// Returns the lane's new (threshold, cursor) pair; `cnt` of the folded rows is reset by the caller.
template <int E>
__device__ __forceinline__ unsigned long long topk_fold(const FoldArgs &A, const FoldHot H, unsigned rows, float thr, int cnt,
                                                        long long tb, int ntrain, int cur, unsigned has, int col_next, int mode)
{
    constexpr int LE = E - 1, LCAP = 32 * LE;
    const int lane = H.lane, k = H.k;
    const int *__restrict__ train_items = H.train_items;
    while (rows) {
        const int r = __ffs(rows) - 1;
        rows &= rows - 1;
        const int row = H.row0 + r;
        const int n = __shfl_sync(0xffffffffu, cnt, r);
        const long long tbr = __shfl_sync(0xffffffffu, tb, r);
        const int nr = __shfl_sync(0xffffffffu, ntrain, r);
        const int pos = __shfl_sync(0xffffffffu, cur, r);        // (unused since the columns are norm-ordered: the whole list is scanned)
        unsigned long long key[E];
        unsigned long long *gl = H.glist + (size_t)row * LCAP;
#pragma unroll
        for (int i = 0; i < LE; ++i) key[i] = ((has >> r) & 1u) ? __ldcg(gl + lane + 32 * i) : 0ull;
        const unsigned long long raw = lane < n ? lds64(H.buf_s + (uint32_t)(lane * EV_M + row) * 8u) : 0ull;   // (column, score bits)
        // the candidate's item id (the columns are in norm order); candidates that are train items of the user are
        // dropped: the user's sorted list is read 32 ids at a time and every lane looks its id up in the chunks whose
        // range can hold it (5-step shuffle binary search)
        const int col = lane < n ? (H.perm ? __ldg(H.perm + (int)(unsigned)raw) : (int)(unsigned)raw) : -1;
        bool masked = false;
        for (int pos = 0; pos < nr; pos += 32) {
            const int tv = pos + lane < nr ? __ldg(train_items + tbr + pos + lane) : 0x7fffffff;
            const int first = __shfl_sync(0xffffffffu, tv, 0), last = __shfl_sync(0xffffffffu, tv, 31);
            if (!__any_sync(0xffffffffu, col >= first && col <= last)) continue;
            int lo = 0;
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) {
                const int tt = __shfl_sync(0xffffffffu, tv, lo + s - 1);
                if (tt < col) lo += s;
            }
            masked |= __shfl_sync(0xffffffffu, tv, lo) == col;
        }
        key[LE] = (lane < n && !masked) ? topk_key(__uint_as_float((unsigned)(raw >> 32)), col) : 0ull;
        bitonic_sort_asc<E, LE, 1>(key, lane);
        bitonic_merge_desc<E>(key, lane);
        if (mode == FOLD_OUTPUT && A.other_list) {
            const unsigned long long *ol = A.other_list + (size_t)row * LCAP;
            for (int c0 = 0; c0 < k; c0 += 32) {          // the other set's list is descending: read 32 keys reversed
                key[LE] = __ldcg(ol + c0 + 31 - lane);
                bitonic_merge_desc<E>(key, lane);
            }
        }
        // k-th best: element index k-1 = register (k-1)/32, lane (k-1)%32
        unsigned long long kth = 0ull;
#pragma unroll
        for (int i = 0; i < E; ++i) kth = ((k - 1) / 32 == i) ? key[i] : kth;
        kth = __shfl_sync(0xffffffffu, kth, (k - 1) & 31);
        // ceil(k/2)-th best: with the other set's ceil(k/2)-th it bounds the k-th best of the union (see the adoption)
        const int ha = (k + 1) / 2 - 1;
        unsigned long long hth = 0ull;
#pragma unroll
        for (int i = 0; i < E; ++i) hth = (ha / 32 == i) ? key[i] : hth;
        hth = __shfl_sync(0xffffffffu, hth, ha & 31);
        if (mode == FOLD_OUTPUT) {
            const int u = A.u_base + row;
#pragma unroll
            for (int i = 0; i < LE; ++i) {
                const int j = i * 32 + lane;
                if (j < k) {
                    A.out_ids[(size_t)u * k + j] = key[i] ? topk_key_id(key[i]) : -1;
                    if (A.out_scores) A.out_scores[(size_t)u * k + j] = key[i] ? topk_key_score(key[i]) : -INFINITY;
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < LE; ++i) __stcg(gl + lane + 32 * i, key[i]);
        }
        if (lane == r) {
            cur = pos;
            if (kth != 0ull) thr = fmaxf(thr, topk_key_score(kth));
            if (H.thr_pub) {
                if (kth != 0ull) H.thr_pub[row] = ((unsigned long long)H.tag << 32) | __float_as_uint(topk_key_score(kth));
                if (hth != 0ull) A.half_pub[row] = ((unsigned long long)H.tag << 32) | __float_as_uint(topk_key_score(hth));
            }
        }
    }
    __syncwarp();
    return ((unsigned long long)(unsigned)cur << 32) | __float_as_uint(thr);
}

// The ONE out-of-line copy of the fold (the streaming loop stays small enough for the instruction cache).
__device__ __noinline__ unsigned long long topk_fold_dispatch(const FoldArgs &A, const FoldHot H, int E, unsigned rows, float thr, int cnt,
                                                              long long tb, int ntrain, int cur, unsigned has, int col_next, int mode)
{
    if (E == 4) return topk_fold<4>(A, H, rows, thr, cnt, tb, ntrain, cur, has, col_next, mode);
    if (E == 8) return topk_fold<8>(A, H, rows, thr, cnt, tb, ntrain, cur, has, col_next, mode);
    return topk_fold<16>(A, H, rows, thr, cnt, tb, ntrain, cur, has, col_next, mode);
}

template <int D, int NT>
__global__ void __launch_bounds__((EvalSmem<D, NT>::THREADS), 1)
eval_kernel(const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo, const EvalParams P)
{
    using S = EvalSmem<D, NT>;
    constexpr int KA = S::KA;
    constexpr int ACC = S::ACC;
    constexpr uint32_t A_COL0 = ACC * NT;                    // A_hi at columns [A_COL0, A_COL0 + D), A_lo behind it
    constexpr uint32_t TMEM_COLS = 512;                      // accumulator ring (256 / 128 columns) + A (2*D <= 256)
    static_assert(ACC >= 2 && ACC * NT + 2 * D <= 512, "tensor memory: accumulators + the A operand");
    constexpr uint32_t IDESC = ptx::idesc_tf32(EV_M, NT);

    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *sB = smem;                                       // [stage][hi|lo][KA][NT x 128B]
    constexpr int NSETS = S::NSETS;
    unsigned long long *buf = reinterpret_cast<unsigned long long *>(sB + (size_t)P.b_stages * S::B_STAGE);   // [set][EV_CAP][128] candidate keys
    unsigned long long *thr_sh = buf + (size_t)NSETS * EV_CAP * EV_M;                                          // [k-th | ceil(k/2)-th][set][128] (user block tag, score)
    uint64_t *bars = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(buf) + (P.k ? (size_t)S::BUF_BYTES : 0));
    uint64_t *full = bars, *empty = bars + MAX_B_STAGES;
    uint64_t *tfull = bars + 2 * MAX_B_STAGES, *tempty = tfull + ACC_STAGES;
    uint64_t *a_full = tempty + ACC_STAGES, *a_empty = a_full + 1;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(a_empty + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int terms = P.terms;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmB_hi); ptx::prefetch_tmap(&tmB_lo);
        for (int s = 0; s < MAX_B_STAGES; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
        for (int s = 0; s < ACC; ++s) { ptx::mbar_init(&tfull[s], 1); ptx::mbar_init(&tempty[s], 4); }
        ptx::mbar_init(a_full, 4); ptx::mbar_init(a_empty, 1);       // a_full: the four warps of epilogue set 0 have written A
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    if (P.k > 0 && threadIdx.x < NSETS * EV_M) { thr_sh[threadIdx.x] = 0ull; thr_sh[NSETS * EV_M + threadIdx.x] = 0ull; }   // tag 0 = nothing published
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            uint32_t it_b = 0, it_a = 0;
            for (int rb = blockIdx.x; rb < P.n_row_blocks; rb += gridDim.x, ++it_a) {
                for (int t = 0; t < P.n_item_tiles; ++t, ++it_b) {
                    const int s = it_b % P.b_stages;
                    HEAT_MBAR_WAIT(&empty[s], ((it_b / P.b_stages) & 1) ^ 1);
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
        {   // the whole warp walks the tile loop (converged); one elected lane issues the MMAs and the commits
            uint32_t it_b = 0, it_acc = 0, it_a = 0;
            // Descriptors: ONE thread issues every MMA, so its instruction count per tile is on the critical path (24 MMAs
            // per 128x64 tile; building both descriptors from scratch for each made this thread, not the tensor pipe or
            // the epilogue, the limiter: ~35 dependent integer instructions per MMA).  The start-address field is
            // (addr >> 4) in the low 14 bits and every operand lies below 256 KiB, so desc(base + off) = desc(base) + (off >> 4):
            // one base descriptor per operand, compile-time offsets in fully unrolled loops.
            const uint32_t a_tmem0 = tmem_base + A_COL0;      // A_hi; A_lo at + D
            const uint64_t b_desc00 = ptx::smem_desc_k_sw128(ptx::smem_u32(sB));
            for (int rb = blockIdx.x; rb < P.n_row_blocks; rb += gridDim.x, ++it_a) {
                HEAT_MBAR_WAIT(a_full, it_a & 1);
                for (int t = 0; t < P.n_item_tiles; ++t, ++it_b, ++it_acc) {
                    const int s = it_b % P.b_stages, acc = it_acc % ACC;
                    HEAT_MBAR_WAIT(&tempty[acc], ((it_acc / ACC) & 1) ^ 1);
                    HEAT_MBAR_WAIT(&full[s], (it_b / P.b_stages) & 1);
                    ptx::tc_fence_after();
                    const uint64_t b_desc0 = b_desc00 + (uint64_t)(((uint32_t)s * (uint32_t)S::B_STAGE) >> 4);
                    const uint32_t d_tmem = tmem_base + acc * NT;
                    if (ptx::elect_one()) {
                    uint32_t accumulate = 0;
                    // small terms first: A_lo.B_hi, A_hi.B_lo, then A_hi.B_hi
#pragma unroll
                    for (int term = 0; term < 3; ++term) {
                        if (terms != 3 && term < 2) continue;
                        const uint32_t a_off = term == 0 ? D : 0;               // columns: lo behind hi
                        const uint32_t b_off = (term == 1 ? KA : 0) * S::B_ATOM;
#pragma unroll
                        for (int a = 0; a < KA; ++a) {
#pragma unroll
                            for (int kk = 0; kk < 4; ++kk) {      // 4 x (K = 8 tf32 = 32 bytes) per 128-byte atom
                                const uint64_t db = b_desc0 + (uint64_t)((b_off + a * S::B_ATOM + kk * 32) >> 4);
                                ptx::mma_tf32_ts(d_tmem, a_tmem0 + a_off + a * 32 + kk * 8, db, IDESC, accumulate);
                                accumulate = 1;
                            }
                        }
                    }
                    ptx::mma_commit(&empty[s]);       // B stage free once these MMAs have read it
                    ptx::mma_commit(&tfull[acc]);     // accumulator ready for the epilogue
                    }
                    __syncwarp();
                }
                if (ptx::elect_one()) ptx::mma_commit(a_empty);             // A block free for the next user block
                __syncwarp();
            }
        }
    } else {
        // ================= epilogue: NSETS sets of four warps, TMEM lane quadrant = warp % 4 =================
        const int set = (warp - 2) >> 2, q = warp & 3;
        const int row = q * 32 + lane;
        const int k = P.k;
        const int lcap = 32 * (P.sort_e - 1);
        // global scratch of this CTA: set 0's lists, then two generations of set 1's lists (set 0 reads generation g
        // while set 1 may already be writing g+1; the accumulator ring keeps the sets within a few tiles of each other)
        unsigned long long *gl_base = P.glist + (size_t)blockIdx.x * 3 * EV_M * lcap;
        const uint32_t buf_s = ptx::smem_u32(buf) + (uint32_t)set * EV_CAP * EV_M * 8u;
        const uint32_t my_s = buf_s + (uint32_t)row * 8u;                    // slot 0 of this lane's row; slots are 1 KiB apart
        FoldArgs FA;
        FA.other_list = nullptr; FA.out_ids = P.topk_ids; FA.out_scores = P.topk_scores;
        FA.half_pub = thr_sh + (NSETS + set) * EV_M;
        FoldHot FH;
        FH.k = k; FH.lane = lane; FH.row0 = q * 32; FH.buf_s = buf_s; FH.train_items = P.train_items; FH.perm = P.perm;
        FH.thr_pub = (NSETS > 1 && k > 0) ? thr_sh + set * EV_M : nullptr;
        unsigned rbi = 0;
        for (int rb = blockIdx.x; rb < P.n_row_blocks; rb += gridDim.x, ++rbi) {
            const int u = rb * EV_M + row;
            const bool row_ok = u < P.U;
            // ---- the A operand of this user block: set 0's four warps (one per TMEM lane quadrant) split their rows into
            //      tf32 hi / lo and write them into tensor memory -- lane = user row, one element per column.  The MMAs then
            //      read A from TMEM: with both operands in shared memory every K=8 step re-read 4 KB of A next to 2 KB of B,
            //      and that operand traffic, not the tensor pipe, paced the kernel.
            if (set == 0) {
                HEAT_MBAR_WAIT(a_empty, (rbi & 1) ^ 1);                        // the previous block's MMAs have read A
                ptx::tc_fence_after();
                const float *urow = P.user_w + (size_t)(row_ok ? u : 0) * D;
                const uint32_t ta = tmem_base + ((uint32_t)(q * 32) << 16) + A_COL0;
#pragma unroll 1
                for (int c0 = 0; c0 < D; c0 += 32) {
                    uint32_t hi[32], lo[32];
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        const float4 x = row_ok ? __ldg(reinterpret_cast<const float4 *>(urow + c0 + j)) : make_float4(0.f, 0.f, 0.f, 0.f);
                        const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi[j + e]) : "f"(xs[e]));
                            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo[j + e]) : "f"(xs[e] - __uint_as_float(hi[j + e])));
                        }
                    }
                    ptx::tmem_st32(ta + c0, hi);
                    ptx::tmem_st32(ta + D + c0, lo);
                }
                ptx::tmem_st_wait();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(a_full);
            }
            float thr = row_ok ? -INFINITY : INFINITY;
            uint32_t wptr = my_s;
            long long tb = 0;
            int ntrain = 0, cur = 0;
            unsigned has = 0;                                                 // rows of this warp whose list exists (uniform)
            if (row_ok && k > 0 && P.train_indptr) { tb = P.train_indptr[u]; ntrain = (int)(P.train_indptr[u + 1] - tb); }
            FH.glist = gl_base + (size_t)(set == 0 ? 0 : 1 + (rbi & 1)) * EV_M * lcap;
            FA.u_base = rb * EV_M; FH.tag = rbi + 1;
            const unsigned long long *thr_other = thr_sh + (set ^ 1) * EV_M + row;
            float *dense_row = (k == 0 && row_ok) ? P.dense_out + (size_t)u * P.I : nullptr;
            const uint32_t it0 = rbi * (uint32_t)P.n_item_tiles;

            auto fold = [&](unsigned rows, int col_next, int mode) {
                const unsigned long long ret = topk_fold_dispatch(FA, FH, P.sort_e, rows, thr, (int)((wptr - my_s) >> 10), tb, ntrain, cur, has, col_next, mode);
                thr = __uint_as_float((unsigned)ret);
                cur = (int)(unsigned)(ret >> 32);
                if ((rows >> lane) & 1u) wptr = my_s;
                has |= rows;
            };
            // One group of 16 consecutive scores of this lane's user.  Streaming path: ONE warp vote per group
            // (max-of-16 against the row thresholds); only when some row of the warp has a survivor do the lanes look
            // at their max-of-4 sub-groups and append the survivors.
            auto group = [&](uint32_t (&r)[16], int colg) {
                float v[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
                if (k == 0) {
                    if (dense_row) {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (colg + i < P.I) dense_row[colg + i] = v[i];
                    }
                    return;
                }
                if (colg + 16 > P.I) {                              // ragged tail: the padding columns never qualify
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (colg + i >= P.I) v[i] = -INFINITY;
                }
                float m4[4];
#pragma unroll
                for (int h = 0; h < 4; ++h)
                    m4[h] = fmaxf(fmaxf(v[4 * h], v[4 * h + 1]), fmaxf(v[4 * h + 2], v[4 * h + 3]));
                const float m = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
                if (lane == 0) EV_STAT(3, 1);
                if (!__any_sync(0xffffffffu, m > thr)) return;
                if (lane == 0) EV_STAT(2, 1);
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                    if (m4[h] > thr) {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            if (v[4 * h + e] > thr) {
                                sts32(wptr, (uint32_t)(colg + 4 * h + e));
                                sts32(wptr + 4, __float_as_uint(v[4 * h + e]));
                                wptr += EV_M * 8;
                                EV_STAT(0, 1);
                            }
                        }
                    }
                    // a row folds as soon as its buffer may not take four more candidates
                    const unsigned over = __ballot_sync(0xffffffffu, wptr > my_s + (EV_CAP - 4) * EV_M * 8);
                    if (over) { if (lane == 0) EV_STAT(1, __popc(over)); fold(over, colg + 4 * h + 4, FOLD_STREAM); }
                }
            };

            for (int t = (int)((it0 + set) & (NSETS - 1)); t < P.n_item_tiles; t += NSETS) {
                const uint32_t it_acc = it0 + t;
                const int acc = it_acc % ACC;
                HEAT_MBAR_WAIT(&tfull[acc], (it_acc / ACC) & 1);
                ptx::tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * NT;
                const int col0 = t * NT;
                if (NSETS > 1 && k > 0) {
                    // adopt the other set's threshold for the same user block, one ulp lower: an equal score on this
                    // side may carry the smaller item id and must still reach the final merge
                    // Also: this set's ceil(k/2)-th best and the other set's are both exceeded-or-met by ceil(k/2) items of
                    // disjoint column sets, so their minimum bounds the k-th best of the union -- for interleaved tiles
                    // that is close to the exact threshold, which neither list alone knows.
                    const unsigned long long o = *thr_other, oh = thr_other[NSETS * EV_M], mh = thr_sh[(NSETS + set) * EV_M + row];
                    if ((unsigned)(o >> 32) == rbi + 1) thr = fmaxf(thr, float_below(__uint_as_float((unsigned)o)));
                    if ((unsigned)(oh >> 32) == rbi + 1 && (unsigned)(mh >> 32) == rbi + 1)
                        thr = fmaxf(thr, float_below(fminf(__uint_as_float((unsigned)oh), __uint_as_float((unsigned)mh))));
                }
                // 16-column TMEM loads, software-pipelined: the next group is in flight while this one is filtered
                uint32_t ra[16], rb16[16];
                ptx::tmem_ld16_issue(taddr, ra);
                ptx::tmem_ld_wait16(ra);
#pragma unroll 1
                for (int g = 0; g < NT; g += 32) {
                    ptx::tmem_ld16_issue(taddr + g + 16, rb16);
                    group(ra, col0 + g);
                    ptx::tmem_ld_wait16(rb16);
                    if (g + 32 < NT) {
                        ptx::tmem_ld16_issue(taddr + g + 32, ra);
                    } else {
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive(&tempty[acc]);     // accumulator drained into registers
                    }
                    group(rb16, col0 + g + 16);
                    if (g + 32 < NT) ptx::tmem_ld_wait16(ra);
                }
            }
            if (k > 0) {
                const unsigned rows = __ballot_sync(0xffffffffu, row_ok);
                __syncwarp();
                if (NSETS == 1) {
                    fold(rows, 0x7fffffff, FOLD_OUTPUT);
                } else if (set == 1) {
                    fold(rows, 0x7fffffff, FOLD_LAST);
                    __threadfence_block();
                    asm volatile("bar.arrive 1, 256;" ::: "memory");          // set 1's lists of this user block are complete
                } else {
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    FA.other_list = gl_base + (size_t)(1 + (rbi & 1)) * EV_M * lcap;
                    fold(rows, 0x7fffffff, FOLD_OUTPUT);
                    FA.other_list = nullptr;
                }
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
// Lane = rank (two ranks per lane for k = 50): every lane compares its ids with the user's test items (broadcast loads),
// one ballot per half gives the hit masks; the discounts 1/ln(2+r) and their prefix sums (the ideal DCG) come from a
// shared-memory table built once per block -- fp64 logarithms per hit made this kernel 0.41 ms of a 4.5 ms evaluation.
__global__ void metrics_kernel(const int *__restrict__ topk_ids, int k, int U, const long long *__restrict__ test_indptr,
                               const int *__restrict__ test_items, double *sums)
{
    __shared__ double disc[64], idcg[65];                 // disc[r] = 1/ln(2+r); idcg[n] = sum_{r<n} disc[r]
    if (threadIdx.x < 64) disc[threadIdx.x] = 1.0 / log(2.0 + (double)threadIdx.x);
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0;
        idcg[0] = 0.0;
        for (int r = 0; r < 64; ++r) { a += disc[r]; idcg[r + 1] = a; }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    const int kk = k < 50 ? k : 50;
    double acc[7] = {0, 0, 0, 0, 0, 0, 0};
    for (int u = w; u < U; u += nw) {
        const long long tb = test_indptr[u], te = test_indptr[u + 1];
        const long long nt = te - tb;
        if (nt <= 0) continue;
        const int id0 = lane < kk ? topk_ids[(size_t)u * k + lane] : -1;
        const int id1 = lane + 32 < kk ? topk_ids[(size_t)u * k + lane + 32] : -1;
        bool h0 = false, h1 = false;
        for (long long j = tb; j < te; ++j) {
            const int t = __ldg(test_items + j);
            h0 |= t == id0; h1 |= t == id1;
        }
        const unsigned m0 = __ballot_sync(0xffffffffu, h0 && id0 >= 0), m1 = __ballot_sync(0xffffffffu, h1 && id1 >= 0);
        if (lane == 0) {
            const int hits20 = __popc(m0 & 0xfffffu), hits50 = __popc(m0) + __popc(m1);
            double dcg20 = 0.0, dcg50 = 0.0;
            for (unsigned m = m0; m; m &= m - 1) { const int r = __ffs(m) - 1; dcg50 += disc[r]; if (r < 20) dcg20 += disc[r]; }
            for (unsigned m = m1; m; m &= m - 1) dcg50 += disc[32 + __ffs(m) - 1];
            const double i20 = idcg[nt < 20 ? (int)nt : 20], i50 = idcg[nt < 50 ? (int)nt : 50];
            acc[0] += hits20 / ((double)nt + 1e-12); acc[1] += hits50 / ((double)nt + 1e-12);
            acc[2] += dcg20 / (i20 + 1e-12);         acc[3] += dcg50 / (i50 + 1e-12);
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
    float *item_hi = nullptr, *item_lo = nullptr;
    size_t item_elems = 0;
    int *topk_ids = nullptr; float *topk_scores = nullptr; size_t topk_elems = 0;
    long long *train_indptr = nullptr; int *train_items = nullptr; size_t train_n = 0, indptr_n = 0;
    long long *test_indptr = nullptr; int *test_items = nullptr; size_t test_n = 0, test_indptr_n = 0;
    uint64_t train_fp = 0, test_fp = 0;     // fingerprints of the CSRs resident on the device (0 = none)
    size_t topk_scores_elems = 0;
    double *sums = nullptr;
    float *dense = nullptr; size_t dense_elems = 0;
    unsigned long long *glist = nullptr; size_t glist_elems = 0;
    int *perm = nullptr; size_t perm_elems = 0;     // items by descending train-set degree (resident with the train CSR; top-k mode)
    bool perm_ok = false;
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

int split_tables(Ctx *c, EvalScratch *s, bool norm_order /* item operand in s->perm order */)
{
    const size_t D = c->cfg.emb_dim, ie = c->cfg.num_items * D;
    size_t h;
    h = s->item_elems; if (int rc = ensure(&s->item_hi, &h, ie)) return rc;
    h = s->item_elems; if (int rc = ensure(&s->item_lo, &h, ie)) return rc;
    s->item_elems = ie;
    if (norm_order) {
        split_tf32_perm_kernel<<<c->num_sms * 8, 256, 0, c->stream>>>(c->item_w, s->perm, s->item_hi, s->item_lo, (int)c->cfg.num_items, (int)D);
        c->launches += 1;
    } else {
        split_tf32_kernel<<<c->num_sms * 8, 256, 0, c->stream>>>(c->item_w, s->item_hi, s->item_lo, ie);
        c->launches += 1;
    }
    HEAT_CUDA(cudaGetLastError());
    return 0;
}

template <int D, int NT>
int launch_eval(Ctx *c, EvalScratch *s, uint64_t u_begin, uint64_t u_count, EvalParams P)
{
    using S = EvalSmem<D, NT>;
    CUtensorMap mB_hi, mB_lo;
    P.user_w = c->user_w + u_begin * D;
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
        // fold width: 32*E keys = a sorted list of 32*(E-1) >= k keys + EV_CAP candidates, E in {4, 8, 16}
        int E = 4;
        while (32 * (E - 1) < P.k) E *= 2;
        if (E > 16) return fail("evaluation kernel: top-k size too large");
        P.sort_e = E;
        size_t have = s->glist_elems;
        if (int rc = ensure(&s->glist, &have, (size_t)grid * 3 * EV_M * 32 * (E - 1))) return rc;
        s->glist_elems = have;
        P.glist = s->glist;
    }
    kern<<<grid, S::THREADS, smem, c->stream>>>(mB_hi, mB_lo, P);
    c->launches += 1;
    HEAT_CUDA(cudaGetLastError());
    return 0;
}

int dispatch_eval(Ctx *c, EvalScratch *s, uint64_t u_begin, uint64_t u_count, const EvalParams &P)
{
    switch (c->cfg.emb_dim) {
    // item tiles of 128: tcgen05.mma kind::tf32 runs at 59-65 % of its peak with N = 64 and at 99 % with N = 128
    // (heat_probe_tf32_peak variants, profiles/r02_tf32_mma_shapes.md); D = 128 keeps 64 (two B stages of 64 KB)
    case 32: return launch_eval<32, 128>(c, s, u_begin, u_count, P);
    case 64: return launch_eval<64, 128>(c, s, u_begin, u_count, P);
    case 128: return launch_eval<128, 64>(c, s, u_begin, u_count, P);
    default: return fail("evaluation kernel supports emb_dim 32, 64 and 128");
    }
}

// 64-bit fingerprint of a host CSR (every word of both arrays; ~1 ms per 10 MB): an evaluation that is handed the
// same train / test split as the previous one -- the frontend does, every eval_interval epochs -- skips the
// validation, the per-row sort check and the upload
uint64_t csr_fingerprint(const int64_t *indptr, const int32_t *items, uint64_t U)
{
    uint64_t h = 0x9E3779B97F4A7C15ull ^ U;
    for (uint64_t i = 0; i <= U; ++i) h = (h ^ (uint64_t)indptr[i]) * 0x100000001B3ull + (h >> 29);
    const size_t nnz = (size_t)indptr[U];
    size_t i = 0;
    for (; i + 2 <= nnz; i += 2) {
        uint64_t w;
        memcpy(&w, items + i, 8);
        h = (h ^ w) * 0x100000001B3ull + (h >> 29);
    }
    if (i < nnz) h = (h ^ (uint32_t)items[i]) * 0x100000001B3ull + (h >> 29);
    return h ? h : 1;
}

// CSR of items per user -> device, each user's list sorted ascending (the kernel's mask cursor needs it)
int upload_csr(Ctx *c, const int64_t *indptr, const int32_t *items, uint64_t U, long long **d_indptr, size_t *indptr_have,
               int **d_items, size_t *items_have, bool sort_rows, uint64_t *resident_fp)
{
    if (indptr[0] != 0) return fail("CSR indptr must start at 0");
    if (indptr[U] < 0) return fail("CSR indptr must be non-decreasing");
    const uint64_t fp = csr_fingerprint(indptr, items, U);
    if (*resident_fp == fp && *d_indptr && *d_items) return 0;
    *resident_fp = 0;
    const size_t nnz = (size_t)indptr[U];
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
    *resident_fp = fp;
    return 0;
}

int topk_device(Ctx *c, EvalScratch *s, uint32_t k, const int64_t *train_indptr, const int32_t *train_items)
{
    if (!c->user_w) return fail("heat_bind_tables has not been called");
    if (k == 0 || k > 256) return fail("top-k size must be in 1..256");
    const uint64_t U = c->cfg.num_users;
    if (train_indptr) {
        if (!train_items) return fail("train_items is null");
        const uint64_t fp_before = s->train_fp;
        if (int rc = upload_csr(c, train_indptr, train_items, U, &s->train_indptr, &s->indptr_n, &s->train_items, &s->train_n, true, &s->train_fp)) return rc;
        if (s->train_fp != fp_before || !s->perm_ok) {            // a new train CSR: its popularity order
            const size_t I = c->cfg.num_items, nnz = (size_t)train_indptr[U];
            std::vector<uint32_t> deg(I, 0);
            for (size_t i = 0; i < nnz; ++i) {
                if ((uint32_t)train_items[i] >= I) return fail("train_items: id out of range");
                deg[train_items[i]] += 1;
            }
            std::vector<int> perm(I);
            for (size_t i = 0; i < I; ++i) perm[i] = (int)i;
            std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return deg[a] > deg[b]; });
            s->perm_ok = false;
            if (int rc = ensure(&s->perm, &s->perm_elems, I)) return rc;
            HEAT_CUDA(cudaMemcpyAsync(s->perm, perm.data(), I * sizeof(int), cudaMemcpyHostToDevice, c->stream));
            HEAT_CUDA(cudaStreamSynchronize(c->stream));          // `perm` is a host temporary
            s->perm_ok = true;
        }
    }
    if (int rc = ensure(&s->topk_ids, &s->topk_elems, (size_t)U * k)) return rc;
    if (int rc = ensure(&s->topk_scores, &s->topk_scores_elems, (size_t)U * k)) return rc;
    HEAT_CUDA(cudaEventRecord(c->ev0, c->stream));
    // items in popularity order when a train CSR came with the call (HEAT_EVAL_ID_ORDER: experiment switch, id order)
    const bool norm_order = train_indptr && s->perm_ok && std::getenv("HEAT_EVAL_ID_ORDER") == nullptr;
    if (int rc = split_tables(c, s, norm_order)) return rc;
    EvalParams P{};
    P.k = (int)k; P.terms = c->cfg.eval_precision == 1 ? 1 : 3;
    P.perm = norm_order ? s->perm : nullptr;
    P.train_indptr = train_indptr ? s->train_indptr : nullptr;
    P.train_items = train_indptr ? s->train_items : nullptr;
    P.topk_ids = s->topk_ids; P.topk_scores = s->topk_scores;
    if (int rc = dispatch_eval(c, s, 0, U, P)) return rc;
    HEAT_CUDA(cudaEventRecord(c->ev1, c->stream));
    return 0;
}

}  // namespace

#ifdef HEAT_EVAL_STATS
extern "C" void heat_debug_eval_stats(unsigned long long out[4], int reset)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_eval_stats, sizeof(unsigned long long) * 4);
    if (reset) { unsigned long long z[4] = {0, 0, 0, 0}; cudaMemcpyToSymbol(g_eval_stats, z, sizeof(z)); }
}
#endif

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
    if (int rc = upload_csr(c, test_indptr, test_items, U, &s->test_indptr, &s->test_indptr_n, &s->test_items, &s->test_n, false, &s->test_fp)) return rc;
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
    if (int rc = split_tables(c, s, false)) return rc;
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
    cudaFree(s->item_hi); cudaFree(s->item_lo);
    cudaFree(s->topk_ids); cudaFree(s->topk_scores); cudaFree(s->train_indptr); cudaFree(s->train_items);
    cudaFree(s->test_indptr); cudaFree(s->test_items); cudaFree(s->sums); cudaFree(s->dense); cudaFree(s->glist);
    cudaFree(s->perm);
    delete s;
    c->eval_scratch = nullptr;
}

}  // namespace heat
