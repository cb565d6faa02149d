// Exact top-k on the 5th-gen tensor cores: bf16 coarse scan (tcgen05.mma fed by TMA, accumulators in
// TMEM) whose accumulator IS the ranking score, with a fused per-bin-minimum + per-row best-bins
// epilogue kept in registers, followed by an exact fp32 rerank of every row in the selected bins and
// a RIGOROUS certificate.  Queries whose certificate fails are reported so the caller can re-run
// them through the fp32 scan (exact_f32.cu).
//
// Replaces: the all-pairs distance + sort of tests/indexing/test_hnsw.py:128-134 /
// src/vector_db/indexing/ivf.py:176-199 (arithmetic: src/vector_db/util/distance.py:3-4).
//
// Why the answer is still exact.  The reference computes sqrt(sum((q-x)^2)) in fp32.  The coarse
// pass ranks by c(x) = |x|^2 - 2 q~.x~ (q~, x~ = bf16 roundings; |x|^2 enters the MMA as an exact
// three-way bf16 split against ones in the query operand), whose error against the true
// t(x) = |x|^2 - 2 q.x is bounded through the MEASURED rounding residuals of both operands:
//   |q~.x~ - q.x| = |(q~-q).x~ + q.(x~-x)| <= |q~-q| |x~| + |q| |x~-x|        (Cauchy-Schwarz, twice)
//   eps = 2 * (rq * A + |q| * B + 2^-13 |q| A),   rq = |q~-q| (per query, computed from q),
//   A = max |x~|, B = max |x~-x| over the rows (vdb_tc_prepare_rows measures both while it builds the mirror);
// the last term covers the fp32 accumulation of the MMA.  (Round 1 used 2 * 2^-8 |q| max|x|, which covers ONE
// rounding: the worst case of two is (4u + 2u^2)|q||x|, u = 2^-8 -- attained by q = x on bf16 half-way points.
// The residual form is rigorous for any input and, on ordinary data, tighter than either constant because a
// rounding residual is ~0.4 u |x|, not u |x|.)  The kernel keeps, per query, the bins (8 consecutive
// rows) with the smallest bin minimum; every row outside them has c(x) >= tau (see
// tc_select_kernel), hence t(x) >= tau - eps.  All rows inside them are re-scored exactly.  If
// tau - eps > t_k (the k-th exact score) no outside row can enter the top-k: certified.
// Cosine: c(x) = q~ . bf16(-x/|x|), t(x) = -q.x/|x|: the same bound without the factor 2, A and B measured on the
// normalised rows (A ~ 1).
//
// Which bins a list may refuse.  tau is (at most) the worst entry of every FULL per-(slice, column group) list of
// the query, so a bin may be refused by ANY list as soon as its minimum is >= the worst entry of SOME full list of
// the same query: the bound above does not change.  The epilogue therefore tests against the minimum of its own
// threshold, its CTA partner's (shared memory) and every other list's of this launch (one global word per query),
// which keeps most lanes out of the insert code; the hits that remain are parked per lane and inserted by the
// warp together (exact_tc_kernel).
#include <cuda.h>
#include <math.h>

#include "common.cuh"
#include "sm100.cuh"

namespace vdb {

using namespace sm100;

constexpr int kTcThreads = 384;          // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 spare, 4-11 epilogue (two groups)
constexpr int kTileM = 128;              // queries per CTA tile
constexpr int kTileN = 256;              // database rows per accumulator tile
constexpr int kBlockK = 64;              // bf16 elements per 128-byte swizzled row
constexpr int kUmmaK = 16;
constexpr int kABytes = kTileM * kBlockK * 2;    // 16 KB per A k-block
constexpr int kBBytes = kTileN * kBlockK * 2;    // 32 KB per B stage
constexpr int kBin = VDB_TC_BIN;
constexpr int kMaxStages = 8;
constexpr int kNBLMax = 16;              // bins kept per (query, slice, epilogue group): template NBL <= this
constexpr int kNBFinal = 64;  // bins handed to the rerank per query, before trimming
constexpr int kGroups = 2;               // epilogue groups == accumulator stages

constexpr int kA2Bytes = kTileM * 32;            // 4 KB: the L2 norm block of the query tile (16 bf16 per row)
constexpr int kB2Bytes = kTileN * 32;            // 8 KB: the L2 norm block of one row tile

struct TcSmemLayout {
    int a_off, a2_off, b2_off, b_off, bar_off, tau_off, total;
};

// has_norm: the L2 operands carry a 16-column norm block (|x|^2 split / ones) after the D data columns;
// it travels in its own narrow 32-byte-swizzled tiles (4 KB + 2 x 8 KB) instead of a seventh 32 KB stage.
__host__ __device__ inline TcSmemLayout tc_layout(int num_kb, int stages, int has_norm) {
    TcSmemLayout L;
    L.a_off = 0;
    L.a2_off = L.a_off + num_kb * kABytes;
    L.b2_off = L.a2_off + (has_norm ? kA2Bytes : 0);
    L.b_off = L.b2_off + (has_norm ? 2 * kB2Bytes : 0);
    L.bar_off = L.b_off + stages * kBBytes;
    L.tau_off = L.bar_off + (2 * kMaxStages + 12) * 8 + 16;
    L.total = L.tau_off + kGroups * kTileM * 4;                 // threshold exchange between the two epilogue groups
    return L;
}

// Sorted (ascending) register list of the NBL best (score, bin) pairs of one query row.
template <int NBL>
struct BinList {
    float s[NBL];
    uint32_t b[NBL];
    __device__ __forceinline__ void reset() {
#pragma unroll
        for (int i = 0; i < NBL; ++i) { s[i] = INFINITY; b[i] = 0xFFFFFFFFu; }
    }
    __device__ __forceinline__ float tau() const { return s[NBL - 1]; }
    // requires score < tau(): bubble the new pair to its place, the old worst falls off the end
    __device__ __forceinline__ void insert(float score, uint32_t bin) {
#pragma unroll
        for (int i = 0; i < NBL; ++i) {
            const bool lt = score < s[i];
            const float ts = s[i];
            const uint32_t tb = b[i];
            s[i] = lt ? score : ts;
            b[i] = lt ? bin : tb;
            score = lt ? ts : score;
            bin = lt ? tb : bin;
        }
    }
};

__device__ __forceinline__ float min8(const float* v) {
    // three FMNMX3 + one FMNMX (the balanced two-input tree is seven instructions on the half-rate ALU pipe)
    const float a = fminf(fminf(v[0], v[1]), v[2]), b = fminf(fminf(v[3], v[4]), v[5]);
    return fminf(fminf(a, b), fminf(v[6], v[7]));
}

// One 32-column chunk of one accumulator row: the columns ARE the scores (the operand mirror is
// built so that q~ . b~ = |x|^2 - 2 q.x  or  -q.x/|x|).  chunk_minima() reduces the chunk to its four
// bin minima; the caller then compares their minimum with the row's threshold (fast path) and only
// chunks that beat it walk their bins.  The insert code exists ONCE in the kernel (non-unrolled
// loops): the previous fully unrolled form was ~180 KB of SASS and ran out of the instruction cache.
constexpr int kBinsPerChunk = 32 / kBin;
__device__ __forceinline__ float bin_min(const float* v) {
    if (kBin == 8) return min8(v);
    return fminf(fminf(fminf(v[0], v[1]), v[2]), v[3]);        // kBin == 4: FMNMX3 + FMNMX
}
__device__ __forceinline__ void chunk_minima(float (&v)[32], int n_valid, float (&m)[kBinsPerChunk]) {
    if (n_valid < 32) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = (i < n_valid) ? v[i] : INFINITY;
    }
#pragma unroll
    for (int b = 0; b < kBinsPerChunk; ++b) m[b] = bin_min(v + b * kBin);
}
__device__ __forceinline__ float pick_bin(const float (&m)[kBinsPerChunk], int b) {
    float r = m[0];
#pragma unroll
    for (int i = 1; i < kBinsPerChunk; ++i) r = (b == i) ? m[i] : r;
    return r;
}

// Threshold exchange between the two epilogue groups of a CTA (same query row, other half of the columns): a bin
// that cannot enter the PARTNER's full list cannot enter the merged selection either (tc_select_kernel takes
// tau from the worst entry of every full list), so each group tests against min(own tau, partner's tau) and
// inserts ~45 % less often -- and an insert by ONE lane sends its whole warp through the insert code.
// The word carries the unit counter mod 4 in its two low bits (the groups are at most two tiles, hence at most two
// units, apart) and the ordered threshold rounded UP to a multiple of 4: a stale or foreign word reads as +inf.
__device__ __forceinline__ uint32_t tau_word(float tau, int ui) {
    uint32_t u = f32_to_ordered(tau) + 4u;
    if (u > 0xFF800000u) u = 0xFF800000u;                       // +inf stays +inf
    return (u & ~3u) | (uint32_t)(ui & 3);
}
__device__ __forceinline__ float tau_of_word(uint32_t w, int ui) {
    return ((w & 3u) == (uint32_t)(ui & 3)) ? ordered_to_f32(w & ~3u) : INFINITY;
}

// Work unit u -> (query tile mt, slice s, row tiles [t0, t1)).  Brute force: u = s * m_tiles + mt, slices of tiles_per_slice
// row tiles.  LIST-MAJOR IVF (utab != nullptr): entry u of the table = (query tile, first row tile, end row tile) of one
// (inverted list, 128-pair tile); one unit per pair tile, so s = 0 and S = 1.
struct TcUnit { int mt, s, t0, t1; };
__device__ __forceinline__ TcUnit tc_unit(int u, int m_tiles, int tiles_per_slice, int tiles_total, const int4* __restrict__ utab) {
    TcUnit r;
    if (utab) {
        const int4 e = __ldg(utab + u);
        r.mt = e.x; r.s = 0; r.t0 = e.y; r.t1 = e.z;
    } else {
        r.mt = u % m_tiles; r.s = u / m_tiles;
        r.t0 = r.s * tiles_per_slice;
        r.t1 = min(tiles_total, r.t0 + tiles_per_slice);
    }
    return r;
}

// flags: bit 0 = debug, skip the epilogue arithmetic (measures the TMA + MMA pipeline alone)
template <int NBL>
__global__ void __launch_bounds__(kTcThreads, 1)
exact_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                const __grid_constant__ CUtensorMap map_a2, const __grid_constant__ CUtensorMap map_b2,
                int Q, int64_t N, int D, int has_norm, int num_kb, int stages, int m_tiles, int S,
                int tiles_per_slice, int tiles_total, int flags, u64* __restrict__ partial, uint32_t* gtau,
                const int4* __restrict__ utab, const int32_t* __restrict__ units_dev) {
    extern __shared__ __align__(1024) unsigned char smem[];   // 128B-swizzled operand tiles need 1024 B alignment
    if ((smem_u32(smem) & 1023u) != 0) __trap();
    const TcSmemLayout L = tc_layout(num_kb, stages, has_norm);
    unsigned char* sA = smem + L.a_off;
    unsigned char* sA2 = smem + L.a2_off;
    unsigned char* sB2 = smem + L.b2_off;
    unsigned char* sB = smem + L.b_off;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    uint64_t* full = bars;                         // [stages]
    uint64_t* empty = bars + kMaxStages;           // [stages]
    uint64_t* tmem_full = bars + 2 * kMaxStages;   // [2]
    uint64_t* tmem_empty = tmem_full + 2;          // [2]
    uint64_t* a_full = tmem_empty + 2;
    uint64_t* a_empty = a_full + 1;
    uint64_t* n_full = a_empty + 1;                // [2] norm-block ring
    uint64_t* n_empty = n_full + 2;                // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(n_empty + 2);
    uint32_t* sTau = reinterpret_cast<uint32_t*>(smem + L.tau_off);   // [group][row]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int units = units_dev ? __ldg(units_dev) : m_tiles * S;      // list-major IVF: the unit count may live on the device
    // UMMAs in the last k-block: columns past D are zero-filled by TMA, skip the all-zero K steps
    const int last_steps = (D - (num_kb - 1) * kBlockK + kUmmaK - 1) / kUmmaK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_b);
        if (has_norm) { tma_prefetch_desc(&map_a2); tma_prefetch_desc(&map_b2); }
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 256); }
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        for (int i = 0; i < 2; ++i) { mbar_init(&n_full[i], 1); mbar_init(&n_empty[i], 1); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    for (int i = threadIdx.x; i < kGroups * kTileM; i += kTcThreads) sTau[i] = 0xFF800000u;   // +inf (any tag)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ================= TMA producer (one lane) =================
        if (lane == 0) {
            int st = 0, ph = 0, ui = 0, ns = 0, nph = 0;
            for (int u = blockIdx.x; u < units; u += gridDim.x, ++ui) {
                const TcUnit un = tc_unit(u, m_tiles, tiles_per_slice, tiles_total, utab);
                const int mt = un.mt, t0 = un.t0, t1 = un.t1;
                mbar_wait(a_empty, (ui & 1) ^ 1);                 // previous unit's MMAs are done with A
                mbar_expect_tx(a_full, (uint32_t)(num_kb * kABytes + (has_norm ? kA2Bytes : 0)));
                for (int kb = 0; kb < num_kb; ++kb)
                    tma_load_2d(sA + kb * kABytes, &map_a, a_full, kb * kBlockK, mt * kTileM, kEvictLast);
                if (has_norm) tma_load_2d(sA2, &map_a2, a_full, D, mt * kTileM, kEvictLast);
                for (int t = t0; t < t1; ++t) {
                    if (has_norm) {
                        mbar_wait(&n_empty[ns], nph ^ 1);
                        mbar_expect_tx(&n_full[ns], (uint32_t)kB2Bytes);
                        tma_load_2d(sB2 + ns * kB2Bytes, &map_b2, &n_full[ns], 0, t * (kTileN / 4), kEvictNormal);
                        ns ^= 1;
                        if (ns == 0) nph ^= 1;
                    }
                    for (int kb = 0; kb < num_kb; ++kb) {
                        mbar_wait(&empty[st], ph ^ 1);
                        mbar_expect_tx(&full[st], (uint32_t)kBBytes);
                        tma_load_2d(sB + st * kBBytes, &map_b, &full[st], kb * kBlockK, t * kTileN, kEvictNormal);
                        if (++st == stages) { st = 0; ph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer (one lane) =================
        if (lane == 0) {
            const uint32_t idesc = make_idesc_bf16(kTileM, kTileN);
            int st = 0, ph = 0, as = 0, aph = 0, ui = 0, ns = 0, nph = 0;
            for (int u = blockIdx.x; u < units; u += gridDim.x, ++ui) {
                const TcUnit un = tc_unit(u, m_tiles, tiles_per_slice, tiles_total, utab);
                const int t0 = un.t0, t1 = un.t1;
                mbar_wait(a_full, ui & 1);
                tc_fence_after();
                for (int t = t0; t < t1; ++t) {
                    mbar_wait(&tmem_empty[as], aph ^ 1);          // epilogue group `as` drained this accumulator
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + (uint32_t)(as * kTileN);
                    if (has_norm) {                               // |x|^2 first: D = ones . split(|x|^2)
                        mbar_wait(&n_full[ns], nph);
                        tc_fence_after();
                        umma_bf16(d_tmem, make_desc_k_sw32(smem_u32(sA2)), make_desc_k_sw32(smem_u32(sB2 + ns * kB2Bytes)), idesc, 0u);
                        umma_commit(&n_empty[ns]);
                        ns ^= 1;
                        if (ns == 0) nph ^= 1;
                    }
                    for (int kb = 0; kb < num_kb; ++kb) {
                        mbar_wait(&full[st], ph);
                        tc_fence_after();
                        const uint64_t da = make_desc_k_sw128(smem_u32(sA + kb * kABytes));
                        const uint64_t db = make_desc_k_sw128(smem_u32(sB + st * kBBytes));
                        const int steps = (kb == num_kb - 1) ? last_steps : kBlockK / kUmmaK;
                        for (int k = 0; k < steps; ++k)
                            umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((has_norm | kb | k) != 0));
                        umma_commit(&empty[st]);                   // smem stage reusable once these MMAs retire
                        if (++st == stages) { st = 0; ph ^= 1; }
                    }
                    umma_commit(&tmem_full[as]);
                    as ^= 1;
                    if (as == 0) aph ^= 1;
                }
                umma_commit(a_empty);
            }
        }
    } else if (warp >= 4) {
        // ====== epilogue: both groups work on EVERY tile, group eg on columns [eg*128, eg*128+128) ======
        // (The MMA for tile t+2 waits for the epilogue of tile t, so one tile's epilogue must fit in
        // one tile's MMA time; splitting the columns halves it without a third accumulator stage.)
        const int eg = (warp - 4) >> 2;
        const int ew = warp & 3;                       // TMEM lane quarter this warp may touch
        const int row = ew * 32 + lane;                // accumulator row == query within the tile
        constexpr int kCols = kTileN / kGroups;        // columns per group
        const uint32_t tbase = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(eg * kCols);
        BinList<NBL> list;
        volatile uint32_t* my_tau = sTau + eg * kTileM + row;
        const volatile uint32_t* partner_tau = sTau + (eg ^ 1) * kTileM + row;
        int as = 0, aph = 0, ui = 0;
        for (int u = blockIdx.x; u < units; u += gridDim.x, ++ui) {
            const TcUnit un = tc_unit(u, m_tiles, tiles_per_slice, tiles_total, utab);
            const int mt = un.mt, s = un.s, t0 = un.t0, t1 = un.t1;
            list.reset();
            float pend_s = INFINITY;               // parked hit of this lane, not yet in the list
            uint32_t pend_b = 0xFFFFFFFFu;
            float told = INFINITY;                 // own threshold as last published
            *my_tau = tau_word(INFINITY, ui);
            // the same argument holds for EVERY list of the query in this launch (other slices run in other CTAs or
            // ran in an earlier wave): gtau[q] is the ordered minimum of all published list thresholds
            const int qrow = mt * kTileM + row;
            volatile uint32_t* gq = (qrow < Q && !(flags & 8)) ? gtau + qrow : nullptr;
            for (int t = t0; t < t1; ++t) {
                const uint32_t gw = gq ? *gq : 0xFF800000u;      // issued before the wait: the L2 round trip hides under it
                mbar_wait(&tmem_full[as], aph);
                tc_fence_after();
                if (!(flags & 1)) {
                    const float ptau = (flags & 8) ? INFINITY : fminf(tau_of_word(*partner_tau, ui), ordered_to_f32(gw));
                    const uint32_t taddr = tbase + (uint32_t)(as * kTileN);
                    const int64_t left = N - ((int64_t)t * kTileN + eg * kCols);
                    const int n_valid = left < kCols ? (int)left : kCols;      // may be <= 0: every column masked
                    const uint32_t bin_base = (uint32_t)t * (kTileN / kBin) + (uint32_t)(eg * (kCols / kBin));
                    float v[32];
                    tmem_ld_x32(taddr, v);
#pragma unroll 1
                    for (int ch = 0; ch < kCols / 32; ++ch) {
                        float m[kBinsPerChunk];
                        tmem_ld_wait(v);
                        chunk_minima(v, n_valid - ch * 32, m);
                        // v is dead from here on: fetch the next chunk under the threshold test / inserts
                        if (ch + 1 < kCols / 32) tmem_ld_x32(taddr + (uint32_t)((ch + 1) * 32), v);
                        if (flags & 2) continue;
                        float cm = fminf(fminf(m[0], m[1]), fminf(m[2], m[3]));
                        if (kBinsPerChunk == 8) cm = fminf(fminf(cm, fminf(m[4 % kBinsPerChunk], m[5 % kBinsPerChunk])),
                                                           fminf(m[6 % kBinsPerChunk], m[7 % kBinsPerChunk]));
                        // An insert by ONE lane sends the whole warp through the bubble (5 instructions per list slot),
                        // so a lane parks its hit in a one-entry pending slot and the warp inserts only when some
                        // lane's slot is already taken: ~7 hits per bubble instead of one.  (+inf, x) inserts nothing.
                        if (__any_sync(0xFFFFFFFFu, cm < fminf(list.tau(), ptau))) {
#pragma unroll 1
                            for (int b = 0; b < kBinsPerChunk; ++b) {
                                const float mb = pick_bin(m, b);
                                const bool hit = mb < fminf(list.tau(), ptau);
                                if (__any_sync(0xFFFFFFFFu, hit)) {
                                    if (__any_sync(0xFFFFFFFFu, hit && pend_s < INFINITY)) {
                                        list.insert(pend_s, pend_b);
                                        pend_s = INFINITY;
                                    }
                                    if (hit) { pend_s = mb; pend_b = bin_base + (uint32_t)(ch * kBinsPerChunk + b); }
                                }
                            }
                        }
                    }
                    if (list.tau() < told) {
                        told = list.tau();
                        *my_tau = tau_word(told, ui);
                        if (gq) atomicMin(const_cast<uint32_t*>(gq), f32_to_ordered(told));
                    }
                }
                tc_fence_before();
                mbar_arrive(&tmem_empty[as]);
                as ^= 1;
                if (as == 0) aph ^= 1;
            }
            list.insert(pend_s, pend_b);
            const int q = mt * kTileM + row;
            if (q < Q) {
                u64* dst = partial + (((size_t)q * S + s) * kGroups + eg) * NBL;
#pragma unroll
                for (int e = 0; e < NBL; ++e) dst[e] = pack_key(list.s[e], list.b[e]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}


// ------------------------------------------------------------------------------------------
// The same scan on CTA PAIRS (tcgen05 cta_group::2; cosine / no norm block, brute force only) -- built because VERDICT r1
// asked for it, measured, and NOT the default (debug bit 5 of vdb_debug_set_tc_flags selects it; tests keep it correct).
// In a pair each CTA keeps its 128 query rows of A and HALF of every B tile (128 database rows), one MMA of M = 256 covers
// both, and every SM reads half the B bytes from shared memory.  Both CTAs load their halves with TMA (bytes counted on the
// LEADER's barrier), the leader (cluster rank 0) issues the MMAs and commits to the barriers of BOTH CTAs (multicast), each
// CTA's epilogue reads its own 128 TMEM lanes and tells the leader when an accumulator stage is drained.  A unit is (pair
// of query tiles, row slice); everything after the scan is unchanged, and the results are bit-identical.
// Measured (1M x 384, Q = 4096, same process, scripts/dev_tc2.py): step 2.80 ms against 2.42 ms for the single-CTA kernel;
// with the epilogue switched off 2.53 against 2.07 ms.  The premise was wrong: an M = 128, N = 256, K = 16 MMA is 128
// tensor clocks (4096 MAC per clock and SM), so the single-CTA kernel reads 12 KB of operands per 128 clocks = 96 B per
// clock, under the SM's 128, and its MMA-only time of 2.07 ms is already 94 % of what cuBLAS reaches on this GEMM
// (1.94-1.98 ms at the measured 1.59-1.62 PFLOP/s).  The pair form adds a cluster-scope barrier hop to every hand-over
// and halves the number of issuing threads.
constexpr int kBHalfBytes = (kTileN / 2) * kBlockK * 2;      // 16 KB: this CTA's 128 rows of a B k-block
constexpr int kMaxStages2 = 12;
struct Tc2SmemLayout {
    int a_off, b_off, bar_off, tau_off, total;
};
__host__ __device__ inline Tc2SmemLayout tc2_layout(int num_kb, int stages) {
    Tc2SmemLayout L;
    L.a_off = 0;
    L.b_off = num_kb * kABytes;
    L.bar_off = L.b_off + stages * kBHalfBytes;
    L.tau_off = L.bar_off + (2 * kMaxStages2 + 8) * 8 + 16;
    L.total = L.tau_off + kGroups * kTileM * 4;
    return L;
}

template <int NBL>
__global__ void __launch_bounds__(kTcThreads, 1)
exact_tc2_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_bh,
                 int Q, int64_t N, int D, int num_kb, int stages, int m_pairs, int S, int tiles_per_slice, int tiles_total,
                 int flags, u64* __restrict__ partial, uint32_t* gtau) {
    extern __shared__ __align__(1024) unsigned char smem[];
    if ((smem_u32(smem) & 1023u) != 0) __trap();
    const Tc2SmemLayout L = tc2_layout(num_kb, stages);
    unsigned char* sA = smem + L.a_off;
    unsigned char* sB = smem + L.b_off;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    uint64_t* full = bars;                          // [stages]  leader's: both halves of the k-block have landed
    uint64_t* empty = bars + kMaxStages2;           // [stages]  in both CTAs (multicast commit)
    uint64_t* tmem_full = bars + 2 * kMaxStages2;   // [2]       in both CTAs (multicast commit)
    uint64_t* tmem_empty = tmem_full + 2;           // [2]       leader's: 16 warps (8 per CTA) have drained the stage
    uint64_t* a_full = tmem_empty + 2;              //           leader's
    uint64_t* a_empty = a_full + 1;                 //           in both CTAs
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(a_empty + 1);
    uint32_t* sTau = reinterpret_cast<uint32_t*>(smem + L.tau_off);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int cid = (int)(blockIdx.x >> 1), nclusters = (int)(gridDim.x >> 1);
    const int units = m_pairs * S;
    const int last_steps = (D - (num_kb - 1) * kBlockK + kUmmaK - 1) / kUmmaK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_bh);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 16); }
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_2cta(tmem_ptr, 512);
    for (int i = threadIdx.x; i < kGroups * kTileM; i += kTcThreads) sTau[i] = 0xFF800000u;
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                      // the peer's barriers exist before anything arrives on them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ================= TMA producer (one lane, in BOTH CTAs): own A tile, own half of every B k-block =================
        if (lane == 0) {
            int st = 0, ph = 0, ui = 0;
            for (int u = cid; u < units; u += nclusters, ++ui) {
                const int mp = u % m_pairs, s = u / m_pairs;
                const int mt = 2 * mp + (int)rank;
                const int t0 = s * tiles_per_slice, t1 = min(tiles_total, t0 + tiles_per_slice);
                mbar_wait(a_empty, (ui & 1) ^ 1);
                if (rank == 0) mbar_expect_tx(a_full, (uint32_t)(2 * num_kb * kABytes));
                for (int kb = 0; kb < num_kb; ++kb)
                    tma_load_2d_2cta(sA + kb * kABytes, &map_a, a_full, kb * kBlockK, mt * kTileM, kEvictLast);
                for (int t = t0; t < t1; ++t) {
                    for (int kb = 0; kb < num_kb; ++kb) {
                        mbar_wait(&empty[st], ph ^ 1);
                        if (rank == 0) mbar_expect_tx(&full[st], (uint32_t)(2 * kBHalfBytes));
                        tma_load_2d_2cta(sB + st * kBHalfBytes, &map_bh, &full[st], kb * kBlockK,
                                         t * kTileN + (int)rank * (kTileN / 2), kEvictNormal);
                        if (++st == stages) { st = 0; ph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer (one lane of the LEADER) =================
        if (lane == 0 && rank == 0) {
            const uint32_t idesc = make_idesc_bf16(2 * kTileM, kTileN);
            int st = 0, ph = 0, as = 0, aph = 0, ui = 0;
            for (int u = cid; u < units; u += nclusters, ++ui) {
                const int s = u / m_pairs;
                const int t0 = s * tiles_per_slice, t1 = min(tiles_total, t0 + tiles_per_slice);
                mbar_wait(a_full, ui & 1);
                tc_fence_after();
                for (int t = t0; t < t1; ++t) {
                    mbar_wait(&tmem_empty[as], aph ^ 1);          // both CTAs' epilogues drained this accumulator
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + (uint32_t)(as * kTileN);
                    for (int kb = 0; kb < num_kb; ++kb) {
                        mbar_wait(&full[st], ph);
                        tc_fence_after();
                        const uint64_t da = make_desc_k_sw128(smem_u32(sA + kb * kABytes));
                        const uint64_t db = make_desc_k_sw128(smem_u32(sB + st * kBHalfBytes));
                        const int steps = (kb == num_kb - 1) ? last_steps : kBlockK / kUmmaK;
                        for (int k = 0; k < steps; ++k)
                            umma_bf16_2cta(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
                        umma_commit_2cta(&empty[st]);
                        if (++st == stages) { st = 0; ph ^= 1; }
                    }
                    umma_commit_2cta(&tmem_full[as]);
                    as ^= 1;
                    if (as == 0) aph ^= 1;
                }
                umma_commit_2cta(a_empty);
            }
        }
    } else if (warp >= 4) {
        // ====== epilogue (both CTAs, own 128 query rows): as in exact_tc_kernel ======
        const int eg = (warp - 4) >> 2;
        const int ew = warp & 3;
        const int row = ew * 32 + lane;
        constexpr int kCols = kTileN / kGroups;
        const uint32_t tbase = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(eg * kCols);
        BinList<NBL> list;
        volatile uint32_t* my_tau = sTau + eg * kTileM + row;
        const volatile uint32_t* partner_tau = sTau + (eg ^ 1) * kTileM + row;
        int as = 0, aph = 0, ui = 0;
        for (int u = cid; u < units; u += nclusters, ++ui) {
            const int mp = u % m_pairs, s = u / m_pairs;
            const int mt = 2 * mp + (int)rank;
            const int t0 = s * tiles_per_slice, t1 = min(tiles_total, t0 + tiles_per_slice);
            list.reset();
            float pend_s = INFINITY;
            uint32_t pend_b = 0xFFFFFFFFu;
            float told = INFINITY;
            *my_tau = tau_word(INFINITY, ui);
            const int qrow = mt * kTileM + row;
            volatile uint32_t* gq = (qrow < Q && !(flags & 8)) ? gtau + qrow : nullptr;
            for (int t = t0; t < t1; ++t) {
                const uint32_t gw = gq ? *gq : 0xFF800000u;
                mbar_wait(&tmem_full[as], aph);
                tc_fence_after();
                if (!(flags & 1)) {
                    const float ptau = (flags & 8) ? INFINITY : fminf(tau_of_word(*partner_tau, ui), ordered_to_f32(gw));
                    const uint32_t taddr = tbase + (uint32_t)(as * kTileN);
                    const int64_t left = N - ((int64_t)t * kTileN + eg * kCols);
                    const int n_valid = left < kCols ? (int)left : kCols;
                    const uint32_t bin_base = (uint32_t)t * (kTileN / kBin) + (uint32_t)(eg * (kCols / kBin));
                    float v[32];
                    tmem_ld_x32(taddr, v);
#pragma unroll 1
                    for (int ch = 0; ch < kCols / 32; ++ch) {
                        float m[kBinsPerChunk];
                        tmem_ld_wait(v);
                        chunk_minima(v, n_valid - ch * 32, m);
                        if (ch + 1 < kCols / 32) tmem_ld_x32(taddr + (uint32_t)((ch + 1) * 32), v);
                        if (flags & 2) continue;
                        float cm = fminf(fminf(m[0], m[1]), fminf(m[2], m[3]));
                        if (kBinsPerChunk == 8) cm = fminf(fminf(cm, fminf(m[4 % kBinsPerChunk], m[5 % kBinsPerChunk])),
                                                           fminf(m[6 % kBinsPerChunk], m[7 % kBinsPerChunk]));
                        if (__any_sync(0xFFFFFFFFu, cm < fminf(list.tau(), ptau))) {
#pragma unroll 1
                            for (int b = 0; b < kBinsPerChunk; ++b) {
                                const float mb = pick_bin(m, b);
                                const bool hit = mb < fminf(list.tau(), ptau);
                                if (__any_sync(0xFFFFFFFFu, hit)) {
                                    if (__any_sync(0xFFFFFFFFu, hit && pend_s < INFINITY)) {
                                        list.insert(pend_s, pend_b);
                                        pend_s = INFINITY;
                                    }
                                    if (hit) { pend_s = mb; pend_b = bin_base + (uint32_t)(ch * kBinsPerChunk + b); }
                                }
                            }
                        }
                    }
                    if (list.tau() < told) {
                        told = list.tau();
                        *my_tau = tau_word(told, ui);
                        if (gq) atomicMin(const_cast<uint32_t*>(gq), f32_to_ordered(told));
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(&tmem_empty[as], 0u);     // one arrival per warp, on the LEADER's barrier
                as ^= 1;
                if (as == 0) aph ^= 1;
            }
            list.insert(pend_s, pend_b);
            const int q = mt * kTileM + row;
            if (q < Q) {
                u64* dst = partial + (((size_t)q * S + s) * kGroups + eg) * NBL;
#pragma unroll
                for (int e = 0; e < NBL; ++e) dst[e] = pack_key(list.s[e], list.b[e]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                      // the peer may still be reading this CTA's operands / arriving on its barriers
    if (warp == 2) tmem_dealloc_2cta(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------
// helper kernels around the coarse pass

// Query operand: bf16(q) padded to Dp columns; for L2 the three columns after D hold 1.0 so that
// the MMA adds the (hi, mid, lo) bf16 split of |x|^2 stored in the same columns of the row mirror.
// One warp per query row, four columns per lane and step (D % 8 == 0, Dp % 8 == 0): 128-bit loads, 64-bit stores.
// List-major IVF (row_query != nullptr): operand row r is the (query, probed list) pair (row_query[r], row_list[r]) and holds
// the RESIDUAL v = q - centroid[list] (the rows of the list-major mirror are residuals too: |v - r|^2 is the distance, and the
// bf16 error scales with |v||r| instead of |q||x|); columns D..D+2 = 1 (x-norm split), D+3..D+5 = the 3-way bf16 split of
// |v|^2, so the accumulator is the SQUARED DISTANCE, comparable across the pairs of a query.  qstat[r] = (|v|^2, |bf16(v)-v|^2).
__global__ void __launch_bounds__(256)
tc_query_operand_kernel(const float* __restrict__ q, __nv_bfloat16* __restrict__ y, int Q, int D, int Dp,
                        int metric, uint32_t* __restrict__ gtau, int32_t* __restrict__ uncert,
                        int32_t* __restrict__ item_count, const int32_t* __restrict__ row_query,
                        const int32_t* __restrict__ row_list, const float* __restrict__ centroids,
                        float2* __restrict__ qstat) {
    if (blockIdx.x == 0 && threadIdx.x == 0) { uncert[0] = 0; item_count[0] = 0; }   // filled by the post-scan kernels
    const int lane = threadIdx.x & 31;
    const int r = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (r >= Q) return;
    if (lane == 0) gtau[r] = 0xFF800000u;                          // the scan's per-query threshold exchange starts at +inf
    const int rq = row_query ? row_query[r] : r;
    uint2* dst = reinterpret_cast<uint2*>(y + (size_t)r * Dp);
    if (rq < 0) {                                                  // padding of a pair tile
        for (int c4 = lane; c4 < (Dp >> 2); c4 += 32) dst[c4] = make_uint2(0u, 0u);
        if (qstat && lane == 0) qstat[r] = make_float2(0.f, 0.f);
        return;
    }
    const float4* src = reinterpret_cast<const float4*>(q + (size_t)rq * D);
    const float4* cen = row_query ? reinterpret_cast<const float4*>(centroids + (size_t)row_list[r] * D) : nullptr;
    float vsq = 0.f, rsq = 0.f;
    for (int c4 = lane; c4 < (D >> 2); c4 += 32) {
        float4 v = src[c4];
        if (cen) { const float4 c = cen[c4]; v.x -= c.x; v.y -= c.y; v.z -= c.z; v.w -= c.w; }
        const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 pk;
        pk.x = *reinterpret_cast<const uint32_t*>(&lo);
        pk.y = *reinterpret_cast<const uint32_t*>(&hi);
        dst[c4] = pk;
        if (cen) {
            const float r0 = __low2float(lo) - v.x, r1 = __high2float(lo) - v.y, r2 = __low2float(hi) - v.z, r3 = __high2float(hi) - v.w;
            vsq = fmaf(v.x, v.x, vsq); vsq = fmaf(v.y, v.y, vsq); vsq = fmaf(v.z, v.z, vsq); vsq = fmaf(v.w, v.w, vsq);
            rsq = fmaf(r0, r0, rsq); rsq = fmaf(r1, r1, rsq); rsq = fmaf(r2, r2, rsq); rsq = fmaf(r3, r3, rsq);
        }
    }
    if (cen) { vsq = warp_sum(vsq); rsq = warp_sum(rsq); }
    if (metric == VDB_METRIC_L2) {                                 // the 16-column norm block after the D data columns
        for (int c4 = (D >> 2) + lane; c4 < (Dp >> 2); c4 += 32) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c4 == (D >> 2)) {
                v = make_float4(1.f, 1.f, 1.f, 0.f);               // columns D, D+1, D+2 against the split of |x|^2
                if (cen) v.w = __bfloat162float(__float2bfloat16_rn(vsq));                      // column D+3: hi part of |v|^2
            } else if (cen && c4 == (D >> 2) + 1) {
                const float h = __bfloat162float(__float2bfloat16_rn(vsq));
                const float m = __bfloat162float(__float2bfloat16_rn(vsq - h));
                v = make_float4(m, __bfloat162float(__float2bfloat16_rn(vsq - h - m)), 0.f, 0.f);   // columns D+4, D+5
            }
            const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
            uint2 pk;
            pk.x = *reinterpret_cast<const uint32_t*>(&lo);
            pk.y = *reinterpret_cast<const uint32_t*>(&hi);
            dst[c4] = pk;
        }
    }
    if (qstat && lane == 0) qstat[r] = make_float2(vsq, rsq);
}

// Row operand ("mirror"), one warp per row, built once per index mutation:
//   L2     : bf16(-2 x) = -2 bf16(x), row stride D; the bf16 split (hi, mid, lo) of |x|^2 goes to a separate packed
//            norm area [ceil4(N)][16] behind the rows (tc_norm_offset) -- as 16 extra columns of every row it cost 256
//            32-byte TMA row fetches per tile and the TMA unit, not the tensor pipe, set the pace of the L2 scan
//   cosine : bf16(-x/|x|) (zero rows stay zero)
// so that the fp32 accumulator of q~ . row~ is the ranking score itself.  Also emits sqnorm.
__global__ void tc_row_operand_kernel(const float* __restrict__ x, int64_t N, int D, int metric,
                                      float* __restrict__ sqnorm, __nv_bfloat16* __restrict__ mirror,
                                      __nv_bfloat16* __restrict__ norms, float* __restrict__ stats) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t n4 = (N + 3) & ~(int64_t)3;
    float amax = 0.f, bmax = 0.f;                       // squared operand norm / squared rounding residual, max over my rows
    for (int64_t r = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5); r < n4; r += warps) {
        if (r >= N) {                                   // padding rows of the norm area (it is read in 4-row lines)
            if (norms && lane < 16) norms[r * 16 + lane] = __float2bfloat16_rn(0.f);
            continue;
        }
        const float* row = x + r * D;
        float acc = 0.f;
        for (int t = lane; t < D; t += 32) { const float v = row[t]; acc = fmaf(v, v, acc); }
        acc = warp_sum(acc);
        if (sqnorm && lane == 0) sqnorm[r] = acc;
        __nv_bfloat16* out = mirror + r * D;
        // operand value o (x for L2 -- the stored -2 bf16(x) is an exact scaling --, x/|x| for cosine), its bf16
        // rounding o~ and the residual o~ - o, which is exact in fp32
        const float rinv = metric == VDB_METRIC_L2 ? 1.f : (acc > 0.f ? 1.f / sqrtf(acc) : 0.f);
        float a2 = 0.f, b2 = 0.f;
        for (int t = lane; t < D; t += 32) {
            const float o = metric == VDB_METRIC_L2 ? row[t] : row[t] * rinv;
            const __nv_bfloat16 ob = __float2bfloat16_rn(o);
            const float of = __bfloat162float(ob), res = of - o;
            a2 = fmaf(of, of, a2);
            b2 = fmaf(res, res, b2);
            out[t] = metric == VDB_METRIC_L2 ? __float2bfloat16_rn(-2.f * o) : __float2bfloat16_rn(-o);
        }
        if (stats) {
            amax = fmaxf(amax, warp_sum(a2));
            bmax = fmaxf(bmax, warp_sum(b2));
        }
        if (metric == VDB_METRIC_L2 && lane < 16) {
            // 16-element norm row in the byte order the UMMA reads (32-byte swizzle: the two 16-byte halves
            // of row r trade places when bit 2 of r is set), so a 256-row tile is a linear 8 KB copy
            const __nv_bfloat16 h = __float2bfloat16_rn(acc);
            const float r1 = acc - __bfloat162float(h);
            const __nv_bfloat16 m = __float2bfloat16_rn(r1);
            const __nv_bfloat16 l = __float2bfloat16_rn(r1 - __bfloat162float(m));
            // columns 3..5 hold 1.0: the list-major IVF query operand carries the split of |q - c|^2 there (zero for the
            // brute-force scan), which turns the score into the squared distance itself
            const __nv_bfloat16 v = lane == 0 ? h : (lane == 1 ? m : (lane == 2 ? l : __float2bfloat16_rn(lane < 6 ? 1.f : 0.f)));
            norms[r * 16 + (lane ^ ((int)((r >> 2) & 1) << 3))] = v;
        }
    }
    if (stats) {
        // sqrt with a relative allowance for the fp32 summation, max over all warps as raw bits (values >= 0)
        if (lane == 0) {
            atomicMax(reinterpret_cast<unsigned*>(stats), __float_as_uint(sqrtf(amax) * 1.0001f));
            atomicMax(reinterpret_cast<unsigned*>(stats) + 1, __float_as_uint(sqrtf(bmax) * 1.0001f));
        }
    }
}

// The certificate's allowance for one query (file header): eps bounds |coarse score - exact score| for every row,
// slack covers the fp32 rounding of the exact side (rerank arithmetic, |q|^2, the normalisation of cosine rows).
// qn = |q|, rq = |bf16(q) - q|, A = max |x~|, B = max |x~ - x| (stats of vdb_tc_prepare_rows).
__device__ __forceinline__ void tc_bound(int metric, float qn, float rq, float A, float B, float* eps, float* slack) {
    const float e = (rq * A + qn * B) * 1.001f + 1.2207031e-4f * qn * A;          // + 2^-13 |q| A: MMA accumulation
    if (metric == VDB_METRIC_L2) { *eps = 2.f * e; *slack = 4e-6f * (qn + 1.004f * A) * (qn + 1.004f * A); }
    else { *eps = e; *slack = 4e-6f * qn; }
}
__device__ __forceinline__ float bf16_residual(float v) { return __bfloat162float(__float2bfloat16_rn(v)) - v; }

// Post-scan = three small kernels (round 1: select_bins -> per-query rerank -> certify, 0.39 ms of a 2.6 ms step, the
// per-query rerank CTA alternating serial phases with a few loads in flight; a single fused per-query kernel measured
// the same 0.32-0.38 ms for the same reason):
//   tc_select_kernel   one CTA per query: merge the L per-(slice, group) bin lists, keep the NB best bins, trim them,
//                      append the kept bins to ONE flat work list, record tau and the certificate's allowance;
//   tc_rerank_bins_kernel  persistent CTAs stream the flat bin list: every bin is kBin consecutive rows = one contiguous
//                      kBin x 4D-byte run, fetched with a 1-D bulk copy into a shared-memory ring (nothing but loads in
//                      flight, no query-sized imbalance), re-scored with the reference's fp32 arithmetic (the operation
//                      order of rerank_f32_kernel / the fp32 scan, so the three agree bit for bit);
//   tc_final_kernel    one warp per query: k best of the query's keys by (score, row), then the certificate:
//                      every bin that is NOT a candidate has a coarse minimum >= tau (it ranks behind the NB-th merged
//                      bin, or it was dropped from a full list whose worst entry bounds it, or it was trimmed); the query
//                      is exact iff tau - eps - slack > t_k.  Rejected queries go to the list the fp32 rescan reads.
struct TcPostQuery {          // per query, written by tc_select_kernel
    float tau_minus;          // tau - eps - slack
    float qsq;                // |q|^2 in rerank_f32_kernel's summation order
    int keep;                 // bins kept
    int pad;
};

// ONE WARP per query (the CTA-per-query rank search of r01 was 12.5 k warp instructions per query, ALU-bound at 0.07 ms).
// The kept bins need no order -- the rerank re-scores all of them -- so only c_k, the k-th smallest bin minimum, has to be
// FOUND: k rounds of "smallest remaining score" (lane-strided scan of the query's L x NBL keys in shared memory, 32-bit
// butterfly minimum, the lowest lane holding it marks one of its entries).  Then every key is tested against the cutoff
// c_k + 2 eps + 4 slack in parallel and the survivors are compacted into the flat work list.
//   tau = min( scores of the keys NOT kept, worst entry of every FULL list ): a bin that is not a candidate either is one
//   of the keys not kept or was dropped from a full list, whose worst entry (kept or not) bounds it.
// More than NB survivors (never seen on real data: the trim keeps ~20 bins): the extraction simply runs on to NB rounds.
constexpr int kSelWarps = 4;
__global__ void __launch_bounds__(32 * kSelWarps)
tc_select_kernel(const u64* __restrict__ partial, int L, int NBL, int NB, int C, const float* __restrict__ qv, int D, int k,
                 int metric, const float* __restrict__ stats, int no_trim, int Q, TcPostQuery* __restrict__ pq,
                 uint2* __restrict__ items, int32_t* __restrict__ item_count, const int32_t* __restrict__ pair_rows, int P,
                 const float2* __restrict__ qstat, const int32_t* __restrict__ row_list, const float2* __restrict__ list_stats) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int q = (int)blockIdx.x * kSelWarps + warp;
    if (q >= Q) return;
    u64* buf = reinterpret_cast<u64*>(smem_raw) + (size_t)warp * C;       // [C] keys; entry i belongs to lane i % 32
    // brute force: the query's L lists are contiguous.  List-major IVF: operand row pair_rows[q][p] holds the lists of the
    // query's p-th probed list (L lists each); -1 = the probed list is empty
    const int per_row = L * NBL;
    const int total = (pair_rows ? P : 1) * per_row;                       // <= C <= 1024: at most 32 entries per lane
    float a = 0.f, b = 0.f;                                // |q|^2 in the order rerank_f32_kernel uses, and |bf16(q) - q|^2
    float eps = 0.f, slack = 0.f;                          // the certificate's allowance, also the trim's
    float eps_pair = 0.f;                                  // list-major IVF: lane p = allowance of the query's p-th pair
    if (qstat) {
        // list-major IVF: the operand rows are residuals q - centroid, one per probed list, and every probed list has its
        // own row bounds (list_stats, else the global ones), so every PAIR has its own allowance eps_p.  The keys of pair p
        // are lowered by eps_p when they are loaded: from there on every score is a LOWER BOUND of the bin's true minimum,
        // and the certificate needs no further eps (a far, wide list no longer loosens the bound of the near ones).
        for (int p = lane; p < P; p += 32) {
            const int row = pair_rows[(size_t)q * P + p];
            if (row >= 0) {
                const float2 st = qstat[row];
                float A = stats[0], B = stats[1];
                if (list_stats) { const float2 ls = list_stats[row_list[row]]; A = fminf(A, ls.x); B = fminf(B, ls.y); }
                float s;
                tc_bound(metric, sqrtf(st.x) * 1.0001f, sqrtf(st.y) * 1.0001f, A, B, &eps_pair, &s);
                slack = fmaxf(slack, s);
            }
        }
        eps = eps_pair;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            eps = fmaxf(eps, __shfl_xor_sync(0xffffffffu, eps, o));
            slack = fmaxf(slack, __shfl_xor_sync(0xffffffffu, slack, o));
        }
    } else {
        for (int t = lane; t < D; t += 32) {
            const float v = qv[(size_t)q * D + t], r = bf16_residual(v);
            a = fmaf(v, v, a);
            b = fmaf(r, r, b);
        }
        a = warp_sum(a);
        b = warp_sum(b);
        tc_bound(metric, sqrtf(a) * 1.0001f, sqrtf(b) * 1.0001f, stats[0], stats[1], &eps, &slack);
    }
    unsigned full_worst = f32_to_ordered(INFINITY);        // ordered-float minimum over the full lists' worst entries
    for (int i = lane; i < C; i += 32) {
        u64 key = VDB_KEY_MAX;
        int pi = 0;
        if (i < total) {
            int64_t row = q;
            int within = i;
            if (pair_rows) {
                pi = i / per_row;
                within = i - pi * per_row;
                row = pair_rows[(size_t)q * P + pi];
            }
            if (row >= 0) key = partial[(size_t)row * per_row + within];
        }
        const float ep = __shfl_sync(0xffffffffu, eps_pair, pi);                   // 0 in the brute-force scan
        if (key_index(key) == 0xFFFFFFFFu) key = VDB_KEY_MAX;                      // empty slot
        else {
            if (pair_rows)
                key = ((u64)f32_to_ordered(ordered_to_f32((unsigned)(key >> 32)) - ep) << 32) | (key & 0xFFFFFFFFull);
            if ((i % NBL) == NBL - 1) full_worst = min(full_worst, (unsigned)(key >> 32));
        }
        buf[i] = key;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) full_worst = min(full_worst, __shfl_xor_sync(0xffffffffu, full_worst, o));
    __syncwarp();
    // ---- rounds: the smallest remaining ordered score; `taken` = this lane's entries already extracted (bit t: entry lane + 32 t)
    uint32_t taken = 0u;
    unsigned last = 0xFFFFFFFFu;                                     // score of the last extracted key (ordered bits)
    int n_ext = 0;
    auto extract = [&]() -> bool {                                   // one round; false when nothing is left
        unsigned best = 0xFFFFFFFFu;
        int bt = -1;
        for (int i = lane, t = 0; i < total; i += 32, ++t) {
            const unsigned sc = (unsigned)(buf[i] >> 32);
            if (!((taken >> t) & 1u) && sc < best) { best = sc; bt = t; }
        }
        unsigned wbest = best;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) wbest = min(wbest, __shfl_xor_sync(0xffffffffu, wbest, o));
        if (wbest == 0xFFFFFFFFu) return false;                      // only empty slots (score bits of VDB_KEY_MAX) are left
        const unsigned owners = __ballot_sync(0xffffffffu, best == wbest && bt >= 0);
        if (lane == __ffs(owners) - 1) taken |= 1u << bt;
        last = wbest;
        ++n_ext;
        return true;
    };
    const int rounds = k < NB ? k : NB;
    while (n_ext < rounds && extract()) {}
    float cutoff = INFINITY;
    if (!no_trim && n_ext == k) cutoff = ordered_to_f32(last) + 2.f * eps + 4.f * slack;
    // ---- survivors: every key with score <= cutoff (the extracted k are among them)
    auto count_kept = [&](unsigned cut_bits, bool use_taken) {
        int c = 0;
        for (int i = lane, t = 0; i < total; i += 32, ++t) {
            const unsigned sc = (unsigned)(buf[i] >> 32);
            const bool kept = use_taken ? ((taken >> t) & 1u) != 0u : (sc != 0xFFFFFFFFu && sc <= cut_bits);
            c += kept ? 1 : 0;
        }
        return c;
    };
    unsigned cut_bits = cutoff == INFINITY ? 0xFFFFFFFEu : f32_to_ordered(cutoff);
    bool by_mask = false;                                            // survivors named by `taken` instead of the cutoff
    int mine = count_kept(cut_bits, false), keep = mine;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) keep += __shfl_xor_sync(0xffffffffu, keep, o);
    if (keep > NB) {                                                 // rare: take exactly the NB smallest
        while (n_ext < NB && extract()) {}
        by_mask = true;
        mine = count_kept(0u, true);
        keep = n_ext;
    }
    // tau: the smallest score that is not kept, and the full lists' worst entries
    unsigned tmin = full_worst;
    for (int i = lane, t = 0; i < total; i += 32, ++t) {
        const unsigned sc = (unsigned)(buf[i] >> 32);
        const bool kept = by_mask ? ((taken >> t) & 1u) != 0u : (sc != 0xFFFFFFFFu && sc <= cut_bits);
        if (!kept) tmin = min(tmin, sc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tmin = min(tmin, __shfl_xor_sync(0xffffffffu, tmin, o));
    const float tau = tmin == 0xFFFFFFFFu ? INFINITY : ordered_to_f32(tmin);
    // ---- compaction into the flat work list: exclusive prefix of the per-lane counts
    int pre = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, pre, o);
        if (lane >= o) pre += v;
    }
    pre -= mine;
    int base = 0;
    if (lane == 0) {
        base = keep > 0 ? atomicAdd(item_count, keep) : 0;
        TcPostQuery o;
        o.tau_minus = qstat ? tau - slack : tau - eps - slack;       // IVF keys are lower bounds already
        o.qsq = qstat ? 0.f : a;                                 // IVF scores are squared distances: nothing to subtract
        o.keep = keep;
        o.pad = 0;
        pq[q] = o;
    }
    base = __shfl_sync(0xffffffffu, base, 0);
    int j = pre;
    for (int i = lane, t = 0; i < total; i += 32, ++t) {
        const u64 key = buf[i];
        const unsigned sc = (unsigned)(key >> 32);
        const bool kept = by_mask ? ((taken >> t) & 1u) != 0u : (sc != 0xFFFFFFFFu && sc <= cut_bits);
        if (kept) {
            items[base + j] = make_uint2((uint32_t)q * (uint32_t)kNBFinal + (uint32_t)j, key_index(key));
            ++j;
        }
    }
}

constexpr int kRrMaxWarps = 8;             // warps per CTA, each with its own two-slot ring and its own work items
constexpr int kRrSlots = 2;

__host__ __device__ inline size_t rr_slot_bytes(int D) { return (size_t)(kBin + 1) * D * 4; }   // kBin rows + the query

// Every WARP streams its own work items (bin = kBin consecutive rows = one contiguous run, plus the bin's query) through
// a private two-slot ring filled by 1-D bulk copies; no CTA barrier anywhere.  (A CTA-wide ring with one warp per row of
// a bin ran at 3.7 TB/s: eight warps in lockstep paid the shuffle / global-load latency of every bin in sequence.)
template <int METRIC>
__global__ void __launch_bounds__(32 * kRrMaxWarps)
tc_rerank_bins_kernel(const uint2* __restrict__ items, const int32_t* __restrict__ item_count, const float* __restrict__ qv,
                      const float* __restrict__ db, int64_t N, int D, const TcPostQuery* __restrict__ pq,
                      u64* __restrict__ keys) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_s[kRrMaxWarps * kRrSlots];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const size_t slot_floats = (size_t)(kBin + 1) * D;
    float* ring = reinterpret_cast<float*>(smem_raw) + (size_t)warp * kRrSlots * slot_floats;
    uint64_t* full = full_s + warp * kRrSlots;
    const int total = *item_count;
    if (lane == 0) {
        for (int i = 0; i < kRrSlots; ++i) mbar_init(&full[i], 1);
        fence_barrier_init();
    }
    __syncwarp();
    const int nchunk = D >> 2;
    const int gw = (int)blockIdx.x * nwarp + warp, gstride = (int)gridDim.x * nwarp;   // item n of this warp = gw + n * gstride
    const int mine = total > gw ? (total - gw + gstride - 1) / gstride : 0;
    uint2 it_l = make_uint2(0u, 0u);                                  // lane l: item (blk + l) of the current 32-item block
    float qsq_l = 0.f;
    auto load_block = [&](int blk) {                                  // descriptors of items blk .. blk + 31, one per lane
        if (blk + lane < mine) {
            it_l = items[(size_t)gw + (size_t)(blk + lane) * gstride];
            qsq_l = pq[it_l.x / kNBFinal].qsq;
        }
    };
    auto issue = [&](uint2 it, int n) {                               // lane 0 only
        const int sl = n % kRrSlots;
        const int64_t r0 = (int64_t)it.y * kBin;
        const int64_t rows = N - r0 < kBin ? N - r0 : kBin;
        const uint32_t bytes = (uint32_t)rows * (uint32_t)D * 4u;
        float* dst = ring + (size_t)sl * slot_floats;
        mbar_expect_tx(&full[sl], bytes + (uint32_t)D * 4u);
        bulk_load_1d(dst, db + (size_t)r0 * D, bytes, &full[sl]);
        bulk_load_1d(dst + (size_t)kBin * D, qv + (size_t)(it.x / kNBFinal) * D, (uint32_t)D * 4u, &full[sl]);
    };
    load_block(0);
    for (int n = 0; n < kRrSlots && n < mine; ++n) {
        const uint2 it = make_uint2(__shfl_sync(0xffffffffu, it_l.x, n), __shfl_sync(0xffffffffu, it_l.y, n));
        if (lane == 0) issue(it, n);
    }
    uint2 nx_l = it_l;                                                // descriptors of the NEXT block (prefetched)
    float nqsq_l = qsq_l;
    for (int n = 0; n < mine; ++n) {
        const int j = n & 31;
        if (j == 0 && n > 0) { it_l = nx_l; qsq_l = nqsq_l; }
        if (j == 0 && n + 32 < mine) {                                // prefetch the block after this one
            const uint2 keep_it = it_l;
            const float keep_q = qsq_l;
            load_block(n + 32);
            nx_l = it_l; nqsq_l = qsq_l;
            it_l = keep_it; qsq_l = keep_q;
        }
        const uint2 it = make_uint2(__shfl_sync(0xffffffffu, it_l.x, j), __shfl_sync(0xffffffffu, it_l.y, j));
        const float qsq = __shfl_sync(0xffffffffu, qsq_l, j);
        const int sl = n % kRrSlots;
        mbar_wait(&full[sl], (uint32_t)((n / kRrSlots) & 1));
        const float* base = ring + (size_t)sl * slot_floats;
        const float4* q4 = reinterpret_cast<const float4*>(base + (size_t)kBin * D);
        float acc[kBin], xx[kBin];
#pragma unroll
        for (int r = 0; r < kBin; ++r) { acc[r] = 0.f; xx[r] = 0.f; }
        // rerank_f32_kernel's operation order per row: chunk t = lane, lane + 32, ...; four fmas per chunk; butterfly
        for (int t = lane; t < nchunk; t += 32) {
            const float4 qv4 = q4[t];
#pragma unroll
            for (int r = 0; r < kBin; ++r) {
                const float4 v = reinterpret_cast<const float4*>(base + (size_t)r * D)[t];
                if (METRIC == VDB_METRIC_L2) {
                    const float d0 = qv4.x - v.x, d1 = qv4.y - v.y, d2 = qv4.z - v.z, d3 = qv4.w - v.w;
                    acc[r] = fmaf(d0, d0, acc[r]); acc[r] = fmaf(d1, d1, acc[r]);
                    acc[r] = fmaf(d2, d2, acc[r]); acc[r] = fmaf(d3, d3, acc[r]);
                } else {
                    acc[r] = fmaf(qv4.x, v.x, acc[r]); acc[r] = fmaf(qv4.y, v.y, acc[r]);
                    acc[r] = fmaf(qv4.z, v.z, acc[r]); acc[r] = fmaf(qv4.w, v.w, acc[r]);
                    xx[r] = fmaf(v.x, v.x, xx[r]); xx[r] = fmaf(v.y, v.y, xx[r]);
                    xx[r] = fmaf(v.z, v.z, xx[r]); xx[r] = fmaf(v.w, v.w, xx[r]);
                }
            }
        }
        __syncwarp();                                                 // every lane is done reading the slot
        // refill the slot before the reductions: the descriptor of item n + kRrSlots comes from the lane that holds it
        // (current or prefetched block)
        {
            const int m = n + kRrSlots;
            const int jm = m & 31;
            const bool in_next = (m >> 5) != (n >> 5);
            const uint32_t sx = in_next ? nx_l.x : it_l.x, sy = in_next ? nx_l.y : it_l.y;
            const uint2 itn = make_uint2(__shfl_sync(0xffffffffu, sx, jm), __shfl_sync(0xffffffffu, sy, jm));
            if (lane == 0 && m < mine) issue(itn, m);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int r = 0; r < kBin; ++r) {
                acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], o);
                if (METRIC == VDB_METRIC_COSINE) xx[r] += __shfl_xor_sync(0xffffffffu, xx[r], o);
            }
        }
        float a_mine = 0.f, x_mine = 1.f;
#pragma unroll
        for (int r = 0; r < kBin; ++r)
            if (lane == r) { a_mine = acc[r]; x_mine = xx[r]; }
        if (lane < kBin) {
            const int64_t row = (int64_t)it.y * kBin + lane;
            u64 key = VDB_KEY_MAX;
            if (row < N) {
                const float score = (METRIC == VDB_METRIC_L2) ? sqrtf(a_mine) : -(a_mine / (sqrtf(x_mine) * sqrtf(qsq)));
                key = pack_key(score, (uint32_t)row);
            }
            keys[(size_t)it.x * kBin + lane] = key;                   // it.x = q * kNBFinal + j
        }
    }
}

// One warp per query: the k best keys of keep x kBin, outputs, certificate.
template <int METRIC>
__global__ void __launch_bounds__(256)
tc_final_kernel(const u64* __restrict__ keys, const TcPostQuery* __restrict__ pq, int Q, int k, const int64_t* __restrict__ ids,
                float* __restrict__ out_d, int64_t* __restrict__ out_i, int32_t* __restrict__ uncert) {
    const int lane = threadIdx.x & 31;
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (q >= Q) return;
    const TcPostQuery info = pq[q];
    const int R = info.keep * kBin;
    constexpr int PER = kNBFinal * kBin / 32;              // keys per lane (16)
    u64 mine[PER];
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int s = i * 32 + lane;
        mine[i] = s < R ? keys[(size_t)q * kNBFinal * kBin + s] : VDB_KEY_MAX;
    }
    u64 my_out = VDB_KEY_MAX;                               // lane r ends up with the r-th best key
    u64 kth = VDB_KEY_MAX;
    int got = 0;
    for (int r = 0; r < k; ++r) {
        u64 best = mine[0];
#pragma unroll
        for (int i = 1; i < PER; ++i) best = mine[i] < best ? mine[i] : best;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other < best ? other : best;
        }
        if (best == VDB_KEY_MAX) break;
#pragma unroll
        for (int i = 0; i < PER; ++i) mine[i] = mine[i] == best ? VDB_KEY_MAX : mine[i];   // keys are unique (score, row)
        if (lane == r) my_out = best;
        kth = best;
        ++got;
    }
    if (lane < k) {
        if (my_out != VDB_KEY_MAX) {
            const float sc = key_score(my_out);
            out_d[(size_t)q * k + lane] = (METRIC == VDB_METRIC_L2) ? sc : -sc;
            const int64_t idx = (int64_t)key_index(my_out);
            out_i[(size_t)q * k + lane] = ids ? ids[idx] : idx;
        } else {
            out_d[(size_t)q * k + lane] = (METRIC == VDB_METRIC_L2) ? INFINITY : -INFINITY;
            out_i[(size_t)q * k + lane] = -1;
        }
    }
    if (lane == 0) {
        bool ok;
        if (got < k) {
            ok = (info.tau_minus == INFINITY);                       // fewer than k rows exist: all must have been reranked
        } else if (METRIC == VDB_METRIC_L2) {
            const float dk = key_score(kth);
            ok = info.tau_minus > dk * dk - info.qsq;
        } else {
            const float simk = -key_score(kth);
            ok = info.tau_minus > -simk * sqrtf(info.qsq);
        }
        if (!ok) {
            const int slot = atomicAdd(uncert, 1);
            uncert[1 + slot] = q;
        }
    }
}

}  // namespace vdb

// ------------------------------------------------------------------------------------------
// host side
namespace vdb {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// bf16 row-major matrix with `cols` visible columns out of rows of `row_elems` elements; box = box_cols x
// box_rows with the matching swizzle (64 columns -> 128 B, 16 columns -> 32 B); zero fill out of bounds
static int make_map(CUtensorMap* map, const void* base, int64_t rows, int cols, int row_elems, int box_cols, int box_rows,
                    int swizzle = -1) {
    EncodeTiledFn fn = encode_tiled_fn();
    VDB_REQUIRE(fn != nullptr, VDB_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)row_elems * 2};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE,
                    swizzle == 0 ? CU_TENSOR_MAP_SWIZZLE_NONE : (box_cols == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_128B),
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    VDB_REQUIRE(r == CUDA_SUCCESS, VDB_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return VDB_OK;
}

struct TcPlan {
    int Dp, has_norm, num_kb, stages, m_tiles, S, tiles_per_slice, tiles_total, grid;
    size_t smem;
    int pair, m_pairs;             // pair = 1: exact_tc2_kernel (CTA pairs); units = m_pairs * S, grid = 2 * clusters
};

static int tc_row_elems(int D, int metric) { return metric == VDB_METRIC_L2 ? D + 16 : D; }   // QUERY operand row
// element offset of the packed norm area behind the N x D operand rows (128-byte aligned), and the buffer size
static size_t tc_norm_offset(int64_t N, int D) { return ((size_t)N * D + 63) & ~(size_t)63; }
static size_t tc_mirror_elems(int64_t N, int D, int metric) {
    return tc_norm_offset(N, D) + (metric == VDB_METRIC_L2 ? (size_t)((N + 3) & ~(int64_t)3) * 16 : 0);
}

static int g_tc_flags = 0;

static bool plan_tc(int Q, int64_t N, int D, int k, int metric, TcPlan* p) {
    if (D % 8 != 0 || D <= 0 || Q <= 0 || N <= 0) return false;
    if (k > 16) return false;                       // kNBFinal bins of kBin rows must hold k results with margin
    p->Dp = tc_row_elems(D, metric);
    p->has_norm = metric == VDB_METRIC_L2 ? 1 : 0;
    p->num_kb = (D + kBlockK - 1) / kBlockK;
    const int smem_max = max_smem_optin();
    int stages = kMaxStages;
    if ((g_tc_flags >> 8) & 0xF) stages = (g_tc_flags >> 8) & 0xF;       // debug: cap the TMA ring depth
    // + 1024: the 1024-byte aligned extern array makes the compiler reserve 1 KB of static shared memory
    while (stages >= 2 && tc_layout(p->num_kb, stages, p->has_norm).total + 1024 > smem_max) --stages;
    if (stages < 2) return false;
    p->stages = stages;
    p->smem = (size_t)tc_layout(p->num_kb, stages, p->has_norm).total;
    p->m_tiles = (Q + kTileM - 1) / kTileM;
    int64_t tiles_total = (N + kTileN - 1) / kTileN;
    if (tiles_total > (1 << 24)) return false;
    p->tiles_total = (int)tiles_total;
    const int sms = sm_count();
    // Number of slices: minimise waves x time per unit.  A unit's time (in tile-MMA units) is the
    // larger of its MMA time (tps) and its epilogue time: a per-tile fast path plus the bin-list
    // inserts, whose count per row grows only with log(tps) -- short units are insert-bound
    // (calibrated on B200: 1M x 384, S = 37 ran 254 tile-times per 106-tile unit).
    double best_cost = -1.0;
    int best_S = 1;
    for (int S = 1; S <= 64; ++S) {
        int tps = (p->tiles_total + S - 1) / S;
        if (S > 1 && tps < 8) break;
        int S_eff = (p->tiles_total + tps - 1) / tps;
        long long units = (long long)p->m_tiles * S_eff;
        long long grid = units < sms ? units : sms;
        long long waves = (units + grid - 1) / grid;
        double epi = 0.3 * tps + 39.0 * (log((double)tps) + 1.0);
        double unit = (tps > epi ? (double)tps : epi) + 2.0;     // +2: per-unit A load / output overhead
        double cost = (double)waves * unit;
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_S = S; }
    }
    p->tiles_per_slice = (p->tiles_total + best_S - 1) / best_S;
    p->S = (p->tiles_total + p->tiles_per_slice - 1) / p->tiles_per_slice;
    long long units = (long long)p->m_tiles * p->S;
    p->grid = (int)(units < sms ? units : sms);
    p->pair = 0;
    p->m_pairs = (p->m_tiles + 1) / 2;
    // CTA pairs (cosine, at least two query tiles): OFF unless debug bit 5 is set -- measured slower, see exact_tc2_kernel.
    // Same cost model over pair units and sms / 2 clusters.
    if (!p->has_norm && p->m_tiles >= 2 && (g_tc_flags & 32) && sms >= 2) {
        int st2 = kMaxStages2;
        if ((g_tc_flags >> 8) & 0xF) st2 = (g_tc_flags >> 8) & 0xF;
        while (st2 >= 2 && tc2_layout(p->num_kb, st2).total + 1024 > smem_max) --st2;
        if (st2 >= 2) {
            const int clusters = sms / 2;
            double bc = -1.0;
            int bS = 1;
            for (int S = 1; S <= 64; ++S) {
                int tps = (p->tiles_total + S - 1) / S;
                if (S > 1 && tps < 8) break;
                int S_eff = (p->tiles_total + tps - 1) / tps;
                long long un = (long long)p->m_pairs * S_eff;
                long long g = un < clusters ? un : clusters;
                long long waves = (un + g - 1) / g;
                double epi = 0.3 * tps + 39.0 * (log((double)tps) + 1.0);
                double unit = (tps > epi ? (double)tps : epi) + 2.0;
                double cost = (double)waves * unit;
                if (bc < 0 || cost < bc) { bc = cost; bS = S; }
            }
            p->pair = 1;
            p->stages = st2;
            p->smem = (size_t)tc2_layout(p->num_kb, st2).total;
            p->tiles_per_slice = (p->tiles_total + bS - 1) / bS;
            p->S = (p->tiles_total + p->tiles_per_slice - 1) / p->tiles_per_slice;
            long long un = (long long)p->m_pairs * p->S;
            p->grid = 2 * (int)(un < clusters ? un : clusters);
        }
    }
    return true;
}

struct TcWorkspace {
    size_t qbf16, partial, tau, qstat, pq, items, keys, counters, rescan, rescan_bytes, total;
};

// Qa = rows of the query operand (= Q for the brute-force scan; pair-tile rows for list-major IVF, which has no fp32 rescan)
static TcWorkspace tc_workspace(const TcPlan& p, int Q, int64_t Qa, int64_t N, int D, int k, bool rescan) {
    TcWorkspace w;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    w.qbf16 = take((size_t)Qa * p.Dp * 2);
    w.partial = take((size_t)Qa * p.S * kGroups * kNBLMax * 8);
    w.tau = take((size_t)Qa * 4);
    w.qstat = take(rescan ? 0 : (size_t)Qa * 8);             // list-major IVF: (|v|^2, |bf16(v) - v|^2) per operand row
    w.pq = take((size_t)Q * sizeof(TcPostQuery));
    w.items = take((size_t)Q * kNBFinal * sizeof(uint2));
    w.keys = take((size_t)Q * kNBFinal * kBin * 8);
    w.counters = take(256);
    w.rescan_bytes = rescan ? exact_f32_subset_workspace(Q, N, D, k) : 0;      // fp32 rescan of uncertified queries
    w.rescan = take(w.rescan_bytes);
    w.total = off;
    return w;
}

// List-major IVF on the same pipeline (ivf.py:170-199 for a batch): the database operand is the list-major copy of the rows
// (every list padded to whole 256-row tiles), the query operand holds one row per (query, probed list) pair, pairs of a
// list padded to whole 128-row tiles, and a unit table names (pair tile, row tiles of its list).
struct TcIvf {
    int n_probe;                   // probed lists per query (<= 32)
    const int32_t* pair_rows;      // [Q][n_probe] operand row of the query's p-th probe, -1 = empty list
    const int32_t* row_query;      // [Qa] query of an operand row, -1 = padding
    const int32_t* row_list;       // [Qa] inverted list of an operand row
    const float* centroids;        // [K][D]: operand rows and mirror rows are residuals against the list's centroid
    int64_t Qa;                    // operand rows (multiple of 128)
    const int4* utab;              // [n_units] (pair tile, first row tile, end row tile, 0)
    int n_units;                   // units, or their upper bound when units_dev is given
    const int32_t* units_dev;      // device-side unit count (vdb_ivf_group_pairs' meta + 1), or null
    const float2* list_stats;      // [K] per list (max |bf16(x - c)|, max |bf16(x - c) - (x - c)|), or null: the global pair
};

static bool plan_tc_ivf(const TcIvf& v, int64_t N, int D, int k, TcPlan* p) {
    if (!plan_tc((int)(v.Qa < (1 << 30) ? v.Qa : (1 << 30)), N, D, k, VDB_METRIC_L2, p)) return false;
    p->m_tiles = v.n_units;        // units = m_tiles * S; tc_unit() reads the table instead of decoding u
    p->S = 1;
    p->tiles_per_slice = p->tiles_total;
    const int sms = sm_count();
    p->grid = v.n_units < sms ? v.n_units : sms;
    return true;
}

static int tc_search_impl(const float* q, int Q, const float* db, const void* mirror, const float* db_stats,
                          const int64_t* ids, int64_t N, int D, int k, int metric, float* out_d, int64_t* out_i,
                          int32_t* out_uncertified, void* ws, size_t ws_bytes, void* stream, const TcIvf* ivf);

}  // namespace vdb

using namespace vdb;

extern "C" {

void vdb_debug_set_tc_flags(int flags) { g_tc_flags = flags; }

int vdb_exact_topk_tc_supported(int Q, int64_t N, int D, int k, int metric) {
    TcPlan p;
    return plan_tc(Q, N, D, k, metric, &p) && N < (1ll << 31) ? 1 : 0;
}

size_t vdb_tc_mirror_elems(int64_t N, int D, int metric) { return tc_mirror_elems(N, D, metric); }

int vdb_tc_prepare_rows(const float* vecs, int64_t N, int D, int metric, float* sqnorm, void* mirror, float* stats,
                        void* stream) {
    VDB_REQUIRE(N >= 0 && D > 0 && D % 8 == 0, VDB_ERR_INVALID, "vdb_tc_prepare_rows: bad N=%lld D=%d", (long long)N, D);
    VDB_REQUIRE(metric == VDB_METRIC_L2 || metric == VDB_METRIC_COSINE, VDB_ERR_INVALID, "unknown metric %d", metric);
    if (stats) VDB_CUDA(cudaMemsetAsync(stats, 0, 2 * sizeof(float), static_cast<cudaStream_t>(stream)));
    if (N == 0) return VDB_OK;
    VDB_REQUIRE(vecs && mirror, VDB_ERR_INVALID, "vdb_tc_prepare_rows: null pointer");
    int blocks = sm_count() * 8;
    int64_t need = (N + 7) / 8;
    if (need < blocks) blocks = (int)need;
    tc_row_operand_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        vecs, N, D, metric, sqnorm, static_cast<__nv_bfloat16*>(mirror),
        metric == VDB_METRIC_L2 ? static_cast<__nv_bfloat16*>(mirror) + tc_norm_offset(N, D) : nullptr, stats);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

size_t vdb_exact_topk_tc_workspace(int Q, int64_t N, int D, int k, int metric) {
    TcPlan p;
    if (!plan_tc(Q, N, D, k, metric, &p)) return 0;
    return tc_workspace(p, Q, Q, N, D, k, true).total;
}

int vdb_exact_topk_tc(const float* q, int Q, const float* db, const void* mirror, const float* db_stats,
                      const int64_t* ids, int64_t N, int D, int k, int metric, float* out_d,
                      int64_t* out_i, int32_t* out_uncertified, void* ws, size_t ws_bytes, void* stream) {
    return tc_search_impl(q, Q, db, mirror, db_stats, ids, N, D, k, metric, out_d, out_i, out_uncertified, ws, ws_bytes,
                          stream, nullptr);
}

size_t vdb_ivf_topk_tc_workspace(int Q, int64_t Qa, int64_t N, int D, int k, int n_units) {
    TcIvf v{1, nullptr, nullptr, nullptr, nullptr, Qa, nullptr, n_units, nullptr, nullptr};
    TcPlan p;
    if (Q <= 0 || Qa <= 0 || n_units <= 0 || !plan_tc_ivf(v, N, D, k, &p)) return 0;
    return tc_workspace(p, Q, Qa, N, D, k, false).total;
}

int vdb_ivf_topk_tc(const float* q, int Q, const float* lm_vecs, const void* lm_mirror, const float* lm_stats,
                    const int64_t* lm_ids, int64_t N, int D, int k, int n_probe, const int32_t* pair_rows,
                    const int32_t* row_query, const int32_t* row_list, const float* centroids, int64_t Qa,
                    const int32_t* unit_table, int n_units, const int32_t* meta, const float* list_stats, float* out_d,
                    int64_t* out_i, int32_t* out_uncertified, void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(n_probe > 0 && n_probe <= 32, VDB_ERR_UNSUPPORTED, "vdb_ivf_topk_tc: n_probe=%d (at most 32)", n_probe);
    VDB_REQUIRE(pair_rows && row_query && row_list && centroids && unit_table && n_units > 0 && Qa > 0 && Qa % kTileM == 0 && N % kTileN == 0,
                VDB_ERR_INVALID, "vdb_ivf_topk_tc: bad pair / unit tables (Qa=%lld, N=%lld, units=%d)", (long long)Qa,
                (long long)N, n_units);
    TcIvf v{n_probe, pair_rows, row_query, row_list, centroids, Qa, reinterpret_cast<const int4*>(unit_table), n_units,
            meta ? meta + 1 : nullptr, reinterpret_cast<const float2*>(list_stats)};
    return tc_search_impl(q, Q, lm_vecs, lm_mirror, lm_stats, lm_ids, N, D, k, VDB_METRIC_L2, out_d, out_i, out_uncertified,
                          ws, ws_bytes, stream, &v);
}

}  // extern "C"

namespace vdb {

static int tc_search_impl(const float* q, int Q, const float* db, const void* mirror, const float* db_stats,
                          const int64_t* ids, int64_t N, int D, int k, int metric, float* out_d, int64_t* out_i,
                          int32_t* out_uncertified, void* ws, size_t ws_bytes, void* stream, const TcIvf* ivf) {
    VDB_REQUIRE(metric == VDB_METRIC_L2 || metric == VDB_METRIC_COSINE, VDB_ERR_INVALID, "unknown metric %d", metric);
    VDB_REQUIRE(Q > 0 && N > 0 && D > 0 && k > 0, VDB_ERR_INVALID, "vdb_exact_topk_tc: bad shape");
    VDB_REQUIRE(q && db && mirror && db_stats && out_d && out_i && out_uncertified && ws, VDB_ERR_INVALID,
                "vdb_exact_topk_tc: null pointer");
    VDB_REQUIRE(N < (1ll << 31), VDB_ERR_UNSUPPORTED, "N too large for int32 candidate rows");
    TcPlan p;
    VDB_REQUIRE(ivf ? plan_tc_ivf(*ivf, N, D, k, &p) : plan_tc(Q, N, D, k, metric, &p), VDB_ERR_UNSUPPORTED,
                "tensor-core scan does not support D=%d k=%d", D, k);
    const int64_t Qa = ivf ? ivf->Qa : Q;                            // rows of the query operand
    const TcWorkspace w = tc_workspace(p, Q, Qa, N, D, k, ivf == nullptr);
    VDB_REQUIRE(ws_bytes >= w.total, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, w.total);
    VDB_REQUIRE((reinterpret_cast<uintptr_t>(mirror) & 15) == 0 && (reinterpret_cast<uintptr_t>(ws) & 255) == 0,
                VDB_ERR_INVALID, "operand mirror / workspace must be 16 / 256-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* base = static_cast<char*>(ws);
    __nv_bfloat16* qb = reinterpret_cast<__nv_bfloat16*>(base + w.qbf16);
    u64* partial = reinterpret_cast<u64*>(base + w.partial);
    float* tau = reinterpret_cast<float*>(base + w.tau);

    // per-list depth: lists of 8 are enough when there are many of them (L * 8 >= 4 * kNBFinal), debug bit 2 forces 16
    const int nbl = ((g_tc_flags & 4) || p.S * kGroups * 8 < 4 * kNBFinal) ? 16 : 8;
    uint32_t* gtau = reinterpret_cast<uint32_t*>(tau);            // per-query threshold exchange of the scan
    tc_query_operand_kernel<<<(unsigned)((Qa + 7) / 8), 256, 0, st>>>(q, qb, (int)Qa, D, p.Dp, metric, gtau, out_uncertified,
                                                                     reinterpret_cast<int32_t*>(base + w.counters),
                                                                     ivf ? ivf->row_query : nullptr, ivf ? ivf->row_list : nullptr,
                                                                     ivf ? ivf->centroids : nullptr,
                                                                     ivf ? reinterpret_cast<float2*>(base + w.qstat) : nullptr);
    CUtensorMap map_a, map_b;
    CUtensorMap map_a2, map_b2;
    int rc = make_map(&map_a, qb, Qa, D, p.Dp, kBlockK, kTileM);
    if (rc) return rc;
    rc = make_map(&map_b, mirror, N, D, D, kBlockK, kTileN);
    if (rc) return rc;
    // norm block (L2 only; cosine passes valid but unused maps): columns D..D+15 of the same operands
    rc = make_map(&map_a2, qb, Qa, p.Dp, p.Dp, 16, kTileM);
    if (rc) return rc;
    // norm area viewed as lines of 4 rows (64 bf16 = 128 B): a 256-row tile is 64 line fetches, no swizzle (pre-applied)
    rc = make_map(&map_b2, p.has_norm ? static_cast<const void*>(static_cast<const __nv_bfloat16*>(mirror) + tc_norm_offset(N, D)) : mirror,
                  p.has_norm ? (N + 3) / 4 : N, p.has_norm ? 64 : D, p.has_norm ? 64 : D, 64, 64, p.has_norm ? 0 : -1);
    if (rc) return rc;
    if (p.pair && !ivf) {
        CUtensorMap map_bh;
        rc = make_map(&map_bh, mirror, N, D, D, kBlockK, kTileN / 2);
        if (rc) return rc;
        ProfScope prof(kProfExactTC, st);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)p.grid);
        cfg.blockDim = dim3(kTcThreads);
        cfg.dynamicSmemBytes = p.smem;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        const int Qi = (int)Qa;
        if (nbl == 16) {
            VDB_CUDA(cudaFuncSetAttribute(exact_tc2_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
            VDB_CUDA(cudaLaunchKernelEx(&cfg, exact_tc2_kernel<16>, map_a, map_bh, Qi, N, D, p.num_kb, p.stages, p.m_pairs, p.S,
                                        p.tiles_per_slice, p.tiles_total, g_tc_flags, partial, gtau));
        } else {
            VDB_CUDA(cudaFuncSetAttribute(exact_tc2_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
            VDB_CUDA(cudaLaunchKernelEx(&cfg, exact_tc2_kernel<8>, map_a, map_bh, Qi, N, D, p.num_kb, p.stages, p.m_pairs, p.S,
                                        p.tiles_per_slice, p.tiles_total, g_tc_flags, partial, gtau));
        }
    } else {
        ProfScope prof(kProfExactTC, st);
        if (nbl == 16) {
            VDB_CUDA(cudaFuncSetAttribute(exact_tc_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
            exact_tc_kernel<16><<<p.grid, kTcThreads, p.smem, st>>>(map_a, map_b, map_a2, map_b2, (int)Qa, N, D, p.has_norm, p.num_kb, p.stages, p.m_tiles, p.S,
                                                                   p.tiles_per_slice, p.tiles_total, g_tc_flags, partial, gtau,
                                                                   ivf ? ivf->utab : nullptr, ivf ? ivf->units_dev : nullptr);
        } else {
            VDB_CUDA(cudaFuncSetAttribute(exact_tc_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
            exact_tc_kernel<8><<<p.grid, kTcThreads, p.smem, st>>>(map_a, map_b, map_a2, map_b2, (int)Qa, N, D, p.has_norm, p.num_kb, p.stages, p.m_tiles, p.S,
                                                                  p.tiles_per_slice, p.tiles_total, g_tc_flags, partial, gtau,
                                                                  ivf ? ivf->utab : nullptr, ivf ? ivf->units_dev : nullptr);
        }
    }
    VDB_CUDA(cudaGetLastError());
    const int L = p.S * kGroups;                                     // lists per operand row
    const int lists = L * (ivf ? ivf->n_probe : 1);                  // lists the select kernel merges per query
    const int C = next_pow2(lists * nbl < 64 ? 64 : lists * nbl);
    VDB_REQUIRE(C <= 1024, VDB_ERR_UNSUPPORTED, "too many bin lists per query (%d x %d)", lists, nbl);
    // bins handed to the rerank per query (the post-scan kernel drops those that provably cannot matter);
    // debug bits 12..19 override the count, bit 4 disables the trimming
    int nbf = kNBFinal;
    if ((g_tc_flags >> 12) & 0xFF) nbf = (g_tc_flags >> 12) & 0xFF;
    if (nbf > kNBFinal) nbf = kNBFinal;
    if (nbf < k) nbf = k;
    TcPostQuery* pq = reinterpret_cast<TcPostQuery*>(base + w.pq);
    uint2* items = reinterpret_cast<uint2*>(base + w.items);
    u64* keys = reinterpret_cast<u64*>(base + w.keys);
    int32_t* item_count = reinterpret_cast<int32_t*>(base + w.counters);
    {
        ProfScope prof(kProfRerank, st);
        const size_t sel_smem = (size_t)kSelWarps * C * 8;
        VDB_CUDA(cudaFuncSetAttribute(tc_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sel_smem));
        tc_select_kernel<<<(Q + kSelWarps - 1) / kSelWarps, 32 * kSelWarps, sel_smem, st>>>(
            partial, L, nbl, nbf, C, q, D, k, metric, db_stats, (g_tc_flags & 16) ? 1 : 0, Q, pq, items, item_count,
            ivf ? ivf->pair_rows : nullptr, ivf ? ivf->n_probe : 1,
            ivf ? reinterpret_cast<const float2*>(base + w.qstat) : nullptr, ivf ? ivf->row_list : nullptr,
            ivf ? ivf->list_stats : nullptr);
        // one CTA per SM: as many warps (each with a two-slot ring of bin + query) as shared memory holds, at most 8
        int rr_warps = (int)((size_t)(max_smem_optin() - 2048) / (kRrSlots * rr_slot_bytes(D)));
        if (rr_warps > kRrMaxWarps) rr_warps = kRrMaxWarps;
        VDB_REQUIRE(rr_warps >= 1, VDB_ERR_UNSUPPORTED, "D=%d: a rerank ring does not fit shared memory", D);
        const size_t rr_smem = (size_t)rr_warps * kRrSlots * rr_slot_bytes(D);
        const int rr_grid = sm_count();
        const int rr_threads = 32 * rr_warps;
        const int fin_blocks = (Q + 7) / 8;
        if (metric == VDB_METRIC_L2) {
            VDB_CUDA(cudaFuncSetAttribute(tc_rerank_bins_kernel<VDB_METRIC_L2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rr_smem));
            tc_rerank_bins_kernel<VDB_METRIC_L2><<<rr_grid, rr_threads, rr_smem, st>>>(items, item_count, q, db, N, D, pq, keys);
            tc_final_kernel<VDB_METRIC_L2><<<fin_blocks, 256, 0, st>>>(keys, pq, Q, k, ids, out_d, out_i, out_uncertified);
        } else {
            VDB_CUDA(cudaFuncSetAttribute(tc_rerank_bins_kernel<VDB_METRIC_COSINE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rr_smem));
            tc_rerank_bins_kernel<VDB_METRIC_COSINE><<<rr_grid, rr_threads, rr_smem, st>>>(items, item_count, q, db, N, D, pq, keys);
            tc_final_kernel<VDB_METRIC_COSINE><<<fin_blocks, 256, 0, st>>>(keys, pq, Q, k, ids, out_d, out_i, out_uncertified);
        }
    }
    count_launch(5);
    VDB_CUDA(cudaGetLastError());
    // list-major IVF: a brute-force rescan would look at lists the query did not probe -- the caller re-runs the listed
    // queries through vdb_ivf_scan
    if (ivf) return VDB_OK;
    // queries the certificate rejected: exact fp32 scan, driven by the device-side list (no host round trip)
    return exact_f32_subset(q, Q, db, ids, N, D, k, metric, out_uncertified + 1, out_uncertified, out_d, out_i,
                            base + w.rescan, w.rescan_bytes, st);
}

}  // namespace vdb
