// PQ encode on the tensor cores, f16-accumulator form (r02): codes[i][j] = argmin_c sqrt(sum((C_j[c]-x_ij)^2)), first
// min wins (src/vector_db/inference/pq.py:77-89), for dsub = 8 and K <= 256.
//
// Why a second form.  pq_tc.cu keeps fp32 accumulators and folds the 256 scores of a (row, sub-space) with one FMNMX3
// per score; every min/max flavour shares ONE half-rate ALU pipe (scripts/ubench/minmax.cu), so that kernel cannot go
// below 1.32 ms per 2M x 384 rows and runs at 2.8.  The only way under one instruction per score is to read the scores
// from TMEM already PACKED two per register (tcgen05.ld .pack::16b needs f16 accumulators, which tcgen05.mma accepts
// only with f16 operands -- bf16 operands trap, scripts/ubench/f16acc.cu) and fold them with the 3-input packed minimum
// (VHMNMX): 0.5 instructions per score.
//
// Making f16 scores good enough.  f16 keeps 11 bits RELATIVE, so the score is arranged to be small where it matters:
// the accumulator holds the whole squared distance S^2 |x - c|^2 (S = 2^e scales the sub-space so that S max|c| is in
// [2, 4)), which is ~1-4 around the minimum (ulp 2^-10 .. 2^-9) instead of |c|^2 - 2 x.c ~ 16.  Two MMAs per 128-column
// half, ordered so that the intermediate f16 rounding is harmless (measured: the stored value is rn16(rn16(P0) + P1)):
//   x S  = x1 + x2 + rest (f16 pieces, rest <= 2^-22)      -2 c S = c1 + c2 + rest
//   P0 = [x2 | x1] . [c1 | c2]                              -- the small cross terms first: |P0| <= 2^-9 |x||c|
//   P1 = [x1 | 1 1 1 X1 X2 X3 0 0] . [c1 | N1 N2 N3 1 1 1 0 0]   -- x1.c1 + |c|^2 + |x|^2 (3-way f16 splits of the norms)
// so the only rounding of size is the last one, half an ulp of the distance itself.  Candidates = every centroid whose
// f16 score is within 1.5 ulp16(min) + 2 eps of the row minimum (eps = 1.2e-5 (|x|+|c|)^2 + 1e-6 bounds the dropped
// split terms, the first rounding, the MMA's accumulation and the reference's own fp32 rounding, in scaled units).
//
// Fold: one pass, TWO families of packed minima, 32 VHMNMX per 64 scores:
//   A[a]  (16 registers) = min over the 16 columns of block a, even and odd columns apart     (32 minima)
//   B[i]  ( 8 registers) = min over the columns = 2i, 2i+1 (mod 16), even and odd apart        (16 minima)
// Two different centroids differ in (block, parity) or in (pair-within-block, parity), so "exactly one A minimum and
// exactly one B minimum within the margin" names the arg-min (16 a + 2 i + parity) -- no second pass, no per-score
// compare.  The margin test is a packed compare v = (V <= thr) per minimum (HSET2) and sum v 2^i (HFMA2), which builds
// the candidate bit masks in f16 arithmetic.
//
// The ~1 % of (row, sub-space) pairs with several candidates (a superset: the cross product of the families) must not be
// resolved by the epilogue lane itself -- a quarter of the warp-steps would stall on L2 latency.  The lane pushes
// (row, sub-space, masks) to a shared-memory queue; a RESOLVER warp evaluates the candidates with the reference's exact
// arithmetic (numpy's pairwise sum, sqrt, first index wins) from the fp32 rows and codebooks (L2) and writes that one
// code byte itself.  Every code is written straight to global memory as its own byte (no staging buffer, so the two
// writers never touch the same byte and need no ordering).  Rows the f16 range cannot hold (S^2 |x|^2 > 30000, NaN) go
// to the resolver with all-ones masks: correct, slow, and rare on data the codebook was trained for.
//
// Measured (2M x 384 rows, m = 48, K = 256, B200; scripts/dev_pq16.py times both kernels in one process): 2.05 ms against
// 2.91 ms for pq_tc.cu on the same box, codes equal on all 96 M (also at scales 0.001 .. 30, K = 5 / 37 / 100, m = 1 / 13 / 16, N
// not a multiple of anything), ~1.06 % of the (row, sub-space) pairs go through the resolver, DRAM read 3.08 GB = the
// algorithmic bytes.  ncu (profiles/r02_ncu_summary.json, pq_encode_tc16): ALU pipe 74 % busy, issue slots 70 %, tensor pipe
// 37 % -- the 137 packed minima of a warp-step are a third of its ~430 instructions; the rest of the ALU load is the mask
// arithmetic (HSET2, LOP3, POPC / FLO) and the split warps' conversions.
//
// How it got there, each step measured (r02):
//   2.63 first version (two epilogue groups, one per stage; split = one warp per row quarter of every step)
//   2.51 x16 TMEM loads instead of x32 (no spills in the loop), the resolver takes whole batches of 32
//   2.43 the tail of step s software-pipelined into the folds of step s + 1
//   2.24 producers wait with a suspend-time hint (their probe loops were 15 % of the issued instructions); packed compare
//        (HSET2) instead of the saturating-FMA indicator
//   2.15 THREE epilogue groups taking the steps in turn (step s: group s mod 3, stage s & 1), no software pipelining (96
//        registers at 640 threads).  A parity wait is only right when the barrier is at most one phase behind: with one
//        "full" barrier per stage the wait for use u passed on the completion of use u - 2 and the kernel dead-locked
//        after a few rounds -- eight barriers per stage, used in turn
//   2.05 ONE accumulator hand-over per step (two N = 256 MMAs, one commit) instead of two 128-column halves: every
//        mbarrier probe / tcgen05.mma / tcgen05.commit is a ~50-100 clock trip for the single issuing thread
// Measured and NOT kept: four groups, each a 128-column half of a stage, the halves of a step meeting through an 8-byte
// summary per row in shared memory (2.6 ms: 294 instructions per half-step, a quarter of them barrier probes); four groups
// in turn (2.24 ms: 80 registers, spills; with single-buffered TMEM loads to save registers 2.59 ms); hinted waits on the epilogue's own barriers (3.1 ms: late wake-ups); dropping the
// per-step commits of the operand / codebook slots (no change -- the step's "full" barrier frees the operand slot now).
// With the epilogue, the MMAs, the split arithmetic and the row loads all switched off the ring alone took 0.9-1.4 ms:
// the step is 128 clocks of tensor time, so barrier round trips are of the order of the step.  Row slices are loaded
// L2-evict-normal: a 32-byte slice shares its 64-byte DRAM burst with the next sub-space's slice, and evict-first lost
// it before its turn (5.0 GB read instead of 3.08).  TMEM reads are not the limit (scripts/ubench/ldtm.cu: packed x16 /
// x32 loads move 970-1250 B of TMEM cells per clock and SM with two warps per sub-partition, a step needs 128 KB).
//
// Roofline: HBM, algorithmic bytes per vector = 4 D + m (rows read once, codes written once).
#include <cuda.h>
#include <cuda_fp16.h>
#include <math.h>

#include "common.cuh"
#include "pq_common.cuh"
#include "sm100.cuh"
#include "tma_host.cuh"

namespace vdb {

using namespace sm100;

constexpr int kE16Groups = 3;                // epilogue groups, taking the CTA's steps in turn (step s: group s mod 3, stage s & 1)
constexpr int kE16Threads = 128 + 128 * kE16Groups + 128;   // warp 0 TMA, 1-2 MMA issuers (2: TMEM alloc), 3 resolver, then the epilogue groups, then 4 split warps
constexpr int kE16Rows = 128;               // rows per tile (UMMA M)
constexpr int kE16K = 256;                  // centroids per sub-space (UMMA N)
constexpr int kE16XBytes = kE16Rows * 32;   // 4 KB: 128 rows x 8 fp32 (TMA landing tile, not swizzled)
constexpr int kE16ABytes = kE16Rows * 32;   // 4 KB: 128 rows x 16 f16 (32-byte swizzle)
constexpr int kE16BBytes = kE16K * 32;      // 8 KB: 256 centroids x 16 f16 (32-byte swizzle)
constexpr int kE16AsBytes = 2 * kE16ABytes;                  // A0 | A1  (8 KB per step)
constexpr int kE16BsBytes = 2 * kE16BBytes;                  // B0 | B1  (16 KB per sub-space, shared by the two tiles of a pair)
constexpr int kE16MaxX = 32, kE16MaxA = 15, kE16MaxBs = 8;   // operand slots <= 15: see the full barriers' second use
constexpr int kE16QCap = 512;               // resolver queue (16-byte items)
constexpr float kE16Big = 60000.f;          // score of padding centroids / overflowing rows (f16 max 65504)
constexpr float kE16Limit = 30000.f;        // a row minimum at or above this cannot be trusted: all-ones masks

// Work order as in pq_tc.cu: a CTA takes PAIRS of row tiles and walks the sub-spaces j-major over the pair, so the codebook
// operands of sub-space j are fetched once for two steps; step s uses accumulator stage s & 1 = its tile within the pair.
// Three rings.  ROWS: the fp32 slice of a step as TMA lands it (4 KB), free again as soon as the split warp has read it;
// OPERANDS: the step's two f16 A tiles (8 KB), free when its MMAs are done; CODEBOOK: B0 | B1 of a sub-space (16 KB, two
// steps).  (One 12 KB slot per step for rows + operands held the landing tile through the split AND the MMA: the slot's
// turn-around -- HBM latency + split + MMA, ~5800 clocks -- over 11 slots was 524 clocks per step, a 1.36 ms floor per
// 2M x 384 rows with the epilogue and the MMAs switched off.)
struct E16Smem {
    int x_off, as_off, bs_off, par_off, q_off, bar_off, total;
};
__host__ __device__ inline E16Smem e16_layout(int nx, int na, int nbs, int m) {
    E16Smem L;
    L.bs_off = 0;                                              // 16 KB slots first: 1024-byte aligned
    L.as_off = nbs * kE16BsBytes;                              // 8 KB slots: 1024-byte aligned as well
    L.x_off = L.as_off + na * kE16AsBytes;                     // 4 KB slots
    L.par_off = L.x_off + nx * kE16XBytes;                     // [m] float2 (scale S, S max|c|)
    L.q_off = (L.par_off + m * 8 + 15) & ~15;
    L.bar_off = L.q_off + kE16QCap * 16;
    L.total = L.bar_off + (2 * kE16MaxX + 2 * kE16MaxA + 2 * kE16MaxBs + 20) * 8 + 32;
    return L;
}

__device__ __forceinline__ uint32_t pack_f16x2(float lo, float hi) {           // round to nearest even
    uint32_t r;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ float f16_lo(uint32_t p) { return __half2float(__ushort_as_half((unsigned short)(p & 0xFFFFu))); }
__device__ __forceinline__ float f16_hi(uint32_t p) { return __half2float(__ushort_as_half((unsigned short)(p >> 16))); }
__device__ __forceinline__ float f16_round(float v) { return __half2float(__float2half_rn(v)); }

// kind::f16 instruction descriptor with f16 operands AND f16 accumulators (c_format 0, a_format 0, b_format 0)
__host__ __device__ constexpr uint32_t make_idesc_f16_f16acc(int M, int N) {
    return ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// Operand tensor [m][2][256] rows of 16 f16: part 0 = [c1 | c2], part 1 = [c1 | N1 N2 N3 1 1 1 0 0] with
// -2 c S = c1 + c2 + rest and |c|^2 S^2 = N1 + N2 + N3 + rest; S = 2^e with S max|c| in [2, 4).  Padded to 256 centroids
// per sub-space (padding scores kE16Big: can never win).  par[j] = (S, S max|c| rounded up).
__global__ void pq_tc16_operands_kernel(const float* __restrict__ cb, int m, int K, uint32_t* __restrict__ ops,
                                        float2* __restrict__ par) {
    const int j = blockIdx.x;
    __shared__ float red[256];
    __shared__ float s_scale;
    float mx = 0.f;
    for (int c = threadIdx.x; c < K; c += blockDim.x) {
        const float* src = cb + ((size_t)j * K + c) * 8;
        float s = 0.f;
#pragma unroll
        for (int t = 0; t < 8; ++t) s = fmaf(src[t], src[t], s);
        mx = fmaxf(mx, s);
    }
    red[threadIdx.x] = mx;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] = fmaxf(red[threadIdx.x], red[threadIdx.x + o]);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float cm = sqrtf(red[0]) * 1.0000005f;            // max |c|, rounded up
        float S = 1.f;
        if (cm > 0.f && cm < 3.0e38f) {
            int e;
            frexpf(cm, &e);                                      // cm = f * 2^e, f in [0.5, 1)  ->  cm * 2^(2 - e) in [2, 4)
            int se = 2 - e;
            se = se < -100 ? -100 : (se > 100 ? 100 : se);
            S = ldexpf(1.f, se);
        }
        s_scale = S;
        par[j] = make_float2(S, cm * S);
    }
    __syncthreads();
    const float S = s_scale;
    for (int c = threadIdx.x; c < 256; c += blockDim.x) {
        uint32_t* o0 = ops + (((size_t)j * 2 + 0) * 256 + c) * 8;      // 8 words = 16 f16
        uint32_t* o1 = ops + (((size_t)j * 2 + 1) * 256 + c) * 8;
        uint32_t p1[4] = {0u, 0u, 0u, 0u}, p2[4] = {0u, 0u, 0u, 0u};
        float s = kE16Big;
        if (c < K) {
            const float* src = cb + ((size_t)j * K + c) * 8;
            s = 0.f;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const float v0 = src[2 * t] * S, v1 = src[2 * t + 1] * S, w0 = -2.f * v0, w1 = -2.f * v1;
                p1[t] = pack_f16x2(w0, w1);
                p2[t] = pack_f16x2(w0 - f16_lo(p1[t]), w1 - f16_hi(p1[t]));
                s = fmaf(v0, v0, s);
                s = fmaf(v1, v1, s);
            }
            if (!(s < kE16Big)) s = kE16Big;                   // a codebook f16 cannot hold: the resolver decides every row
        }
        const float n1 = f16_round(s), r1 = s - n1, n2 = f16_round(r1), n3 = f16_round(r1 - n2);
#pragma unroll
        for (int t = 0; t < 4; ++t) { o0[t] = p1[t]; o0[4 + t] = p2[t]; o1[t] = p1[t]; }
        o1[4] = pack_f16x2(n1, n2);
        o1[5] = pack_f16x2(n3, 1.f);
        o1[6] = pack_f16x2(1.f, 1.f);
        o1[7] = 0u;
    }
}

// ---------------------------------------------------------------- packed f16x2 helpers
__device__ __forceinline__ uint32_t hmin2u(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("min.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
// three-input packed minimum: one VHMNMX (ptxas fuses the two-input pair)
__device__ __forceinline__ uint32_t hmin3u(uint32_t a, uint32_t b, uint32_t c) { return hmin2u(hmin2u(a, b), c); }
__device__ __forceinline__ uint32_t hfma2_sat(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("fma.rn.sat.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t hfma2u(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t hle2u(uint32_t a, uint32_t b) {        // (a <= b) per half as 1.0 / 0.0
    uint32_t r;
    asm("set.le.f16x2.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ uint32_t hadd2u(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
constexpr uint32_t kInf2 = 0x7C007C00u;     // (+inf, +inf)

// 32 lanes x 32 consecutive f16 accumulator columns -> 16 registers of packed pairs (column 2i low, 2i + 1 high)
__device__ __forceinline__ void tmem_ld_x16_pack16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait_u(uint32_t (&r)[16]) {    // the loaded registers pass through the wait
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :
                 : "memory");
}
// One 32-column chunk CH (16 packed registers = blocks 2 CH, 2 CH + 1) folded into the two families: 16 instructions.
template <int CH>
__device__ __forceinline__ void fold_chunk16(const uint32_t (&v)[16], uint32_t (&A)[16], uint32_t (&B)[8]) {
#pragma unroll
    for (int b = 0; b < 2; ++b) {
        const uint32_t* w = v + 8 * b;
        A[2 * CH + b] = hmin2u(hmin3u(hmin3u(w[0], w[1], w[2]), hmin3u(w[3], w[4], w[5]), w[6]), w[7]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) B[i] = hmin3u(B[i], v[i], v[i + 8]);
}

// The reference's own arithmetic over the candidates the masks name (pq.py:79-80): centroid 16 a + 2 i + p is a candidate
// when bit (16 p + a) of mA and bit (8 p + i) of mB are set; first minimum wins.  Candidates are taken four at a time so
// that their centroid loads (L2) are in flight together -- an item has 2-4 candidates, i.e. one round trip.
struct CandIter {
    uint32_t mA, mB, ra, rb, rb_full;
    int p;
    __device__ __forceinline__ void init(uint32_t a, uint32_t b) {
        mA = a; mB = b; p = 0;
        rb_full = mB & 0xFFu;
        rb = rb_full;
        ra = rb_full ? (mA & 0xFFFFu) : 0u;
    }
    __device__ __forceinline__ int next() {
        for (;;) {
            if (ra == 0u) {
                if (p == 1) return -1;
                p = 1;
                rb_full = (mB >> 8) & 0xFFu;
                rb = rb_full;
                ra = rb_full ? (mA >> 16) : 0u;
                continue;
            }
            if (rb == 0u) { ra &= ra - 1u; rb = rb_full; continue; }
            const int i = __ffs(rb) - 1;
            rb &= rb - 1u;
            return (__ffs(ra) - 1) * 16 + 2 * i + p;
        }
    }
};
__device__ __noinline__ int pq16_resolve(const float* __restrict__ xrow, const float* __restrict__ cbj, int K, uint32_t mA,
                                         uint32_t mB) {
    const float4 xa = __ldg(reinterpret_cast<const float4*>(xrow)), xb = __ldg(reinterpret_cast<const float4*>(xrow) + 1);
    float best = INFINITY;
    int code = 0;
    CandIter it;
    it.init(mA, mB);
    for (;;) {
        int c[4];
        float4 ca[4], cc[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            c[t] = it.next();
            if (c[t] >= K) c[t] = -2;                              // padding centroid: skip, keep going
            if (c[t] >= 0) {
                const float4* cp = reinterpret_cast<const float4*>(cbj + (size_t)c[t] * 8);
                ca[t] = __ldg(cp);
                cc[t] = __ldg(cp + 1);
            }
        }
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            if (c[t] >= 0) {
                float d[8] = {ca[t].x - xa.x, ca[t].y - xa.y, ca[t].z - xa.z, ca[t].w - xa.w,
                              cc[t].x - xb.x, cc[t].y - xb.y, cc[t].z - xb.z, cc[t].w - xb.w};
                const float dist = sqrtf(np_sumsq<8>(d));
                if (dist < best || (dist == best && c[t] < code)) { best = dist; code = c[t]; }
            }
        }
        if (c[3] == -1) break;
    }
    return code;
}

__global__ void __launch_bounds__(kE16Threads, 1)
pq_encode_tc16_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_b,
                      const float* __restrict__ x, int64_t N, int D, int m, int K, const float* __restrict__ cb,
                      const float2* __restrict__ par, int nx, int na, int nbs, int64_t rounds, uint8_t* __restrict__ codes,
                      unsigned long long* __restrict__ stats, int T, uint64_t x_hint) {
    extern __shared__ __align__(1024) unsigned char smem[];
    if ((smem_u32(smem) & 1023u) != 0) __trap();
    const E16Smem L = e16_layout(nx, na, nbs, m);
    unsigned char* sX = smem + L.x_off;
    unsigned char* sAs = smem + L.as_off;
    unsigned char* sBs = smem + L.bs_off;
    float2* sPar = reinterpret_cast<float2*>(smem + L.par_off);
    uint4* sQ = reinterpret_cast<uint4*>(smem + L.q_off);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    uint64_t* x_full = bars;                        // [nx]: TMA landed the fp32 row slice
    uint64_t* x_empty = bars + kE16MaxX;            // [nx]: the step's split warp has read it
    uint64_t* a_full = bars + 2 * kE16MaxX;         // [na]: A0 / A1 written by the step's split warp
    uint64_t* b_full = a_full + 2 * kE16MaxA;       // [nbs]: TMA landed B0, B1  (the operand slot is freed by the step's full barrier)
    uint64_t* b_empty = b_full + kE16MaxBs;         // [nbs]: the MMAs of BOTH steps of the sub-space are done
    // EIGHT "full" barriers per stage, used by the stage's uses in turn.  A parity wait is only right when the barrier
    // is at most one phase behind the awaited one.  With three groups taking the steps in turn, the group that waits for use
    // u of a stage last consumed use u - 3 of it, so uses up to u - 3 have been produced -- but u - 1 and u - 2 (other
    // groups' steps) may still be pending: with one barrier per stage the wait for use u then passes on the completion of
    // use u - 2 (same parity).  With several barriers the same-parity predecessor is many uses back, long produced.  The
    // same barrier tells the split warps that the step's operand slot is free again (na <= 15 steps later, before the
    // barrier's next use 16 steps later can complete).
    uint64_t* tmem_full = b_empty + kE16MaxBs;      // [2 accumulator stages][use & 7]
    uint64_t* tmem_empty = tmem_full + 16;          // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    volatile uint32_t* q_head = tmem_ptr + 1;       // items pushed so far
    volatile uint32_t* q_tail = tmem_ptr + 2;       // items the resolver has taken
    volatile uint32_t* q_done = tmem_ptr + 3;       // epilogue groups that have finished
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t hint = 2000u;                                          // suspend-time hint of the producers' waits (ns)

    for (int i = threadIdx.x; i < m; i += kE16Threads) sPar[i] = par[i];
    for (int i = threadIdx.x; i < kE16QCap; i += kE16Threads) sQ[i] = make_uint4(0u, 0u, 0u, 0u);
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < nx; ++i) { mbar_init(&x_full[i], 1); mbar_init(&x_empty[i], 1); }
        for (int i = 0; i < na; ++i) mbar_init(&a_full[i], 1);
        for (int i = 0; i < nbs; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 2); }
        for (int i = 0; i < 16; ++i) mbar_init(&tmem_full[i], 1);
        for (int i = 0; i < 2; ++i) mbar_init(&tmem_empty[i], 4);
        *q_head = 0u; *q_tail = 0u; *q_done = 0u;
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            int xs = 0, bs = 0;
            uint32_t xph = 0, bph = 0;
            for (int64_t round = blockIdx.x; round < rounds; round += gridDim.x) {
                for (int j = 0; j < m; ++j) {
                    mbar_wait_relaxed(&b_empty[bs], bph ^ 1, hint);
                    unsigned char* bb = sBs + (size_t)bs * kE16BsBytes;
                    mbar_expect_tx(&b_full[bs], (uint32_t)kE16BsBytes);
                    tma_load_2d(bb, &map_b, &b_full[bs], 0, (j * 2) * 256, kEvictLast);
                    tma_load_2d(bb + kE16BBytes, &map_b, &b_full[bs], 0, (j * 2 + 1) * 256, kEvictLast);
                    if (++bs == nbs) { bs = 0; bph ^= 1; }
                    for (int t = 0; t < T; ++t) {
                        mbar_wait_relaxed(&x_empty[xs], xph ^ 1, hint);
                        mbar_expect_tx(&x_full[xs], (uint32_t)kE16XBytes);
                        // rows past N (a round's missing tiles included) land as zeros and are never stored
                        tma_load_2d(sX + (size_t)xs * kE16XBytes, &map_x, &x_full[xs], j * 8, (int)((round * T + t) * kE16Rows), x_hint);
                        if (++xs == nx) { xs = 0; xph ^= 1; }
                    }
                }
            }
        }
    } else if (warp < 3) {
        // ================= MMA issuers: warp 1 owns accumulator stage 0 (tile 0 of every pair), warp 2 stage 1 =================
        if (lane == 0) {
            const int as = warp - 1;
            const uint32_t idesc = make_idesc_f16_f16acc(kE16Rows, kE16K);
            uint32_t use = 0u;                         // own steps so far = use number of the stage
            int xs = as % na, bs = 0, aph = 0;         // own steps s = as, as + 2, ...: operand slot s mod na, codebook slot (s / T) mod nbs
            uint32_t xph = 0, bph = 0;
            for (int64_t round = blockIdx.x; round < rounds; round += gridDim.x) {
                for (int j = 0; j < m; ++j) {
                    mbar_wait_relaxed(&b_full[bs], bph, hint);
                  for (int u = as; u < T; u += 2) {
                    mbar_wait_relaxed(&a_full[xs], xph, hint);
                    const uint32_t a0 = smem_u32(sAs + (size_t)xs * kE16AsBytes);
                    const uint32_t b0 = smem_u32(sBs + (size_t)bs * kE16BsBytes);
                    const uint64_t d_a0 = make_desc_k_sw32(a0), d_a1 = make_desc_k_sw32(a0 + kE16ABytes);
                    // ONE hand-over per step: the issuing thread's own latency (every mbarrier probe, tcgen05.mma and
                    // tcgen05.commit is a ~50-100 clock trip for a single thread) was the ring's floor with two 128-column
                    // halves per step -- 13 such operations per step and issuer, ~500 clocks per step of the CTA
                    mbar_wait(&tmem_empty[as], aph ^ 1);        // the hand-over is on the critical path: plain probing
                    tc_fence_after();
                    {
                        const uint32_t d_tmem = tmem_base + (uint32_t)(as * kE16K);
                        umma_bf16(d_tmem, d_a0, make_desc_k_sw32(b0), idesc, 0u);                      // x2.c1 + x1.c2 (small, first)
                        umma_bf16(d_tmem, d_a1, make_desc_k_sw32(b0 + kE16BBytes), idesc, 1u);         // + x1.c1 + |c|^2 + |x|^2
                        umma_commit(&tmem_full[as * 8 + (int)(use & 7u)]);     // also frees the step's operand slot (split warps)
                    }
                    if (u + 2 >= T) umma_commit(&b_empty[bs]);                   // this issuer's last step on the sub-space
                    aph ^= 1;
                    ++use;
                    xs += 2;
                    if (xs >= na) { xs -= na; xph ^= 1; }
                  }
                    if (++bs == nbs) { bs = 0; bph ^= 1; }
                }
            }
        }
    } else if (warp >= 4 + 4 * kE16Groups) {
        // ================= operand split: warp 12 + w converts steps s = w (mod 4), all 128 rows (four rows per lane, their
        // dependency chains interleave):  v = S x;  x1 = f16(v), x2 = f16(v - x1);  X = 3-way f16 split of |v|^2;
        // A0 row = [x2 | x1], A1 row = [x1 | 1 1 1 X1 X2 X3 0 0].  A row the f16 range cannot hold becomes
        // [0 | 0], [0 | 1 1 1 Big 0 0 0 0].  (One warp per row quarter of EVERY step serialised the ~600-clock
        // load -> convert -> store -> fence chain of a step: 2.6 ms per 2M x 384 rows.) =================
        const int w = warp - (4 + 4 * kE16Groups);
        int xs = 0, as = 0;
        uint32_t xph = 0, aph = 0;
        uint32_t step = 0;
        for (int64_t round = blockIdx.x; round < rounds; round += gridDim.x) {
            for (int sidx = 0; sidx < T * m; ++sidx, ++step) {
                if ((int)(step & 3u) == w) {
                    const float S = sPar[sidx / T].x;
                    if (step >= (uint32_t)na) {                    // the operand slot is free: the MMAs of step - na are done
                        const uint32_t prev = step - (uint32_t)na, puse = prev >> 1;
                        mbar_wait_relaxed(&tmem_full[(prev & 1u) * 8u + (puse & 7u)], (puse >> 3) & 1u, hint);
                    }
                    mbar_wait_relaxed(&x_full[xs], xph, hint);
                    const unsigned char* base = sX + (size_t)xs * kE16XBytes;
                    unsigned char* abase = sAs + (size_t)as * kE16AsBytes;
#pragma unroll
                    for (int it = 0; it < 4; ++it) {
                        const int r = it * 32 + lane;
                        const uint32_t sw = (uint32_t)(((r >> 2) & 1) << 4);    // 32-byte swizzle: the 16-byte halves of a row trade places
                        const float4 a = *reinterpret_cast<const float4*>(base + r * 32);
                        const float4 b = *reinterpret_cast<const float4*>(base + r * 32 + 16);
                        const float v[8] = {a.x * S, a.y * S, a.z * S, a.w * S, b.x * S, b.y * S, b.z * S, b.w * S};
                        float xx = 0.f;
#pragma unroll
                        for (int t = 0; t < 8; ++t) xx = fmaf(v[t], v[t], xx);
                        uint4 p, q;
                        p.x = pack_f16x2(v[0], v[1]); p.y = pack_f16x2(v[2], v[3]); p.z = pack_f16x2(v[4], v[5]); p.w = pack_f16x2(v[6], v[7]);
                        q.x = pack_f16x2(v[0] - f16_lo(p.x), v[1] - f16_hi(p.x));
                        q.y = pack_f16x2(v[2] - f16_lo(p.y), v[3] - f16_hi(p.y));
                        q.z = pack_f16x2(v[4] - f16_lo(p.z), v[5] - f16_hi(p.z));
                        q.w = pack_f16x2(v[6] - f16_lo(p.w), v[7] - f16_hi(p.w));
                        const float n1 = f16_round(xx), r1 = xx - n1, n2 = f16_round(r1), n3 = f16_round(r1 - n2);
                        uint4 c = make_uint4(0x3C003C00u, pack_f16x2(1.f, n1), pack_f16x2(n2, n3), 0u);
                        if (!(xx <= kE16Limit)) {                           // too large for f16 (or NaN): the resolver decides this row
                            p = make_uint4(0u, 0u, 0u, 0u);
                            q = p;
                            c = make_uint4(0x3C003C00u, pack_f16x2(1.f, kE16Big), 0u, 0u);
                        }
                        unsigned char* a0 = abase + r * 32;
                        *reinterpret_cast<uint4*>(a0 + (0u ^ sw)) = q;                      // A0 = [x2 | x1]
                        *reinterpret_cast<uint4*>(a0 + (16u ^ sw)) = p;
                        *reinterpret_cast<uint4*>(a0 + kE16ABytes + (0u ^ sw)) = p;         // A1 = [x1 | constants and |x|^2]
                        *reinterpret_cast<uint4*>(a0 + kE16ABytes + (16u ^ sw)) = c;
                    }
                    fence_proxy_async();               // generic-proxy writes -> visible to the MMA's async-proxy reads
                    __syncwarp();
                    if (lane == 0) { mbar_arrive(&a_full[as]); mbar_arrive(&x_empty[xs]); }
                }
                if (++xs == nx) { xs = 0; xph ^= 1; }
                if (++as == na) { as = 0; aph ^= 1; }
            }
        }
    } else if (warp == 3) {
        // ================= resolver: drains the queue of (row, sub-space, candidate masks) items =================
        uint32_t tail = 0u;
        unsigned long long n_items = 0ull, n_full = 0ull;
        for (;;) {
            const uint32_t done = *q_done;
            const uint32_t head = *q_head;
            const uint32_t avail = head - tail;
            // whole batches of 32 (a call of the resolve routine costs the same ~1000 warp instructions for 1 item or 32 --
            // polling for single items made this warp 24 % of the SM's issued instructions); the rest when both groups are done
            if (avail < 32u && done != (uint32_t)kE16Groups) { __nanosleep(512); continue; }
            if (avail == 0u) break;                    // `done` was read BEFORE `head`: nothing can arrive after both groups finished
            const uint32_t n = avail < 32u ? avail : 32u;
            uint4 it = make_uint4(0u, 0u, 0u, 0u);
            if ((uint32_t)lane < n) {
                const uint32_t seq = tail + (uint32_t)lane;
                volatile uint4* slot = reinterpret_cast<volatile uint4*>(sQ + (seq % kE16QCap));
                while (slot->w != seq + 1u) {}         // the pusher writes the tag last
                __threadfence_block();
                it.x = slot->x; it.y = slot->y; it.z = slot->z;
            }
            __syncwarp();
            tail += n;
            if (lane == 0) *q_tail = tail;             // the slots may be reused from here on
            if ((uint32_t)lane < n) {
                const int64_t row = (int64_t)it.x;
                const int j = (int)(it.z >> 16);
                const int code = pq16_resolve(x + row * D + j * 8, cb + (size_t)j * K * 8, K, it.y, it.z & 0xFFFFu);
                codes[row * m + j] = (uint8_t)code;
                ++n_items;
                if (it.y == 0xFFFFFFFFu) ++n_full;
            }
        }
        if (stats) {
            if (n_items) atomicAdd(stats, n_items);
            if (n_full) atomicAdd(stats + 1, n_full);
        }
    } else {
        // ================= epilogue: three groups of four warps, one thread per row; the groups take the CTA's steps in
        // turn.  A step = fold the 256 packed scores (8 TMEM loads), release the stage, then the TAIL: row minimum ->
        // threshold -> candidate masks -> code / push (~170 mostly dependent instructions; three warps per scheduler
        // overlap one another's chains) =================
        const int eg = (warp - 4) >> 2;
        const int ew = warp & 3;
        const int rt = ew * 32 + lane;                 // row within the tile == TMEM lane
        const uint32_t taddr0 = tmem_base + ((uint32_t)(ew * 32) << 16);

        struct Tail {                                  // state of the step whose tail is pending
            uint8_t* out;                              // its code byte
            uint32_t row;
            int j;
            bool live;                                 // row < N and the step exists
            uint32_t th2;                              // piece 1 -> 2: the threshold in both halves
            bool trust;                                // minimum inside the f16 range
            uint32_t mA, mB;                           // piece 2 -> 3
        };
        // piece 1: row minimum (f16) and the candidate threshold
        auto tail1 = [&](const uint32_t (&A)[16], Tail& t) {
            uint32_t mm = hmin3u(hmin3u(A[0], A[1], A[2]), hmin3u(A[3], A[4], A[5]), hmin3u(A[6], A[7], A[8]));
            mm = hmin3u(mm, hmin3u(A[9], A[10], A[11]), hmin3u(A[12], A[13], A[14]));
            mm = hmin2u(mm, A[15]);
            mm = hmin2u(mm, __byte_perm(mm, mm, 0x1032));                   // both halves = the row minimum
            const float dm = f16_lo(mm);
            const float2 pj = sPar[t.j];                                     // (S, S max|c|)
            const float ad = fabsf(dm);
            const float ulp = ad >= 6.1035156e-5f ? __uint_as_float((__float_as_uint(ad) & 0x7F800000u) - (10u << 23)) : 5.9604645e-8f;
            float sq;
            asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(fmaxf(dm, 0.f)));
            const float xb = pj.y + sq * 1.001f + 1e-3f;                     // |S x| <= |S c*| + S d
            const float eps = 1.2e-5f * (xb + pj.y) * (xb + pj.y) + 1e-6f;
            const float thr = dm + 1.5f * ulp + 2.f * eps;
            unsigned short t16;
            asm("cvt.rp.f16.f32 %0, %1;" : "=h"(t16) : "f"(thr));            // rounded UP
            t.trust = dm < kE16Limit;
            t.th2 = (uint32_t)t16 * 0x00010001u;
        };
        // piece 2: the 0/1 indicators v = (V <= thr) of the 48 minima (packed compare) and their bit masks sum v 2^i (HFMA2)
        auto tail2 = [&](const uint32_t (&A)[16], const uint32_t (&B)[8], Tail& t) {
            uint32_t accA0 = 0u, accA1 = 0u, accB = 0u;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t w = (uint32_t)(0x3C00u + (i << 10)) * 0x00010001u;    // (2^i, 2^i)
                accA0 = hfma2u(hle2u(A[i], t.th2), w, accA0);
                accA1 = hfma2u(hle2u(A[i + 8], t.th2), w, accA1);
                accB = hfma2u(hle2u(B[i], t.th2), w, accB);
            }
            const uint32_t k1024 = 0x64006400u;                             // + 1024: the integer sits in the low mantissa bits
            const uint32_t iA0 = hadd2u(accA0, k1024) & 0x00FF00FFu, iA1 = hadd2u(accA1, k1024) & 0x00FF00FFu;
            const uint32_t iB = hadd2u(accB, k1024) & 0x00FF00FFu;
            t.mA = t.trust ? (iA0 | (iA1 << 8)) : 0xFFFFFFFFu;              // bit (16 p + a): minimum (block a, parity p) within the margin
            t.mB = t.trust ? ((iB & 0xFFu) | ((iB >> 8) & 0xFF00u)) : 0xFFFFu;   // bit (8 p + i)
        };
        // piece 3: the code, or the item for the resolver (the only branch, warp-uniform)
        auto tail3 = [&](Tail& t) {
            const bool uni = __popc(t.mA) == 1 && __popc(t.mB) == 1;
            const int ta = __ffs(t.mA) - 1, tb = __ffs(t.mB) - 1;
            if (t.live && uni) *t.out = (uint8_t)(((ta & 15) << 4) + ((tb & 7) << 1) + (ta >> 4));
            const bool amb = t.live && !uni;
            const uint32_t bal = __ballot_sync(0xffffffffu, amb);
            if (bal) {
                // several candidates (or an untrusted minimum): one reservation per warp, then every lane fills its slot.
                // The resolver never waits for this warp, so waiting for a free slot cannot deadlock.
                const int leader = __ffs(bal) - 1;
                uint32_t h = 0u;
                if (lane == leader) h = atomicAdd(const_cast<uint32_t*>(q_head), (uint32_t)__popc(bal));
                h = __shfl_sync(0xffffffffu, h, leader);
                if (amb) {
                    const uint32_t seq = h + (uint32_t)__popc(bal & ((1u << lane) - 1u));
                    while (seq - *q_tail >= (uint32_t)kE16QCap) {}
                    volatile uint4* slot = reinterpret_cast<volatile uint4*>(sQ + (seq % kE16QCap));
                    slot->x = t.row; slot->y = t.mA; slot->z = t.mB | ((uint32_t)t.j << 16);
                    __threadfence_block();
                    slot->w = seq + 1u;
                }
            }
        };
        // Group eg takes the CTA's steps s = eg, eg + 3, ...; step s lives in accumulator stage s & 1 (its use number s >> 1
        // gives the barrier parity).  A group releases each half of the stage as soon as the half is in registers, so the
        // stage is refilled for step s + 2 (another group's) while this group is still in its tail.
        Tail t;
        t.th2 = 0u; t.mA = t.mB = 1u; t.trust = false;
        int64_t round = blockIdx.x;
        int j = 0, u = eg;                             // step = (round, sub-space j, tile u of the round)
        while (u >= T) { u -= T; ++j; }
        while (j >= m) { j -= m; round += gridDim.x; }
        for (uint32_t sidx = (uint32_t)eg; round < rounds; sidx += (uint32_t)kE16Groups) {
            const int st = (int)(sidx & 1u);
            const uint32_t use = sidx >> 1, ph = (use >> 3) & 1u;      // use number of the stage; its barrier serves every eighth use
            const uint32_t ta = taddr0 + (uint32_t)(st * kE16K);
            uint32_t va[16], vb[16], A[16], B[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) B[i] = kInf2;
            mbar_wait(&tmem_full[st * 8 + (int)(use & 7u)], ph);
            tc_fence_after();
            // eight 32-column chunks, two 16-register buffers (tcgen05.wait::ld waits for ALL outstanding loads, so exactly
            // one is outstanding at every wait)
            tmem_ld_x16_pack16(ta, va);
            tmem_ld_wait_u(va); tmem_ld_x16_pack16(ta + 32u, vb); fold_chunk16<0>(va, A, B);
            tmem_ld_wait_u(vb); tmem_ld_x16_pack16(ta + 64u, va); fold_chunk16<1>(vb, A, B);
            tmem_ld_wait_u(va); tmem_ld_x16_pack16(ta + 96u, vb); fold_chunk16<2>(va, A, B);
            tmem_ld_wait_u(vb); tmem_ld_x16_pack16(ta + 128u, va); fold_chunk16<3>(vb, A, B);
            tmem_ld_wait_u(va); tmem_ld_x16_pack16(ta + 160u, vb); fold_chunk16<4>(va, A, B);
            tmem_ld_wait_u(vb); tmem_ld_x16_pack16(ta + 192u, va); fold_chunk16<5>(vb, A, B);
            tmem_ld_wait_u(va); tmem_ld_x16_pack16(ta + 224u, vb); fold_chunk16<6>(va, A, B);
            tmem_ld_wait_u(vb);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[st]);   // the step's 256 columns are in this warp's registers
            fold_chunk16<7>(vb, A, B);
            const uint32_t row = (uint32_t)((round * T + u) * kE16Rows + rt);    // N < 2^31
            t.row = row;
            t.j = j;
            t.live = (int64_t)row < N;
            t.out = codes + (size_t)row * m + j;
            tail1(A, t);
            tail2(A, B, t);
            tail3(t);
            u += kE16Groups;
            while (u >= T) { u -= T; ++j; }
            while (j >= m) { j -= m; round += gridDim.x; }
        }
        __threadfence_block();
        named_bar_sync(1 + eg, 128);
        if (rt == 0) atomicAdd(const_cast<uint32_t*>(q_done), 1u);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

}  // namespace vdb

using namespace vdb;

namespace vdb {

// Workspace (shared with pq_tc.cu, vdb_pq_encode_tc_workspace): [0, 16 KB m) operands, [16 KB m, +8 m) par,
// last 256 bytes: three 64-bit counters (items the resolver took, of which all-ones, items resolved in place: queue full).
int g_pq16_tiles = 4;     // row tiles per round (even): the codebook operands of a sub-space are fetched once per round
int g_pq16_xhint = 1;     // L2 hint of the row loads: 1 = evict normal, 0 = evict first (a 32-byte row slice shares its 64-byte DRAM
                          // burst with the next sub-space's slice: evict-first lost it before its turn, 3.8 / 5.0 / 7.3 GB read per
                          // 2M x 384 rows at 2 / 4 / 8 tiles per round against the algorithmic 3.07 GB; evict-normal reads 3.08 GB)

int pq_encode_tc16_launch(const float* x, int64_t N, int D, int m, int K, const float* codebooks, uint8_t* codes, char* base,
                          size_t operand_slab, size_t stats_off, cudaStream_t st) {
    uint32_t* opnd = reinterpret_cast<uint32_t*>(base);
    float2* par = reinterpret_cast<float2*>(base + 2 * operand_slab);
    unsigned long long* stats = reinterpret_cast<unsigned long long*>(base + stats_off);
    VDB_CUDA(cudaMemsetAsync(stats, 0, 24, st));
    pq_tc16_operands_kernel<<<m, 256, 0, st>>>(codebooks, m, K, opnd, par);
    count_launch(1);
    CUtensorMap map_x, map_b;
    int rc = tma_map_2d(&map_x, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, x, (uint64_t)D, (uint64_t)N, (uint64_t)D * 4, 8, kE16Rows,
                        CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = tma_map_2d(&map_b, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, opnd, 16, (uint64_t)m * 2 * 256, 32, 16, kE16K, CU_TENSOR_MAP_SWIZZLE_32B);
    if (rc) return rc;
    // ring depths: 4 codebook slots (8 steps), 10 operand slots, as many row slots as shared memory allows
    const int smem_max = max_smem_optin();
    int nbs = 4, na = 10, nx = kE16MaxX;
    while (nx > 8 && e16_layout(nx, na, nbs, m).total + 1024 > smem_max) --nx;
    while (na > 4 && e16_layout(nx, na, nbs, m).total + 1024 > smem_max) --na;
    while (nbs > 2 && e16_layout(nx, na, nbs, m).total + 1024 > smem_max) --nbs;
    VDB_REQUIRE(e16_layout(nx, na, nbs, m).total + 1024 <= smem_max, VDB_ERR_UNSUPPORTED, "m=%d does not fit shared memory", m);
    const size_t smem = (size_t)e16_layout(nx, na, nbs, m).total;
    const int64_t tiles = (N + kE16Rows - 1) / kE16Rows;
    const int T = g_pq16_tiles;
    const int64_t rounds = (tiles + T - 1) / T;
    const int grid = (int)(rounds < sm_count() ? rounds : sm_count());
    VDB_CUDA(cudaFuncSetAttribute(pq_encode_tc16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    {
        ProfScope prof(kProfPqEncode, st);
        pq_encode_tc16_kernel<<<grid, kE16Threads, smem, st>>>(map_x, map_b, x, N, D, m, K, codebooks, par, nx, na, nbs, rounds, codes, stats, T, g_pq16_xhint ? kEvictNormal : kEvictFirst);
    }
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // namespace vdb
