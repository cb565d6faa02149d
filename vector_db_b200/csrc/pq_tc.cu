// PQ encode on the tensor cores (a11): codes[i][j] = argmin_c sqrt(sum((C_j[c]-x_ij)^2)), first min wins
// (src/vector_db/inference/pq.py:77-89), for dsub = 8 and K <= 256.
//
// Per (128-row tile, sub-space j) the whole 128 x 256 score matrix  s(c) = |c|^2 - 2 x.c  is TWO
// tcgen05.mma.kind::f16 instructions (bf16 operands, K = 16 each, fp32 accumulation in TMEM).  bf16 keeps 8
// significant bits, so both operands are split into bf16-exact pieces:
//   x  = x1 + x2 + (rest <= 2^-18 |x|)        -- a helper warp converts the landed fp32 TMA tile
//   -2c = c1 + c2 + (rest <= 2^-18 |2c|)      -- prepared once per call
//   A0 = [x1 | x1]   B0 = [c1 | c2]           -> x1.c1 + x1.c2
//   A1 = [x2 | 1 1 1 0 0 0 0 0]   B1 = [c1 | n1 n2 n3 0 0 0 0 0]   -> x2.c1 + |c|^2   (n = 3-way bf16 split of |c|^2)
// Every product the MMA forms is exact in fp32; the dropped terms x1.c3 + x2.c2 + x2.c3 + x3.c are bounded by
// 3.01 * 2^-18 |x| |2c| = 2.3e-5 |x| |c| (Cauchy-Schwarz on the piece norms).  (An earlier version used four
// kind::tf32 instructions on a hi/lo split: twice the tensor time and twice the shared-memory operand traffic --
// K = 8 tf32 MMAs read 128 B of operands per clock, which starved the epilogue's own shared-memory reads.)
//
// The accumulator is the score.  The epilogue (one thread per row) makes ONE pass over the 256 scores and folds
// them into 16 block minima A[a] (centroids 16a..16a+15) and 16 residue minima B[b] (centroids = b mod 16), one
// FMNMX3 per score.  Two different centroids differ in a or in b, so "exactly one A and exactly one B within the
// margin of the row minimum" proves that exactly one centroid is within the margin and names it (16a + b); the
// margin test therefore runs over 32 minima, not 256 scores, and there is no second pass over TMEM.  The margin is
// RIGOROUS (5.2e-5 |x| max|c| for the split of both compared scores, plus 1e-5 (|x|+max|c|)^2 for the fp32
// accumulation and the reference's own fp32 rounding); rows with several candidates (the cross product of the
// two families, a superset) are resolved with the reference's exact arithmetic (numpy's pairwise-sum order,
// sqrt, first index wins) from the fp32 row still in shared memory and the fp32 codebook (global memory, L2 resident) -- so codes stay
// bit-exact although the heavy lifting is bf16.
//
// Measured (2M x 384 rows, m = 48, B200): 2.90 ms = 690 M vectors/s (m = 16, D = 128: 0.99 ms = 2.0 G vectors/s); ncu:
// ALU pipe 66 %, tensor pipe 25 %, issue slots 55 % -- the fold is one FMNMX3 per score on the half-rate ALU pipe
// (2 clocks per warp instruction), ~580 clocks of the ~1020 a step takes; the rest is the dependent tail of a step
// (norm, margin, the FMA-pipe margin test) and the MMA hand-over, with two epilogue warps per scheduler to hide them.
//
// Tried and measured slower or equal (kept out): a two-pass epilogue (row minimum, then a 256-bit candidate mask:
// 2.5 instructions per score, 5.2 ms); parking each lane's arg-min chunk in shared memory; a 16-warp epilogue with
// the columns of a row split over two warps; four kind::tf32 MMAs on a hi/lo split (3.3 ms: twice the tensor time,
// and K = 8 tf32 MMAs read 128 B of operands per clock, starving the epilogue's shared-memory reads); one 36 KB
// stage per step with the fp32 codebook slice in it (2.90 ms at 6 stages, 1.6 ms + 7.7 ms / stages below that --
// the pair-major order here reaches 10 steps of look-ahead and the same 2.90 ms: past 6 stages the ring is not the
// bound); operand slots released at MMA completion with 128-byte row boxes (3.4 ms); the epilogue groups preparing
// their own next-step operands instead of a split warp (3.6 ms: fence.proxy.async per thread and step); two split
// warps with the highest warp ids (2.96 ms); THREE epilogue groups rotating over the two accumulator stages (3.5 ms
// at 128 registers; needs a full-barrier per (step mod 6), a barrier per stage aliases use u with u + 2); ambiguous
// rows parked in shared memory and resolved once per tile (2.96 ms); r02: the tail of a step software-pipelined under
// the next step's first TMEM load (2.89 ms at 145 registers -- the tail's FMA-pipe work does not reach the idle slots
// of the ALU pipe, the other epilogue warp of the scheduler already fills what it can).
//
// Roofline: HBM, algorithmic bytes per vector = 4 D + m (rows read once, codes written once).
#include <cuda.h>
#include <math.h>

#include "common.cuh"
#include "pq_common.cuh"
#include "sm100.cuh"
#include "tma_host.cuh"

namespace vdb {

using namespace sm100;

constexpr int kEtThreads = 384;            // warp 0 TMA, 1-2 MMA issuers (2: TMEM alloc), 3 operand split, 4-7 / 8-11 epilogue groups
constexpr int kEtRows = 128;               // rows per tile (UMMA M)
constexpr int kEtK = 256;                  // centroids per sub-space (UMMA N)
constexpr int kEtXBytes = kEtRows * 32;    // 4 KB: 128 rows x 8 fp32 (TMA landing tile, not swizzled)
constexpr int kEtABytes = kEtRows * 32;    // 4 KB: 128 rows x 16 bf16 (32-byte swizzle)
constexpr int kEtBBytes = kEtK * 32;       // 8 KB: 256 centroids x 16 bf16 (32-byte swizzle)
constexpr int kEtXaBytes = kEtXBytes + 2 * kEtABytes;       // x | A0 | A1  (12 KB per step)
constexpr int kEtBsBytes = 2 * kEtBBytes;                   // B0 | B1  (16 KB per sub-space, shared by the two tiles of a pair)
constexpr int kEtMaxXa = 16, kEtMaxBs = 8;

// Work order: a CTA takes PAIRS of row tiles and walks the sub-spaces j-major over the pair -- (tile 0, j), (tile 1, j),
// (tile 0, j+1), ... -- so the codebook operands of sub-space j are fetched once for two steps.  Two rings: the ROW
// ring (fp32 row slice as TMA lands it + its two bf16 operand tiles, 12 KB per step) and the CODEBOOK ring (16 KB per
// sub-space, two steps each).  At the same shared memory the look-ahead is 10 steps instead of 6 -- the kernel's
// time follows the ring depth (1.6 ms + 7.7 ms / stages per 2M x 384 rows with one 36 KB stage per step) -- and the
// codebook traffic from L2 halves.  Step s uses accumulator stage s & 1 = its tile within the pair, so each
// epilogue group owns one tile of the pair and stages that tile's codes on its own.
struct EtSmem {
    int xa_off, bs_off, cmax_off, codes_off, bar_off, total;
};
__host__ __device__ inline EtSmem et_layout(int nxa, int nbs, int m) {
    EtSmem L;
    L.bs_off = 0;                                              // 16 KB slots first: 1024-byte aligned
    L.xa_off = nbs * kEtBsBytes;                               // 12 KB slots: 1024-byte aligned as well
    L.cmax_off = L.xa_off + nxa * kEtXaBytes;
    L.codes_off = L.cmax_off + ((m * 4 + 15) & ~15);
    L.bar_off = (L.codes_off + 2 * kEtRows * m + 15) & ~15;    // code staging [2 tiles][128 rows][m]
    L.total = L.bar_off + (3 * kEtMaxXa + 2 * kEtMaxBs + 8) * 8 + 16;
    return L;
}

// two fp32 -> packed bf16x2 (round to nearest even); `lo` lands in the low half = the lower address
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ float bf16_lo(uint32_t p) { return __uint_as_float(p << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t p) { return __uint_as_float(p & 0xFFFF0000u); }
__device__ __forceinline__ float bf16_round(float v) { return bf16_lo(pack_bf16x2(v, 0.f)); }

// Operand tensor [m][2][256] rows of 16 bf16: part 0 = [c1 | c2], part 1 = [c1 | n1 n2 n3 0 0 0 0 0] with
// -2c = c1 + c2 + rest and |c|^2 = n1 + n2 + n3 + rest.  Padded to 256 centroids per sub-space (padding can never win).
__global__ void pq_tc_operands_kernel(const float* __restrict__ cb, int m, int K, uint32_t* __restrict__ ops,
                                      float* __restrict__ cmax) {
    const int j = blockIdx.x;
    __shared__ float red[256];
    float mx = 0.f;
    for (int c = threadIdx.x; c < 256; c += blockDim.x) {
        uint32_t* o0 = ops + (((size_t)j * 2 + 0) * 256 + c) * 8;      // 8 words = 16 bf16
        uint32_t* o1 = ops + (((size_t)j * 2 + 1) * 256 + c) * 8;
        float s = 0.f;
        uint32_t p1[4] = {0u, 0u, 0u, 0u}, p2[4] = {0u, 0u, 0u, 0u};
        if (c < K) {
            const float* src = cb + ((size_t)j * K + c) * 8;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const float v0 = src[2 * t], v1 = src[2 * t + 1], w0 = -2.f * v0, w1 = -2.f * v1;
                p1[t] = pack_bf16x2(w0, w1);
                p2[t] = pack_bf16x2(w0 - bf16_lo(p1[t]), w1 - bf16_hi(p1[t]));
                s = fmaf(v0, v0, s);
                s = fmaf(v1, v1, s);
            }
            mx = fmaxf(mx, s);
        } else {
            s = 3.0e38f;
        }
        const float n1 = bf16_round(s), r1 = s - n1, n2 = bf16_round(r1), n3 = bf16_round(r1 - n2);
#pragma unroll
        for (int t = 0; t < 4; ++t) { o0[t] = p1[t]; o0[4 + t] = p2[t]; o1[t] = p1[t]; }
        o1[4] = pack_bf16x2(n1, n2);
        o1[5] = pack_bf16x2(n3, 0.f);
        o1[6] = 0u;
        o1[7] = 0u;
    }
    red[threadIdx.x] = mx;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] = fmaxf(red[threadIdx.x], red[threadIdx.x + o]);
        __syncthreads();
    }
    if (threadIdx.x == 0) cmax[j] = sqrtf(red[0]) * 1.0000005f;     // |c| from the fma-order |c|^2, rounded up
}

__device__ __forceinline__ float min3f(float a, float b, float c) { return fminf(fminf(a, b), c); }   // one FMNMX3

// minimum of v[O .. O+15]: 8 instructions
template <int O>
__device__ __forceinline__ float min16(const float (&v)[32]) {
    const float a = min3f(v[O + 0], v[O + 1], v[O + 2]), b = min3f(v[O + 3], v[O + 4], v[O + 5]);
    const float c = min3f(v[O + 6], v[O + 7], v[O + 8]), d = min3f(v[O + 9], v[O + 10], v[O + 11]);
    const float e = min3f(v[O + 12], v[O + 13], v[O + 14]);
    return fminf(min3f(a, b, c), min3f(d, e, v[O + 15]));
}
__device__ __forceinline__ float min16(const float (&v)[16]) {
    const float a = min3f(v[0], v[1], v[2]), b = min3f(v[3], v[4], v[5]);
    const float c = min3f(v[6], v[7], v[8]), d = min3f(v[9], v[10], v[11]);
    const float e = min3f(v[12], v[13], v[14]);
    return fminf(min3f(a, b, c), min3f(d, e, v[15]));
}

// One 32-column chunk CH of a row's 256 scores folded into two families of minima (1 instruction per score):
//   A[a] = min over the 16 consecutive centroids 16a .. 16a+15          (a = 2 CH, 2 CH + 1)
//   B[b] = min over the 16 centroids with index = b (mod 16)            (running over all chunks)
// Two different centroids differ in a or in b, so "exactly one A and exactly one B within the margin" proves
// that exactly one centroid is within the margin -- and names it (16 a + b) -- without a second pass over the
// scores and without a per-score compare: the margin test runs over 32 minima instead of 256 scores.
template <int CH>
__device__ __forceinline__ void fold_chunk(const float (&v)[32], float (&A)[16], float (&B)[16]) {
    A[2 * CH] = min16<0>(v);
    A[2 * CH + 1] = min16<16>(v);
#pragma unroll
    for (int i = 0; i < 16; ++i) B[i] = min3f(B[i], v[i], v[i + 16]);
}

// bit i of the result: v[i] <= thr.  Written as setp + predicated or (2 SASS instructions per element);
// the plain C++ form compiles to FSETP + SEL + LOP3.
__device__ __forceinline__ uint32_t le_mask16(const float (&v)[16], float thr) {
    uint32_t mk[4] = {0u, 0u, 0u, 0u};          // four independent chains: the predicated ORs of one chain serialise
#pragma unroll
    for (int i = 0; i < 16; ++i)
        asm("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\t@p or.b32 %0, %0, %3;\n\t}"
            : "+r"(mk[i & 3]) : "f"(v[i]), "f"(thr), "r"(1u << i));
    return (mk[0] | mk[1]) | (mk[2] | mk[3]);
}

// sum over i of sat((v[i] - thr) * S) * (1 + 32 i), all on the FMA pipe (FFMA.SAT + FFMA per element, four chains)
__device__ __forceinline__ float gt_accumulate16(const float (&v)[16], float S, float nts) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        float u;
        asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(u) : "f"(v[i]), "f"(S), "f"(nts));
        acc[i & 3] = fmaf(u, (float)(1 + 32 * i), acc[i & 3]);
    }
    return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

__device__ __forceinline__ float4 lds128(uint32_t saddr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
    return v;
}

__global__ void __launch_bounds__(kEtThreads, 1)
pq_encode_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_b,
                    const float* __restrict__ x, int64_t N, int D, int m, int K, const float* __restrict__ cb,
                    const float* __restrict__ cmax, int nxa, int nbs, int64_t pairs, uint8_t* __restrict__ codes) {
    extern __shared__ __align__(1024) unsigned char smem[];
    if ((smem_u32(smem) & 1023u) != 0) __trap();
    const EtSmem L = et_layout(nxa, nbs, m);
    unsigned char* sXa = smem + L.xa_off;
    unsigned char* sBs = smem + L.bs_off;
    float* sCmax = reinterpret_cast<float*>(smem + L.cmax_off);
    uint8_t* sCodes = smem + L.codes_off;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    uint64_t* x_full = bars;                        // [nxa]: TMA landed the fp32 row slice
    uint64_t* a_full = bars + kEtMaxXa;             // [nxa]: A0 / A1 written by the split warp
    uint64_t* xa_empty = bars + 2 * kEtMaxXa;       // [nxa]: MMAs done with A0 / A1 + epilogue done with x
    uint64_t* b_full = bars + 3 * kEtMaxXa;         // [nbs]: TMA landed B0, B1
    uint64_t* b_empty = b_full + kEtMaxBs;          // [nbs]: the MMAs of BOTH steps of the sub-space are done
    uint64_t* tmem_full = b_empty + kEtMaxBs;       // [2 accumulator stages][2 column halves]
    uint64_t* tmem_empty = tmem_full + 4;           // [2][2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // constant half of every A1 row: (1,1,1,0,0,0,0,0) bf16 against the norm pieces; with the 32-byte swizzle the
    // two 16-byte halves of row r trade places when bit 2 of r is set
    for (int i = threadIdx.x; i < nxa * kEtRows; i += kEtThreads) {
        const int sidx = i / kEtRows, r = i % kEtRows;
        unsigned char* a1 = sXa + (size_t)sidx * kEtXaBytes + kEtXBytes + kEtABytes;
        *reinterpret_cast<uint4*>(a1 + r * 32 + (16 ^ (((r >> 2) & 1) << 4))) = make_uint4(0x3F803F80u, 0x00003F80u, 0u, 0u);
    }
    for (int i = threadIdx.x; i < m; i += kEtThreads) sCmax[i] = cmax[i];
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < nxa; ++i) {
            mbar_init(&x_full[i], 1); mbar_init(&a_full[i], 1); mbar_init(&xa_empty[i], 1 + 4);   // MMA commit + 4 epilogue warps
        }
        for (int i = 0; i < nbs; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 2); }
        for (int i = 0; i < 4; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 128); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    fence_proxy_async();                 // the generic-proxy writes of the A1 constants must be visible to the MMA (async proxy)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            int xs = 0, bs = 0;
            uint32_t xph = 0, bph = 0;
            for (int64_t pair = blockIdx.x; pair < pairs; pair += gridDim.x) {
                for (int j = 0; j < m; ++j) {
                    mbar_wait(&b_empty[bs], bph ^ 1);
                    unsigned char* bb = sBs + (size_t)bs * kEtBsBytes;
                    mbar_expect_tx(&b_full[bs], (uint32_t)kEtBsBytes);
                    tma_load_2d(bb, &map_b, &b_full[bs], 0, (j * 2) * 256, kEvictLast);
                    tma_load_2d(bb + kEtBBytes, &map_b, &b_full[bs], 0, (j * 2 + 1) * 256, kEvictLast);
                    if (++bs == nbs) { bs = 0; bph ^= 1; }
#pragma unroll
                    for (int t = 0; t < 2; ++t) {
                        mbar_wait(&xa_empty[xs], xph ^ 1);
                        mbar_expect_tx(&x_full[xs], (uint32_t)kEtXBytes);
                        // rows past N (a pair's missing second tile included) land as zeros and are never stored
                        tma_load_2d(sXa + (size_t)xs * kEtXaBytes, &map_x, &x_full[xs], j * 8, (int)((pair * 2 + t) * kEtRows), kEvictFirst);
                        if (++xs == nxa) { xs = 0; xph ^= 1; }
                    }
                }
            }
        }
    } else if (warp < 3) {
        // ================= MMA issuers: warp 1 owns accumulator stage 0 (tile 0 of every pair), warp 2 stage 1 (tile 1).
        // With ONE in-order issuer the MMAs of a step whose accumulator half was already free queued behind the
        // other stage's wait for its epilogue group. =================
        if (lane == 0) {
            // Each 256-column accumulator is produced and released in two 128-column halves: the epilogue hands half 0
            // back while it is still folding half 1, so the MMAs of the group's next step run under epilogue work and
            // only the short second half sits in the release -> MMA -> fold dependency chain.
            const int as = warp - 1;
            const uint32_t idesc = make_idesc_bf16(kEtRows, kEtK / 2);
            int xs = as % nxa, bs = 0, aph = 0;        // own steps s = as, as + 2, ...: row slot s mod nxa, codebook slot (s / 2) mod nbs
            uint32_t xph = 0, bph = 0;
            for (int64_t pair = blockIdx.x; pair < pairs; pair += gridDim.x) {
                for (int j = 0; j < m; ++j) {
                    mbar_wait(&b_full[bs], bph);
                    mbar_wait(&a_full[xs], xph);
                    const uint32_t a0 = smem_u32(sXa + (size_t)xs * kEtXaBytes) + kEtXBytes;
                    const uint32_t b0 = smem_u32(sBs + (size_t)bs * kEtBsBytes);
                    const uint64_t d_a0 = make_desc_k_sw32(a0), d_a1 = make_desc_k_sw32(a0 + kEtABytes);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        mbar_wait(&tmem_empty[as * 2 + h], aph ^ 1);
                        tc_fence_after();
                        const uint32_t d_tmem = tmem_base + (uint32_t)(as * kEtK + h * (kEtK / 2));
                        const uint32_t boff = (uint32_t)(h * (kEtBBytes / 2));       // centroids 128 h .. 128 h + 127
                        umma_bf16(d_tmem, d_a0, make_desc_k_sw32(b0 + boff), idesc, 0u);                   // x1.c1 + x1.c2
                        umma_bf16(d_tmem, d_a1, make_desc_k_sw32(b0 + kEtBBytes + boff), idesc, 1u);       // x2.c1 + |c|^2
                        umma_commit(&tmem_full[as * 2 + h]);
                    }
                    umma_commit(&xa_empty[xs]);
                    umma_commit(&b_empty[bs]);
                    aph ^= 1;
                    xs += 2;
                    if (xs >= nxa) { xs -= nxa; xph ^= 1; }
                    if (++bs == nbs) { bs = 0; bph ^= 1; }
                }
            }
        }
    } else if (warp == 3) {
        // ================= operand split (one row per lane, 32 rows per pass): the landed fp32 row ->
        // x1 = bf16(x), x2 = bf16(x - x1); A0 row = [x1 | x1], A1 row = [x2 | ones] =================
        int xs = 0;
        uint32_t xph = 0;
        for (int64_t pair = blockIdx.x; pair < pairs; pair += gridDim.x) {
            for (int sidx = 0; sidx < 2 * m; ++sidx) {
                mbar_wait(&x_full[xs], xph);
                unsigned char* base = sXa + (size_t)xs * kEtXaBytes;
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int r = it * 32 + lane;
                    const float4 a = *reinterpret_cast<const float4*>(base + r * 32);
                    const float4 b = *reinterpret_cast<const float4*>(base + r * 32 + 16);
                    uint4 p, q;
                    p.x = pack_bf16x2(a.x, a.y); p.y = pack_bf16x2(a.z, a.w); p.z = pack_bf16x2(b.x, b.y); p.w = pack_bf16x2(b.z, b.w);
                    q.x = pack_bf16x2(a.x - bf16_lo(p.x), a.y - bf16_hi(p.x));
                    q.y = pack_bf16x2(a.z - bf16_lo(p.y), a.w - bf16_hi(p.y));
                    q.z = pack_bf16x2(b.x - bf16_lo(p.z), b.y - bf16_hi(p.z));
                    q.w = pack_bf16x2(b.z - bf16_lo(p.w), b.w - bf16_hi(p.w));
                    unsigned char* a0 = base + kEtXBytes + r * 32;
                    *reinterpret_cast<uint4*>(a0) = p;
                    *reinterpret_cast<uint4*>(a0 + 16) = p;
                    *reinterpret_cast<uint4*>(a0 + kEtABytes + (((r >> 2) & 1) << 4)) = q;
                }
                fence_proxy_async();                   // generic-proxy writes -> visible to the MMA's async-proxy reads
                __syncwarp();
                if (lane == 0) mbar_arrive(&a_full[xs]);
                if (++xs == nxa) { xs = 0; xph ^= 1; }
            }
        }
    } else {
        // ================= epilogue: one thread per row; group eg owns accumulator stage eg = tile eg of every pair =================
        // The step's row slot stays resident until the epilogue releases it, so the exact fp32 row slice is read back
        // from shared memory; the fp32 centroids of the rare ambiguous row come from global memory (L2 resident).
        const int eg = (warp - 4) >> 2;
        const int ew = warp & 3;
        const int rt = ew * 32 + lane;                 // row within the tile == TMEM lane
        const int et = threadIdx.x - 128 - eg * 128;   // 0..127 within the group
        const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(eg * kEtK);
        uint8_t* sCodesG = sCodes + (size_t)eg * kEtRows * m;
        int xs = eg % nxa;                             // own steps s = eg, eg + 2, ...: row slot s mod nxa
        int phase = 0;
        for (int64_t pair = blockIdx.x; pair < pairs; pair += gridDim.x) {
            for (int j = 0; j < m; ++j) {
                const uint32_t sx = smem_u32(sXa + (size_t)xs * kEtXaBytes) + (uint32_t)(rt * 32);
                mbar_wait(&tmem_full[eg * 2], phase);  // columns 0..127 done => the step's operands were in shared memory too
                tc_fence_after();
                // ONE pass over the scores.  Two 32-column register buffers: while one chunk is folded the next one's
                // TMEM load is in flight (tcgen05.wait::ld waits for ALL outstanding loads, so exactly one is
                // outstanding at every wait).
                float va[32], vb[32], A[16], B[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) B[i] = INFINITY;
                tmem_ld_x32(taddr, va);
                const float4 xa = lds128(sx), xb = lds128(sx + 16);     // issued here, consumed after the fold
                tmem_ld_wait(va); tmem_ld_x32(taddr + 32u, vb); fold_chunk<0>(va, A, B);
                tmem_ld_wait(vb); tmem_ld_x32(taddr + 64u, va); fold_chunk<1>(vb, A, B);
                tmem_ld_wait(va); tmem_ld_x32(taddr + 96u, vb); fold_chunk<2>(va, A, B);
                tmem_ld_wait(vb);
                tc_fence_before();
                mbar_arrive(&tmem_empty[eg * 2]);              // columns 0..127 are in registers
                mbar_wait(&tmem_full[eg * 2 + 1], phase);
                phase ^= 1;
                tc_fence_after();
                tmem_ld_x32(taddr + 128u, va); fold_chunk<3>(vb, A, B);
                tmem_ld_wait(va); tmem_ld_x32(taddr + 160u, vb); fold_chunk<4>(va, A, B);
                tmem_ld_wait(vb); tmem_ld_x32(taddr + 192u, va); fold_chunk<5>(vb, A, B);
                tmem_ld_wait(va); tmem_ld_x32(taddr + 224u, vb); fold_chunk<6>(va, A, B);
                tmem_ld_wait(vb);
                tc_fence_before();
                mbar_arrive(&tmem_empty[eg * 2 + 1]);          // columns 128..255 are in registers
                fold_chunk<7>(vb, A, B);
                const float xs8[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
                float xx = 0.f;
#pragma unroll
                for (int t = 0; t < 8; ++t) xx = fmaf(xs8[t], xs8[t], xx);
                __syncwarp();
                if (lane == 0) mbar_arrive(&xa_empty[xs]);     // this warp no longer reads the step's row slice
                const float xn = sqrtf(xx) * 1.0000005f, cm_j = sCmax[j];
                // rigorous candidate margin (see file header): operand split error + fp32 accumulation / reference rounding
                const float thr = min16(B) + 5.2e-5f * xn * cm_j + 1e-5f * (xn + cm_j) * (xn + cm_j);
                // Margin test on the FMA pipe (the fold above owns the half-rate ALU pipe): u = sat((V - thr) * S) is
                // exactly 1 for V > thr and 0 for V <= thr once S >= 2 / ulp(thr); acc = sum u * (1 + 32 i) then holds
                // the count of minima outside the margin (low 5 bits) and the sum of their indices.
                const uint32_t tb = __float_as_uint(thr), te = (tb >> 23) & 0xFFu;
                const float S = __uint_as_float((279u - te) << 23), nts = -thr * S;
                const float accA = gt_accumulate16(A, S, nts), accB = gt_accumulate16(B, S, nts);
                const int ia = __float2int_rn(accA), ib = __float2int_rn(accB);
                int code = 0;
                if (te >= 27u && (ia & 31) == 15 && (ib & 31) == 15) {
                    code = (120 - (ia >> 5)) * 16 + 120 - (ib >> 5);      // the one index that is missing from each sum
                } else if (const uint32_t mA = le_mask16(A, thr), mB = le_mask16(B, thr); mA && mB) {
                    // several candidates (a superset: the cross product of the two families): the reference's own
                    // arithmetic decides (pq.py:79-80), first minimum wins
                    float best = INFINITY;
                    const float* cbj = cb + (size_t)j * K * 8;
                    for (uint32_t ra = mA; ra; ra &= ra - 1) {
                        const int c0 = (__ffs(ra) - 1) * 16;
                        for (uint32_t rb = mB; rb; rb &= rb - 1) {
                            const int c = c0 + __ffs(rb) - 1;
                            if (c < K) {
                                const float4* cp = reinterpret_cast<const float4*>(cbj + (size_t)c * 8);
                                const float4 ca = __ldg(cp), cc = __ldg(cp + 1);
                                float d[8] = {ca.x - xs8[0], ca.y - xs8[1], ca.z - xs8[2], ca.w - xs8[3],
                                              cc.x - xs8[4], cc.y - xs8[5], cc.z - xs8[6], cc.w - xs8[7]};
                                const float dist = sqrtf(np_sumsq<8>(d));
                                if (dist < best) { best = dist; code = c; }
                            }
                        }
                    }
                }
                sCodesG[rt * m + j] = (uint8_t)code;
                xs += 2;
                if (xs >= nxa) xs -= nxa;
            }
            // the tile's codes are one contiguous run of 128*m bytes in global memory
            named_bar_sync(1 + eg, 128);
            const int64_t row0 = (pair * 2 + eg) * kEtRows;
            if (row0 < N) {
                const int64_t nbytes = (N - row0 < kEtRows ? N - row0 : kEtRows) * m;
                uint8_t* dst = codes + row0 * m;
                if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0 && (reinterpret_cast<uintptr_t>(sCodesG) & 15) == 0) {
                    const int64_t n16 = nbytes >> 4;
                    for (int64_t i = et; i < n16; i += 128)
                        reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(sCodesG)[i];
                    for (int64_t i = (n16 << 4) + et; i < nbytes; i += 128) dst[i] = sCodesG[i];
                } else {
                    for (int64_t i = et; i < nbytes; i += 128) dst[i] = sCodesG[i];
                }
            }
            named_bar_sync(1 + eg, 128);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

}  // namespace vdb

using namespace vdb;

static size_t et_ws_operand(int m) { return (size_t)m * 256 * 8 * 4; }

namespace vdb {
int pq_encode_tc16_launch(const float* x, int64_t N, int D, int m, int K, const float* codebooks, uint8_t* codes, char* base,
                          size_t operand_slab, size_t stats_off, cudaStream_t st);     // pq_tc16.cu
extern int g_pq16_tiles, g_pq16_xhint;
}
// 1 (default) = f16-accumulator kernel (pq_tc16.cu), 0 = the fp32-accumulator kernel of this file
static int g_pq_tc_mode = 1;

extern "C" {

void vdb_debug_set_pq_tc_mode(int mode) {      // bit 0: kernel; bits 4-7: row tiles per round (0 = keep); bit 8: row loads evict-first
    g_pq_tc_mode = mode & 1;
    if ((mode >> 4) & 15) vdb::g_pq16_tiles = ((mode >> 4) & 15) & ~1;
    vdb::g_pq16_xhint = ((mode >> 8) & 1) ^ 1;      // bit 8 set: evict-first (the measured-worse setting)
}

// counters of the last f16-accumulator encode that used this workspace: out[0] = (row, sub-space) items the resolver warp
// decided, out[1] = of which with all-ones masks (row out of f16 range), out[2] = pairs resolved by the epilogue lane itself
// because the queue was full.  Synchronises the stream.
int vdb_debug_pq_tc_stats(const void* ws, int m, unsigned long long* out, void* stream) {
    VDB_REQUIRE(ws && out, VDB_ERR_INVALID, "vdb_debug_pq_tc_stats: null pointer");
    const size_t off = 3 * et_ws_operand(m) + (((size_t)m * 4 + 255) & ~(size_t)255);
    VDB_CUDA(cudaMemcpyAsync(out, static_cast<const char*>(ws) + off, 24, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
    VDB_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    return VDB_OK;
}

int vdb_pq_encode_tc_supported(int64_t N, int D, int m, int K) {
    return (N > 0 && N < (1ll << 31) && m > 0 && D == 8 * m && K > 0 && K <= 256 && m <= 512) ? 1 : 0;
}

size_t vdb_pq_encode_tc_workspace(int m) { return 3 * et_ws_operand(m) + (((size_t)m * 4 + 255) & ~(size_t)255) + 256; }

int vdb_pq_encode_tc(const float* x, int64_t N, int D, int m, int K, const float* codebooks, uint8_t* codes, void* ws,
                     size_t ws_bytes, void* stream) {
    VDB_REQUIRE(vdb_pq_encode_tc_supported(N, D, m, K), VDB_ERR_UNSUPPORTED,
                "tensor-core encode needs dsub == 8, K <= 256, N < 2^31 (N=%lld D=%d m=%d K=%d)", (long long)N, D, m, K);
    VDB_REQUIRE(x && codebooks && codes && ws, VDB_ERR_INVALID, "vdb_pq_encode_tc: null pointer");
    VDB_REQUIRE(ws_bytes >= vdb_pq_encode_tc_workspace(m), VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes,
                vdb_pq_encode_tc_workspace(m));
    VDB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(ws) & 255) == 0 &&
                    (reinterpret_cast<uintptr_t>(codebooks) & 15) == 0,
                VDB_ERR_INVALID, "rows / codebooks / workspace must be 16 / 16 / 256-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* base = static_cast<char*>(ws);
    if (g_pq_tc_mode == 1)
        return pq_encode_tc16_launch(x, N, D, m, K, codebooks, codes, base, et_ws_operand(m),
                                     3 * et_ws_operand(m) + (((size_t)m * 4 + 255) & ~(size_t)255), st);
    uint32_t* opnd = reinterpret_cast<uint32_t*>(base);
    float* cmax = reinterpret_cast<float*>(base + 3 * et_ws_operand(m));
    pq_tc_operands_kernel<<<m, 256, 0, st>>>(codebooks, m, K, opnd, cmax);
    count_launch(1);
    CUtensorMap map_x, map_b;
    int rc = tma_map_2d(&map_x, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, x, (uint64_t)D, (uint64_t)N, (uint64_t)D * 4, 8, kEtRows,
                        CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = tma_map_2d(&map_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, opnd, 16, (uint64_t)m * 2 * 256, 32, 16, kEtK, CU_TENSOR_MAP_SWIZZLE_32B);
    if (rc) return rc;
    // ring depths: 5 codebook slots (10 steps), as many row slots as shared memory allows
    const int smem_max = max_smem_optin();
    int nbs = 5, nxa = kEtMaxXa;
    while (nxa > 4 && et_layout(nxa, nbs, m).total + 1024 > smem_max) --nxa;
    while (nbs > 2 && et_layout(nxa, nbs, m).total + 1024 > smem_max) --nbs;
    VDB_REQUIRE(et_layout(nxa, nbs, m).total + 1024 <= smem_max, VDB_ERR_UNSUPPORTED, "m=%d does not fit shared memory", m);
    const size_t smem = (size_t)et_layout(nxa, nbs, m).total;
    const int64_t tiles = (N + kEtRows - 1) / kEtRows;
    const int64_t pairs = (tiles + 1) / 2;
    const int grid = (int)(pairs < sm_count() ? pairs : sm_count());
    VDB_CUDA(cudaFuncSetAttribute(pq_encode_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    {
        ProfScope prof(kProfPqEncode, st);
        pq_encode_tc_kernel<<<grid, kEtThreads, smem, st>>>(map_x, map_b, x, N, D, m, K, codebooks, cmax, nxa, nbs, pairs, codes);
    }
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"
