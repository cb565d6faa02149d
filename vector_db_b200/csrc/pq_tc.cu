// PQ encode on the tensor cores (a11): codes[i][j] = argmin_c sqrt(sum((C_j[c]-x_ij)^2)), first min wins
// (src/vector_db/inference/pq.py:77-89), for dsub = 8 and K <= 256.
//
// dsub = 8 is exactly one tf32 UMMA K step, so per (128-row tile, sub-space j) the whole 128 x 256
// score matrix  s(c) = |c|^2 - 2 x.c  is FOUR tcgen05.mma.kind::tf32 instructions.  The tensor core keeps
// 10 mantissa bits of each operand, so both operands are split into tf32-exact parts (3xTF32 style):
//   x = hi(x) + lo(x)        -- a helper warp rewrites the landed TMA tile in place and writes lo(x) beside it
//   -2c = hi(-2c) + lo(-2c)  -- prepared once per call
// and the score is hi(x).hi(-2c) + hi(x).lo(-2c) + lo(x).hi(-2c) + ones.split3(|c|^2): every operand the MMA
// sees is exactly representable, the dropped term and the re-rounding of the lo parts are ~2^-20 |x||c|.
// The accumulator is the score; the epilogue (one thread per row) takes its minimum, then collects
// every centroid whose score is within a RIGOROUS margin of it (1.2e-5 |x| max|c| for the split, plus
// 1e-5 (|x|+max|c|)^2 for the fp32 accumulation and the reference's own fp32 rounding); a single candidate
// is the answer, several candidates are resolved with the reference's exact arithmetic (numpy's pairwise-sum
// order, sqrt, first index wins) -- so codes stay bit-exact although the heavy lifting is tf32.  With a
// single-pass tf32 product the margin had to be 0.0081 |x| max|c| and a fifth of all rows took the slow path.
//
// Tried and measured slower (kept out): parking each lane's arg-min chunk in shared memory to skip the second
// TMEM pass (the predicated 128-bit stores cost the shared-memory bandwidth of the whole score tile), and a
// 16-warp epilogue with the columns of a row split over two warps (two pair barriers per step ate the gain).
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

constexpr int kEtThreads = 384;            // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 spare, 4-7 / 8-11 epilogue groups
constexpr int kEtRows = 128;               // rows per tile (UMMA M)
constexpr int kEtK = 256;                  // centroids per sub-space (UMMA N)
constexpr int kEtABytes = kEtRows * 32;    // 4 KB: 128 rows x 8 fp32
constexpr int kEtBBytes = kEtK * 32;       // 8 KB: 256 centroids x 8 fp32
constexpr int kEtStageBytes = 2 * kEtABytes + 3 * kEtBBytes;   // x | x_lo | hi(-2C) | lo(-2C) | norm split  (32 KB)
constexpr int kEtMaxStages = 8;

struct EtSmem {
    int stage_off, ones_off, codes_off, bar_off, total;
};
__host__ __device__ inline EtSmem et_layout(int stages, int m) {
    EtSmem L;
    L.stage_off = 0;
    L.ones_off = stages * kEtStageBytes;                       // 32 KB stages keep 1024-byte alignment
    L.codes_off = L.ones_off + kEtABytes;
    L.bar_off = (L.codes_off + kEtRows * m + 15) & ~15;        // code staging [128 rows][m]
    L.total = L.bar_off + (3 * kEtMaxStages + 8) * 8 + 16;
    return L;
}

// tf32 keeps 10 explicit mantissa bits: the tensor core ignores the low 13 bits of an fp32 operand
__device__ __forceinline__ float tf32_hi(float v) { return __uint_as_float(__float_as_uint(v) & 0xFFFFE000u); }

// Operand tensor [m][3][256][8]: part 0 = hi(-2C), part 1 = lo(-2C) = -2C - hi (exact), part 2 = three-way
// tf32 split of |c|^2 in columns 0..2.  Padded to 256 centroids per sub-space (padding can never win).
__global__ void pq_tc_operands_kernel(const float* __restrict__ cb, int m, int K, float* __restrict__ ops,
                                      float* __restrict__ cmax) {
    const int j = blockIdx.x;
    __shared__ float red[256];
    float mx = 0.f;
    for (int c = threadIdx.x; c < 256; c += blockDim.x) {
        float* oh = ops + (((size_t)j * 3 + 0) * 256 + c) * 8;
        float* ol = ops + (((size_t)j * 3 + 1) * 256 + c) * 8;
        float* on = ops + (((size_t)j * 3 + 2) * 256 + c) * 8;
        float s = 0.f;
        if (c < K) {
            const float* src = cb + ((size_t)j * K + c) * 8;
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const float v = src[t], w = -2.f * v, h = tf32_hi(w);
                oh[t] = h; ol[t] = w - h;
                s = fmaf(v, v, s);
            }
            mx = fmaxf(mx, s);
        } else {
#pragma unroll
            for (int t = 0; t < 8; ++t) { oh[t] = 0.f; ol[t] = 0.f; }
            s = 3.0e38f;
        }
        const float h = tf32_hi(s);
        const float r1 = s - h;
        const float mid = tf32_hi(r1);
        const float lo = r1 - mid;
        on[0] = h; on[1] = mid; on[2] = lo;
#pragma unroll
        for (int t = 3; t < 8; ++t) on[t] = 0.f;
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

__global__ void __launch_bounds__(kEtThreads, 1)
pq_encode_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_b,
                    const float* __restrict__ x, int64_t N, int D, int m,
                    int K, const float* __restrict__ cb, const float* __restrict__ cmax, int stages, int64_t tiles,
                    uint8_t* __restrict__ codes) {
    extern __shared__ __align__(1024) unsigned char smem[];
    if ((smem_u32(smem) & 1023u) != 0) __trap();
    const EtSmem L = et_layout(stages, m);
    unsigned char* sStage = smem + L.stage_off;
    float* sOnes = reinterpret_cast<float*>(smem + L.ones_off);
    uint8_t* sCodes = smem + L.codes_off;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    uint64_t* full = bars;                          // [stages]
    uint64_t* empty = bars + kEtMaxStages;          // [stages]
    uint64_t* xlo_full = bars + 2 * kEtMaxStages;   // [stages]: x_lo written by the split warp
    uint64_t* tmem_full = bars + 3 * kEtMaxStages;  // [2 accumulator stages][2 column halves]
    uint64_t* tmem_empty = tmem_full + 4;           // [2][2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // constant A tile of the norm MMA: every row is (1,1,1,0,0,0,0,0); with the 32-byte swizzle the two
    // 16-byte halves of row r trade places when bit 2 of r is set
    if (threadIdx.x < kEtRows) {
        const int r = threadIdx.x;
        const int first = ((r >> 2) & 1) ? 4 : 0;
#pragma unroll
        for (int t = 0; t < 8; ++t) sOnes[r * 8 + t] = (t >= first && t < first + 3) ? 1.f : 0.f;
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < stages; ++i) {
            mbar_init(&full[i], 1); mbar_init(&empty[i], 1 + 4); mbar_init(&xlo_full[i], 1);   // empty: MMA commit + 4 epilogue warps
        }
        for (int i = 0; i < 4; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 128); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    fence_proxy_async();                 // the generic-proxy writes of sOnes must be visible to the MMA (async proxy)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            int st = 0, ph = 0;
            for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
                for (int j = 0; j < m; ++j) {
                    mbar_wait(&empty[st], ph ^ 1);
                    unsigned char* base = sStage + (size_t)st * kEtStageBytes;
                    mbar_expect_tx(&full[st], (uint32_t)(kEtABytes + 3 * kEtBBytes));
                    tma_load_2d(base, &map_x, &full[st], j * 8, (int)(tile * kEtRows), kEvictFirst);
#pragma unroll
                    for (int part = 0; part < 3; ++part)
                        tma_load_2d(base + 2 * kEtABytes + part * kEtBBytes, &map_b, &full[st], 0, (j * 3 + part) * 256,
                                    kEvictLast);
                    if (++st == stages) { st = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            // Each 256-column accumulator is produced and released in two 128-column halves: the epilogue frees half 0
            // while it is still masking half 1, so the next-but-one step's MMAs start (and finish) under epilogue work
            // instead of after it -- with whole-stage hand-over the epilogue warps waited ~20 % of the time for tmem_full.
            const uint32_t idesc = make_idesc_tf32(kEtRows, kEtK / 2);
            const uint64_t d_ones = make_desc_k_sw32(smem_u32(sOnes));
            int st = 0, ph = 0, as = 0, aph = 0;
            for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
                for (int j = 0; j < m; ++j) {
                    mbar_wait(&full[st], ph);
                    mbar_wait(&xlo_full[st], ph);
                    const uint32_t base = smem_u32(sStage + (size_t)st * kEtStageBytes);
                    const uint64_t d_x = make_desc_k_sw32(base), d_xlo = make_desc_k_sw32(base + kEtABytes);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        mbar_wait(&tmem_empty[as * 2 + h], aph ^ 1);
                        tc_fence_after();
                        const uint32_t d_tmem = tmem_base + (uint32_t)(as * kEtK + h * (kEtK / 2));
                        const uint32_t boff = (uint32_t)(h * (kEtBBytes / 2));       // centroids 128 h .. 128 h + 127
                        const uint64_t d_hi = make_desc_k_sw32(base + 2 * kEtABytes + boff);
                        const uint64_t d_lo = make_desc_k_sw32(base + 2 * kEtABytes + kEtBBytes + boff);
                        // x.(-2c) to ~2^-20: hi(x).hi + hi(x).lo + lo(x).hi  (every operand is tf32-exact)
                        umma_tf32(d_tmem, d_x, d_hi, idesc, 0u);
                        umma_tf32(d_tmem, d_x, d_lo, idesc, 1u);
                        umma_tf32(d_tmem, d_xlo, d_hi, idesc, 1u);
                        umma_tf32(d_tmem, d_ones, make_desc_k_sw32(base + 2 * kEtABytes + 2 * kEtBBytes + boff), idesc, 1u);
                        umma_commit(&tmem_full[as * 2 + h]);
                    }
                    umma_commit(&empty[st]);
                    if (++st == stages) { st = 0; ph ^= 1; }
                    as ^= 1;
                    if (as == 0) aph ^= 1;
                }
            }
        }
    } else if (warp == 3) {
        // ================= operand split: x -> hi(x) in place, x_lo = x - hi(x) beside it; elementwise on the landed
        // tile, so the swizzled layout does not matter =================
        int st = 0, ph = 0;
        for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
            for (int j = 0; j < m; ++j) {
                mbar_wait(&full[st], ph);
                float4* src_w = reinterpret_cast<float4*>(sStage + (size_t)st * kEtStageBytes);
                const float4* src = src_w;
                float4* dst = reinterpret_cast<float4*>(sStage + (size_t)st * kEtStageBytes + kEtABytes);
#pragma unroll
                for (int i = 0; i < kEtABytes / 16 / 32; ++i) {
                    const float4 v = src[i * 32 + lane];
                    const float4 h = make_float4(tf32_hi(v.x), tf32_hi(v.y), tf32_hi(v.z), tf32_hi(v.w));
                    src_w[i * 32 + lane] = h;           // operands are tf32-exact: no dependence on the MMA's input rounding mode
                    dst[i * 32 + lane] = make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w);
                }
                fence_proxy_async();                   // generic-proxy writes -> visible to the MMA's async-proxy reads
                __syncwarp();
                if (lane == 0) mbar_arrive(&xlo_full[st]);
                if (++st == stages) { st = 0; ph ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ================= epilogue: one thread per row; group eg owns accumulator stage eg (every other step) =================
        // The step's operand stage stays resident until the epilogue releases it, so the exact fp32 row
        // slice and the centroids (-2c, exact) are read back from shared memory, never from global.
        const int eg = (warp - 4) >> 2;
        const int ew = warp & 3;
        const int rt = ew * 32 + lane;                 // row within the tile == TMEM lane
        const int et = threadIdx.x - 128;              // 0..255
        const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(eg * kEtK);
        const int swz = ((rt >> 2) & 1) << 4;          // 32-byte swizzle: 16-byte halves of row r swap when bit 2 of r is set
        int64_t step = 0;
        int st = 0, phase = 0;
        for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
            for (int j = 0; j < m; ++j, ++step, st = (st + 1 == stages ? 0 : st + 1)) {
                if ((int)(step & 1) != eg) continue;
                mbar_wait(&tmem_full[eg * 2], phase);  // columns 0..127 done => the step's operands were in shared memory too
                tc_fence_after();
                const unsigned char* sA = sStage + (size_t)st * kEtStageBytes;
                const unsigned char* sBh = sA + 2 * kEtABytes;      // hi(-2c)
                const unsigned char* sBl = sBh + kEtBBytes;         // lo(-2c)
                float xs[8];
                {
                    float4 a = *reinterpret_cast<const float4*>(sA + rt * 32 + (0 ^ swz));
                    float4 b = *reinterpret_cast<const float4*>(sA + rt * 32 + (16 ^ swz));
                    const float4 al = *reinterpret_cast<const float4*>(sA + kEtABytes + rt * 32 + (0 ^ swz));
                    const float4 bl = *reinterpret_cast<const float4*>(sA + kEtABytes + rt * 32 + (16 ^ swz));
                    a.x += al.x; a.y += al.y; a.z += al.z; a.w += al.w;      // hi + lo = x exactly
                    b.x += bl.x; b.y += bl.y; b.z += bl.z; b.w += bl.w;
                    xs[0] = a.x; xs[1] = a.y; xs[2] = a.z; xs[3] = a.w; xs[4] = b.x; xs[5] = b.y; xs[6] = b.z; xs[7] = b.w;
                }
                float xx = 0.f;
#pragma unroll
                for (int t = 0; t < 8; ++t) xx = fmaf(xs[t], xs[t], xx);
                const float xn = sqrtf(xx) * 1.0000005f, cm_j = __ldg(cmax + j);
                // rigorous candidate margin (see file header): operand split error + fp32 accumulation / reference rounding
                const float margin = 1.2e-5f * xn * cm_j + 1e-5f * (xn + cm_j) * (xn + cm_j);
                // ONE pass over the scores.  Two 32-column register buffers: while one chunk is folded the next one's
                // TMEM load is in flight (tcgen05.wait::ld waits for ALL outstanding loads, so exactly one is
                // outstanding at every wait).  Each accumulator half goes back to the MMA as soon as its last chunk
                // is in registers.
                float va[32], vb[32], A[16], B[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) B[i] = INFINITY;
                tmem_ld_x32(taddr, va);
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
                const float thr = min16(B) + margin;
                const uint32_t mA = le_mask16(A, thr), mB = le_mask16(B, thr);
                int code = 0;
                if (__popc(mA) == 1 && __popc(mB) == 1) {
                    code = (__ffs(mA) - 1) * 16 + __ffs(mB) - 1;
                } else if (mA && mB) {
                    // several candidates (a superset: the cross product of the two families): the reference's own
                    // arithmetic decides (pq.py:79-80), first minimum wins
                    float best = INFINITY;
                    for (uint32_t ra = mA; ra; ra &= ra - 1) {
                        const int c0 = (__ffs(ra) - 1) * 16;
                        for (uint32_t rb = mB; rb; rb &= rb - 1) {
                            const int c = c0 + __ffs(rb) - 1;
                            if (c < K) {
                                const int cs = ((c >> 2) & 1) << 4;
                                float4 a = *reinterpret_cast<const float4*>(sBh + c * 32 + (0 ^ cs));
                                float4 b = *reinterpret_cast<const float4*>(sBh + c * 32 + (16 ^ cs));
                                const float4 al = *reinterpret_cast<const float4*>(sBl + c * 32 + (0 ^ cs));
                                const float4 bl = *reinterpret_cast<const float4*>(sBl + c * 32 + (16 ^ cs));
                                // hi + lo = -2c exactly, so c = -0.5 * (hi + lo) exactly
                                a.x += al.x; a.y += al.y; a.z += al.z; a.w += al.w;
                                b.x += bl.x; b.y += bl.y; b.z += bl.z; b.w += bl.w;
                                float d[8] = {-0.5f * a.x - xs[0], -0.5f * a.y - xs[1], -0.5f * a.z - xs[2], -0.5f * a.w - xs[3],
                                              -0.5f * b.x - xs[4], -0.5f * b.y - xs[5], -0.5f * b.z - xs[6], -0.5f * b.w - xs[7]};
                                const float dist = sqrtf(np_sumsq<8>(d));
                                if (dist < best) { best = dist; code = c; }
                            }
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[st]);    // this warp no longer reads the operand stage
                sCodes[rt * m + j] = (uint8_t)code;
            }
            // the tile's codes are one contiguous run of 128*m bytes in global memory
            named_bar_sync(1, 256);
            const int64_t row0 = tile * kEtRows;
            const int64_t nbytes = (N - row0 < kEtRows ? N - row0 : kEtRows) * m;
            uint8_t* dst = codes + row0 * m;
            if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
                const int64_t n16 = nbytes >> 4;
                for (int64_t i = et; i < n16; i += 256)
                    reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(sCodes)[i];
                for (int64_t i = (n16 << 4) + et; i < nbytes; i += 256) dst[i] = sCodes[i];
            } else {
                for (int64_t i = et; i < nbytes; i += 256) dst[i] = sCodes[i];
            }
            named_bar_sync(1, 256);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

}  // namespace vdb

using namespace vdb;

static size_t et_ws_operand(int m) { return (size_t)m * 256 * 8 * 4; }

extern "C" {

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
    float* opnd = reinterpret_cast<float*>(base);
    float* cmax = reinterpret_cast<float*>(base + 3 * et_ws_operand(m));
    pq_tc_operands_kernel<<<m, 256, 0, st>>>(codebooks, m, K, opnd, cmax);
    count_launch(1);
    CUtensorMap map_x, map_b;
    int rc = tma_map_2d(&map_x, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, x, (uint64_t)D, (uint64_t)N, (uint64_t)D * 4, 8, kEtRows,
                        CU_TENSOR_MAP_SWIZZLE_32B);
    if (rc) return rc;
    rc = tma_map_2d(&map_b, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, opnd, 8, (uint64_t)m * 3 * 256, 32, 8, kEtK, CU_TENSOR_MAP_SWIZZLE_32B);
    if (rc) return rc;
    int stages = kEtMaxStages;
    const int smem_max = max_smem_optin();
    while (stages >= 2 && et_layout(stages, m).total + 1024 > smem_max) --stages;
    VDB_REQUIRE(stages >= 2, VDB_ERR_UNSUPPORTED, "m=%d does not fit shared memory", m);
    const size_t smem = (size_t)et_layout(stages, m).total;
    const int64_t tiles = (N + kEtRows - 1) / kEtRows;
    const int grid = (int)(tiles < sm_count() ? tiles : sm_count());
    VDB_CUDA(cudaFuncSetAttribute(pq_encode_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    {
        ProfScope prof(kProfPqEncode, st);
        pq_encode_tc_kernel<<<grid, kEtThreads, smem, st>>>(map_x, map_b, x, N, D, m, K, codebooks, cmax, stages, tiles,
                                                           codes);
    }
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"
