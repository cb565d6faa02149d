// Product quantisation: encode (a11), Lloyd step (a12), LUT build and ADC scan (a13).
//
// Reference arithmetic replaced:
//   encode  src/vector_db/inference/pq.py:77-89  argmin_c sqrt(sum((C_j[c]-x_j)^2)), first min wins.
//           numpy evaluates sum() of the ROUNDED squares with its pairwise_sum order (8 strided
//           accumulators + a fixed tree for 8 <= n <= 128, a plain loop below 8); np_sumsq() below
//           reproduces that order with non-fused multiplies so codes are bit-exact, not just close.
//   train   pq.py:7-12 (scipy kmeans2, iter=100, minit='points'); values are parity-unpinned
//           (SURVEY.md 8c), the step kernel restates one Lloyd round.
//   ADC     no reference symbol; sub-space layout pq.py:48-62, squared form of pq.py:77-80.
#include "common.cuh"
#include "pq_common.cuh"

namespace vdb {

// generic-dsub variant working from memory (x in global/L1, centroid row in shared)
__device__ __forceinline__ float np_sumsq_mem(const float* __restrict__ c, const float* __restrict__ x, int n) {
    if (n < 8) {
        float r = 0.f;
        for (int t = 0; t < n; ++t) { float d = c[t] - x[t]; r = __fadd_rn(r, __fmul_rn(d, d)); }
        return r;
    }
    float r[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) { float d = c[t] - x[t]; r[t] = __fmul_rn(d, d); }
    const int main_end = n - (n % 8);
    for (int i = 8; i < main_end; i += 8) {
#pragma unroll
        for (int t = 0; t < 8; ++t) { float d = c[i + t] - x[i + t]; r[t] = __fadd_rn(r[t], __fmul_rn(d, d)); }
    }
    float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                          __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
    for (int i = main_end; i < n; ++i) { float d = c[i] - x[i]; res = __fadd_rn(res, __fmul_rn(d, d)); }
    return res;
}

constexpr int kEncThreads = 256;

// One thread = one row; sub-spaces are walked in order with the sub-space codebook staged in
// shared memory (broadcast reads).  DS = dsub when it is one of the register-resident sizes,
// 0 = generic (row slice re-read from L1 per centroid).
template <int DS>
__global__ void __launch_bounds__(kEncThreads)
pq_encode_kernel(const float* __restrict__ x, int64_t N, int D, int m, int K, int dsub,
                 const float* __restrict__ cb, uint8_t* __restrict__ codes) {
    extern __shared__ __align__(16) float cbs[];   // K * dsub
    const int64_t tiles = (N + kEncThreads - 1) / kEncThreads;
    for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int64_t row = tile * kEncThreads + threadIdx.x;
        for (int j = 0; j < m; ++j) {
            __syncthreads();
            const float* src = cb + (size_t)j * K * dsub;
            for (int i = threadIdx.x; i < K * dsub; i += kEncThreads) cbs[i] = src[i];
            __syncthreads();
            if (row < N) {
                const float* xr = x + row * D + (size_t)j * dsub;
                float best = INFINITY;
                int bi = 0;
                if (DS > 0) {
                    float xs[DS > 0 ? DS : 1];
#pragma unroll
                    for (int t = 0; t < DS; ++t) xs[t] = xr[t];
                    for (int c = 0; c < K; ++c) {
                        float d[DS > 0 ? DS : 1];
#pragma unroll
                        for (int t = 0; t < DS; ++t) d[t] = cbs[c * DS + t] - xs[t];
                        float dist = sqrtf(np_sumsq<(DS > 0 ? DS : 1)>(d));
                        if (dist < best) { best = dist; bi = c; }
                    }
                } else {
                    for (int c = 0; c < K; ++c) {
                        float dist = sqrtf(np_sumsq_mem(cbs + (size_t)c * dsub, xr, dsub));
                        if (dist < best) { best = dist; bi = c; }
                    }
                }
                codes[row * m + j] = (uint8_t)bi;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// Lloyd step.  grid = (row blocks, m).  Assignment uses the squared direct-difference distance
// (first minimum wins); per-CTA sums/counts are accumulated in shared memory and flushed with
// global atomics; kmeans_update_kernel then divides (an empty cluster keeps its centroid).
__global__ void __launch_bounds__(256)
kmeans_assign_kernel(const float* __restrict__ x, int64_t N, int D, int m, int K, int dsub,
                     const float* __restrict__ cb, int32_t* __restrict__ labels,
                     float* __restrict__ sums, float* __restrict__ counts) {
    extern __shared__ __align__(16) float sm[];
    float* cbs = sm;                    // K*dsub
    float* ssum = cbs + (size_t)K * dsub;   // K*dsub
    float* scnt = ssum + (size_t)K * dsub;  // K
    const int j = blockIdx.y;
    const float* src = cb + (size_t)j * K * dsub;
    for (int i = threadIdx.x; i < K * dsub; i += blockDim.x) { cbs[i] = src[i]; ssum[i] = 0.f; }
    for (int i = threadIdx.x; i < K; i += blockDim.x) scnt[i] = 0.f;
    __syncthreads();
    for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < N; row += (int64_t)gridDim.x * blockDim.x) {
        const float* xr = x + row * D + (size_t)j * dsub;
        float best = INFINITY;
        int bi = 0;
        for (int c = 0; c < K; ++c) {
            float a = 0.f;
            for (int t = 0; t < dsub; ++t) { float d = xr[t] - cbs[c * dsub + t]; a = fmaf(d, d, a); }
            if (a < best) { best = a; bi = c; }
        }
        if (labels) labels[(size_t)j * N + row] = bi;
        for (int t = 0; t < dsub; ++t) atomicAdd(ssum + bi * dsub + t, xr[t]);
        atomicAdd(scnt + bi, 1.f);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < K * dsub; i += blockDim.x)
        if (ssum[i] != 0.f) atomicAdd(sums + (size_t)j * K * dsub + i, ssum[i]);
    for (int i = threadIdx.x; i < K; i += blockDim.x)
        if (scnt[i] != 0.f) atomicAdd(counts + (size_t)j * K + i, scnt[i]);
}

__global__ void kmeans_update_kernel(float* __restrict__ cb, const float* __restrict__ sums,
                                     const float* __restrict__ counts, int total, int dsub) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < total) {
        float n = counts[i / dsub];
        if (n > 0.f) cb[i] = sums[i] / n;
    }
}

// ------------------------------------------------------------------------------------------
// LUT[q][j][c] = sum_t (q_jt - C_j[c][t])^2, sequential non-fused fp32 (the oracle's order).
__global__ void pq_lut_kernel(const float* __restrict__ q, int D, int m, int K, int dsub,
                              const float* __restrict__ cb, float* __restrict__ lut) {
    const int qi = blockIdx.x, j = blockIdx.y;
    const float* qs = q + (size_t)qi * D + (size_t)j * dsub;
    for (int c = threadIdx.x; c < K; c += blockDim.x) {
        const float* cr = cb + ((size_t)j * K + c) * dsub;
        float a = 0.f;
        for (int t = 0; t < dsub; ++t) { float d = qs[t] - cr[t]; a = __fadd_rn(a, __fmul_rn(d, d)); }
        lut[((size_t)qi * m + j) * K + c] = a;
    }
}

}  // namespace vdb

namespace vdb {

// ------------------------------------------------------------------------------------------
// ADC scan, fixed-point.
//
// Semantics (mirrored by oracle.adc_scan_topk):  per query, e = 23 - exponent(max LUT entry) and
// lutq = rint(lut * 2^e) (uint32, < 2^23 per entry);  d_int(q,i) = sum_j lutq[q][j][code_ij] -- an
// EXACT integer sum, so the order in which sub-spaces are visited cannot change the result;
// results are the R smallest by (d_int, row), reported as float(d_int) * 2^-e.  That freedom is
// what makes the scan fast: shared-memory lookups of random codes are bank-conflict bound (the fp32
// j-ascending kernel measured 10.6 wavefronts per 128-bit lookup instead of 4), and the fix is to let
// lane l walk the sub-spaces of each block of 8 in the rotated order j = (t + l) mod 8 over a LUT
// laid out so that sub-space j only occupies 16-byte bank group j mod 8 -- the 8 lanes of every
// quarter-warp phase then hit 8 different bank groups whatever their codes are.
//
// One persistent CTA walks work units (db slice, query group of G).  The group's LUTs live in
// shared memory as slots of G uint32 -- slot((j,c)) = ((j/8)*256 + c)*8 + (j%8) -- so one 128-bit
// shared load (G=4) serves four queries per code byte.  One thread = one row: 128-bit code loads,
// (score,row) keys that beat the query's running threshold go to that query's BlockSelect.  Units
// are ordered slice-major so the CTAs resident at any time stream the same code slice out of L2.
constexpr int kAdcThreads = 384;

template <int G>
struct LutVec;
template <> struct LutVec<1> { typedef uint32_t type; };
template <> struct LutVec<2> { typedef uint2 type; };
template <> struct LutVec<4> { typedef uint4 type; };

template <int G>
__device__ __forceinline__ void lut_add(uint32_t (&acc)[G], const uint32_t* lutS, uint32_t slot) {
    typedef typename LutVec<G>::type V;
    const V v = *reinterpret_cast<const V*>(lutS + (size_t)slot * G);
    const uint32_t* pv = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
    for (int g = 0; g < G; ++g) acc[g] += pv[g];
}

// 64-bit rotate right by 8*r bits (r = 0..7) of the little-endian byte string (lo, hi)
__device__ __forceinline__ void rotr_bytes(uint32_t& lo, uint32_t& hi, int r) {
    if (r & 4) { const uint32_t t = lo; lo = hi; hi = t; }
    const int sh = (r & 3) * 8;
    const uint32_t nlo = __funnelshift_r(lo, hi, sh);
    const uint32_t nhi = __funnelshift_r(hi, lo, sh);
    lo = nlo; hi = nhi;
}

// per query: e = 23 - exponent(max entry); lutq = rint(lut * 2^e).  One CTA per query.
__global__ void __launch_bounds__(256)
adc_quantize_lut_kernel(const float* __restrict__ lut, int per_query, uint32_t* __restrict__ lutq, int32_t* __restrict__ expo) {
    __shared__ float red[8];
    __shared__ int e_s;
    const float* src = lut + (size_t)blockIdx.x * per_query;
    float mx = 0.f;
    for (int i = threadIdx.x; i < per_query; i += blockDim.x) mx = fmaxf(mx, src[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        float m = red[0];
        for (int w = 1; w < 8; ++w) m = fmaxf(m, red[w]);
        int ex = 0;
        if (m > 0.f && m < INFINITY) frexpf(m, &ex); else ex = 23;
        e_s = 23 - ex;
        expo[blockIdx.x] = e_s;
    }
    __syncthreads();
    const int e = e_s;
    uint32_t* dst = lutq + (size_t)blockIdx.x * per_query;
    for (int i = threadIdx.x; i < per_query; i += blockDim.x) {
        const float v = rintf(ldexpf(src[i], e));
        dst[i] = v >= 8388608.f ? 8388608u : (v > 0.f ? (uint32_t)v : 0u);
    }
}

// MB = m/8 when m is a multiple of 8 and small enough to unroll, 0 = generic (any m, byte loads).
template <int G, int MB>
__global__ void __launch_bounds__(kAdcThreads)
adc_scan_kernel(const uint8_t* __restrict__ codes, int64_t N, int m, const uint32_t* __restrict__ lutq, int Q, int R,
                int C, int64_t slice_rows, int S, int groups, u64* __restrict__ partial) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int mb = (m + 7) >> 3;                                     // blocks of 8 sub-spaces (last one zero padded)
    u64* sel_buf = reinterpret_cast<u64*>(smem_raw);                 // G * C
    u64* sel_thr = sel_buf + (size_t)G * C;                          // G (padded to 2)
    uint32_t* lutS = reinterpret_cast<uint32_t*>(sel_thr + ((G + 1) & ~1));   // mb*256*8 slots of G
    int* sel_cnt = reinterpret_cast<int*>(lutS + (size_t)mb * 2048 * G);
    __shared__ unsigned compact_mask;
    const int tid = threadIdx.x;
    const int rot = tid & 7;
    const int units = S * groups;

    for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int s = u / groups, grp = u - s * groups;
        const int q0 = grp * G;
        const int nq = min(G, Q - q0);
        __syncthreads();   // previous unit fully drained
        // stage the group's LUTs into the rotated-slot layout
        for (int i = tid; i < mb * 2048; i += kAdcThreads) {
            const int j = i >> 8, c = i & 255;
            const uint32_t slot = (uint32_t)(((j >> 3) * 256 + c) * 8 + (j & 7));
#pragma unroll
            for (int g = 0; g < G; ++g)
                lutS[(size_t)slot * G + g] = (g < nq && j < m) ? lutq[((size_t)(q0 + g) * m + j) * 256 + c] : 0u;
        }
        for (int i = tid; i < G * C; i += kAdcThreads) sel_buf[i] = VDB_KEY_MAX;
        if (tid < G) { sel_cnt[tid] = 0; sel_thr[tid] = VDB_KEY_MAX; }
        __syncthreads();

        const int64_t r_begin = (int64_t)s * slice_rows;
        const int64_t r_end = min(N, r_begin + slice_rows);
        for (int64_t rb = r_begin; rb < r_end; rb += kAdcThreads) {
            const int64_t r = rb + tid;
            if (r < r_end) {
                uint32_t acc[G];
#pragma unroll
                for (int g = 0; g < G; ++g) acc[g] = 0u;
                if (MB > 0) {
                    uint32_t w[MB > 0 ? 2 * MB : 2];
                    if ((MB & 1) == 0) {
                        const uint4* p = reinterpret_cast<const uint4*>(codes + r * m);
#pragma unroll
                        for (int i = 0; i < MB / 2; ++i) {
                            const uint4 v = ld_stream_u4(p + i);
                            w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
                        }
                    } else {
                        const uint2* p = reinterpret_cast<const uint2*>(codes + r * m);
#pragma unroll
                        for (int i = 0; i < MB; ++i) { const uint2 v = __ldg(p + i); w[2 * i] = v.x; w[2 * i + 1] = v.y; }
                    }
#pragma unroll
                    for (int b = 0; b < MB; ++b) {
                        uint32_t lo = w[2 * b], hi = w[2 * b + 1];
                        rotr_bytes(lo, hi, rot);                         // byte t of (lo,hi) is now sub-space (t + rot) & 7
#pragma unroll
                        for (int t = 0; t < 8; ++t) {
                            const uint32_t c = ((t < 4 ? lo : hi) >> (8 * (t & 3))) & 0xFFu;
                            lut_add<G>(acc, lutS, (uint32_t)(b * 2048) + c * 8u + (uint32_t)((t + rot) & 7));
                        }
                    }
                } else {
                    const uint8_t* p = codes + r * m;
                    for (int b = 0; b < mb; ++b) {
                        for (int t = 0; t < 8; ++t) {
                            const int jj = (t + rot) & 7, j = b * 8 + jj;
                            const uint32_t c = j < m ? (uint32_t)__ldg(p + j) : 0u;   // padded sub-spaces hold zeros
                            lut_add<G>(acc, lutS, (uint32_t)(b * 2048) + c * 8u + (uint32_t)jj);
                        }
                    }
                }
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    if (g < nq) {
                        const u64 key = ((u64)acc[g] << 32) | (u64)(uint32_t)r;
                        if (key < *(volatile u64*)(sel_thr + g)) {
                            int p = atomicAdd(sel_cnt + g, 1);
                            sel_buf[(size_t)g * C + R + p] = key;
                        }
                    }
                }
            }
            // compaction check (CTA-uniform): at most kAdcThreads pushes per query since the last one
            __syncthreads();
            if (tid == 0) {
                unsigned mk = 0;
                for (int g = 0; g < G; ++g)
                    if (sel_cnt[g] + kAdcThreads > C - R) mk |= 1u << g;
                compact_mask = mk;
            }
            __syncthreads();
            const unsigned mk = compact_mask;
            if (mk) {
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    if (mk & (1u << g)) {
                        block_bitonic_sort(sel_buf + (size_t)g * C, C);
                        if (tid == 0) { sel_cnt[g] = 0; sel_thr[g] = sel_buf[(size_t)g * C + R - 1]; }
                        __syncthreads();
                    }
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int g = 0; g < G; ++g) {
            if (g < nq) {
                block_bitonic_sort(sel_buf + (size_t)g * C, C);
                u64* dst = partial + ((size_t)(q0 + g) * S + s) * R;
                for (int i = tid; i < R; i += kAdcThreads) dst[i] = sel_buf[(size_t)g * C + i];
                __syncthreads();
            }
        }
    }
}

// R best of L fixed-point keys per query -> (float(d_int) * 2^-e, id).  One CTA per query.
__global__ void __launch_bounds__(256)
adc_finalize_kernel(const u64* __restrict__ keys, int L, int R, int C, const int32_t* __restrict__ expo,
                    const int64_t* __restrict__ ids, float* __restrict__ out_d, int64_t* __restrict__ out_i) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* buf = reinterpret_cast<u64*>(smem_raw);
    __shared__ int cnt;
    __shared__ u64 thr;
    BlockSelect sel;
    sel.init(buf, &cnt, &thr, C, R, blockDim.x);
    __syncthreads();
    const u64* src = keys + (size_t)blockIdx.x * L;
    for (int base = 0; base < L; base += blockDim.x) {
        int i = base + threadIdx.x;
        if (i < L) {
            u64 key = src[i];
            if (key != VDB_KEY_MAX) sel.push(key);
        }
        sel.maybe_compact();
    }
    sel.finalize();
    const int e = expo[blockIdx.x];
    for (int i = threadIdx.x; i < R; i += blockDim.x) {
        const u64 key = buf[i];
        const size_t o = (size_t)blockIdx.x * R + i;
        if (key == VDB_KEY_MAX) {
            out_d[o] = INFINITY;
            out_i[o] = -1;
        } else {
            out_d[o] = ldexpf((float)(uint32_t)(key >> 32), -e);
            const int64_t idx = (int64_t)(uint32_t)(key & 0xFFFFFFFFull);
            out_i[o] = ids ? ids[idx] : idx;
        }
    }
}

struct AdcPlan {
    int G, C, S, groups, grid;
    int64_t slice_rows;
    size_t smem;
};

static bool plan_adc(int Q, int64_t N, int m, int R, AdcPlan* p) {
    const size_t smem_max = (size_t)max_smem_optin() - 64;
    // pending area of at least one batch of pushes; twice that when it is cheap (fewer compactions)
    int C = next_pow2(R + kAdcThreads);
    if (C < 512) C = 512;
    const int mb = (m + 7) / 8;
    const int gs[3] = {4, 2, 1};
    for (int t = 0; t < 3; ++t) {
        const int G = gs[t];
        if (G > 1 && Q < G && t < 2 && Q <= gs[t + 1]) continue;   // do not waste LUT space on absent queries
        size_t smem = (size_t)G * C * 8 + (size_t)((G + 1) & ~1) * 8 + (size_t)mb * 2048 * G * 4 + (size_t)G * 4 + 16;
        if (smem > smem_max) continue;
        p->G = G; p->C = C; p->smem = smem;
        p->groups = (Q + G - 1) / G;
        // slices: ~16 MB of codes each (L2-resident while every CTA streams it), at least enough
        // units to fill the machine, at most what keeps S*R partial keys per query reasonable.
        int64_t S = (N * (int64_t)m + (16ll << 20) - 1) / (16ll << 20);
        int64_t fill = (2ll * sm_count() + p->groups - 1) / p->groups;
        if (S < fill) S = fill;
        int64_t max_by_rows = (N + 4095) / 4096;
        if (max_by_rows < 1) max_by_rows = 1;
        if (S > max_by_rows) S = max_by_rows;
        int64_t max_by_keys = (1ll << 18) / R;
        if (max_by_keys < 1) max_by_keys = 1;
        if (S > max_by_keys) S = max_by_keys;
        if (S < 1) S = 1;
        p->slice_rows = (N + S - 1) / S;
        if (p->slice_rows < 1) p->slice_rows = 1;
        p->S = (int)((N + p->slice_rows - 1) / p->slice_rows);
        if (p->S < 1) p->S = 1;
        int64_t units = (int64_t)p->S * p->groups;
        p->grid = (int)(units < sm_count() ? units : sm_count());
        // more than one CTA per SM when shared memory allows it
        int per_sm = (int)(smem_max / smem);
        if (per_sm > 4) per_sm = 4;
        if (per_sm > 1) {
            int64_t g2 = (int64_t)sm_count() * per_sm;
            p->grid = (int)(units < g2 ? units : g2);
        }
        return true;
    }
    return false;
}

template <int G>
static int launch_adc(const AdcPlan& p, const uint8_t* codes, int64_t N, int m, const uint32_t* lutq, int Q, int R,
                      u64* partial, cudaStream_t st) {
#define VDB_ADC_CASE(MBV)                                                                                       \
    {                                                                                                           \
        auto kern = adc_scan_kernel<G, MBV>;                                                                    \
        VDB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));         \
        kern<<<p.grid, kAdcThreads, p.smem, st>>>(codes, N, m, lutq, Q, R, p.C, p.slice_rows, p.S, p.groups, partial); \
    }
    ProfScope prof(kProfAdcScan, st);
    const bool aligned16 = (reinterpret_cast<uintptr_t>(codes) & 15) == 0;
    if (aligned16 && m == 16) VDB_ADC_CASE(2)
    else if (aligned16 && m == 32) VDB_ADC_CASE(4)
    else if (aligned16 && m == 48) VDB_ADC_CASE(6)
    else if (aligned16 && m == 64) VDB_ADC_CASE(8)
    else if (aligned16 && m == 96) VDB_ADC_CASE(12)
    else if ((reinterpret_cast<uintptr_t>(codes) & 7) == 0 && m == 8) VDB_ADC_CASE(1)
    else if ((reinterpret_cast<uintptr_t>(codes) & 7) == 0 && m == 24) VDB_ADC_CASE(3)
    else VDB_ADC_CASE(0)
#undef VDB_ADC_CASE
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

static size_t adc_lutq_bytes(int Q, int m) { return (((size_t)Q * m * 256 * 4) + 255) & ~(size_t)255; }
static size_t adc_expo_bytes(int Q) { return (((size_t)Q * 4) + 255) & ~(size_t)255; }

}  // namespace vdb

using namespace vdb;

extern "C" {

int vdb_pq_encode(const float* x, int64_t N, int D, int m, int K, const float* codebooks, uint8_t* codes, void* stream) {
    VDB_REQUIRE(N >= 0 && D > 0 && m > 0 && K > 0, VDB_ERR_INVALID, "vdb_pq_encode: bad shape");
    VDB_REQUIRE(D % m == 0, VDB_ERR_INVALID, "dim must be divisible by chunks");
    VDB_REQUIRE(K <= 256, VDB_ERR_UNSUPPORTED, "K=%d > 256 does not fit uint8 codes", K);
    const int dsub = D / m;
    VDB_REQUIRE(dsub <= 128, VDB_ERR_UNSUPPORTED, "dsub=%d > 128", dsub);
    if (N == 0) return VDB_OK;
    VDB_REQUIRE(x && codebooks && codes, VDB_ERR_INVALID, "vdb_pq_encode: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    size_t smem = (size_t)K * dsub * 4;
    int64_t tiles = (N + kEncThreads - 1) / kEncThreads;
    int grid = (int)(tiles < 8ll * sm_count() ? tiles : 8ll * sm_count());
#define VDB_ENC_CASE(DSV)                                                                                  \
    {                                                                                                      \
        VDB_CUDA(cudaFuncSetAttribute(pq_encode_kernel<DSV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        pq_encode_kernel<DSV><<<grid, kEncThreads, smem, st>>>(x, N, D, m, K, dsub, codebooks, codes);      \
    }
    ProfScope prof(kProfPqEncode, st);
    switch (dsub) {
        case 1: VDB_ENC_CASE(1) break;
        case 2: VDB_ENC_CASE(2) break;
        case 4: VDB_ENC_CASE(4) break;
        case 8: VDB_ENC_CASE(8) break;
        case 16: VDB_ENC_CASE(16) break;
        default: VDB_ENC_CASE(0) break;
    }
#undef VDB_ENC_CASE
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

size_t vdb_pq_kmeans_workspace(int m, int K, int dsub) { return (size_t)m * K * (dsub + 1) * 4; }

int vdb_pq_kmeans_step(const float* x, int64_t N, int D, int m, int K, float* codebooks, int32_t* labels, void* ws,
                       size_t ws_bytes, void* stream) {
    VDB_REQUIRE(N > 0 && D > 0 && m > 0 && K > 0 && D % m == 0, VDB_ERR_INVALID, "vdb_pq_kmeans_step: bad shape");
    const int dsub = D / m;
    VDB_REQUIRE(x && codebooks, VDB_ERR_INVALID, "vdb_pq_kmeans_step: null pointer");
    size_t need = vdb_pq_kmeans_workspace(m, K, dsub);
    VDB_REQUIRE(ws && ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
    size_t smem = ((size_t)2 * K * dsub + K) * 4;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "K*dsub too large for shared memory");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* sums = static_cast<float*>(ws);
    float* counts = sums + (size_t)m * K * dsub;
    VDB_CUDA(cudaMemsetAsync(ws, 0, need, st));
    VDB_CUDA(cudaFuncSetAttribute(kmeans_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t rb = (N + 255) / 256;
    int gx = (int)(rb < 2ll * sm_count() ? rb : 2ll * sm_count());
    if (gx * m < sm_count() && rb > gx) gx = (int)(rb < sm_count() ? rb : sm_count());
    kmeans_assign_kernel<<<dim3(gx, m), 256, smem, st>>>(x, N, D, m, K, dsub, codebooks, labels, sums, counts);
    int total = m * K * dsub;
    kmeans_update_kernel<<<(total + 255) / 256, 256, 0, st>>>(codebooks, sums, counts, total, dsub);
    count_launch(2);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

int vdb_pq_lut(const float* q, int Q, int D, int m, int K, const float* codebooks, float* lut, void* stream) {
    VDB_REQUIRE(Q >= 0 && D > 0 && m > 0 && K > 0 && D % m == 0, VDB_ERR_INVALID, "vdb_pq_lut: bad shape");
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(q && codebooks && lut, VDB_ERR_INVALID, "vdb_pq_lut: null pointer");
    VDB_REQUIRE(m <= 65535, VDB_ERR_UNSUPPORTED, "m too large");
    pq_lut_kernel<<<dim3(Q, m), 256, 0, static_cast<cudaStream_t>(stream)>>>(q, D, m, K, D / m, codebooks, lut);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

size_t vdb_pq_adc_topk_workspace(int Q, int64_t N, int m, int R) {
    if (Q <= 0 || R <= 0 || m <= 0) return 0;
    AdcPlan p;
    if (!plan_adc(Q, N, m, R, &p)) return 0;
    return adc_lutq_bytes(Q, m) + adc_expo_bytes(Q) + (size_t)Q * p.S * R * 8 + 256;
}

int vdb_pq_adc_topk(const uint8_t* codes, int64_t N, int m, const float* lut, int Q, int R, const int64_t* ids,
                    float* out_d, int64_t* out_i, void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(Q >= 0 && N >= 0 && m > 0 && R > 0, VDB_ERR_INVALID, "vdb_pq_adc_topk: bad shape");
    VDB_REQUIRE(N < (1ll << 32), VDB_ERR_UNSUPPORTED, "N=%lld exceeds 2^32 rows per call", (long long)N);
    VDB_REQUIRE(m <= 255, VDB_ERR_UNSUPPORTED, "m=%d > 255 overflows the 32-bit fixed-point sum", m);
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(lut && out_d && out_i && (codes || N == 0), VDB_ERR_INVALID, "vdb_pq_adc_topk: null pointer");
    AdcPlan p;
    VDB_REQUIRE(plan_adc(Q, N, m, R, &p), VDB_ERR_UNSUPPORTED, "m=%d R=%d do not fit shared memory", m, R);
    const size_t need = adc_lutq_bytes(Q, m) + adc_expo_bytes(Q) + (size_t)Q * p.S * R * 8;
    VDB_REQUIRE(ws && ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
    VDB_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15) == 0, VDB_ERR_INVALID, "workspace must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* base = static_cast<char*>(ws);
    uint32_t* lutq = reinterpret_cast<uint32_t*>(base);
    int32_t* expo = reinterpret_cast<int32_t*>(base + adc_lutq_bytes(Q, m));
    u64* partial = reinterpret_cast<u64*>(base + adc_lutq_bytes(Q, m) + adc_expo_bytes(Q));
    adc_quantize_lut_kernel<<<Q, 256, 0, st>>>(lut, m * 256, lutq, expo);
    count_launch(1);
    int rc;
    if (p.G == 4) rc = launch_adc<4>(p, codes, N, m, lutq, Q, R, partial, st);
    else if (p.G == 2) rc = launch_adc<2>(p, codes, N, m, lutq, Q, R, partial, st);
    else rc = launch_adc<1>(p, codes, N, m, lutq, Q, R, partial, st);
    if (rc) return rc;
    const int L = p.S * R;
    const int C = select_capacity(R, 256);
    const size_t smem = (size_t)C * 8;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "R=%d too large for the final merge", R);
    VDB_CUDA(cudaFuncSetAttribute(adc_finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    {
        ProfScope prof(kProfFinalize, st);
        adc_finalize_kernel<<<Q, 256, smem, st>>>(partial, L, R, C, expo, ids, out_d, out_i);
    }
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"
