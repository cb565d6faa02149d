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

namespace vdb {

// ------------------------------------------------------------------------------------------
// numpy pairwise_sum order over n rounded squares v[t] = rn(d[t]*d[t])   (n <= 128)
template <int DS>
__device__ __forceinline__ float np_sumsq(const float (&d)[DS]) {
    if (DS < 8) {
        float r = 0.f;
#pragma unroll
        for (int t = 0; t < DS; ++t) r = __fadd_rn(r, __fmul_rn(d[t], d[t]));
        return r;
    } else {
        float r[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) r[t] = __fmul_rn(d[t], d[t]);
        constexpr int main_end = DS - (DS % 8);
#pragma unroll
        for (int i = 8; i < main_end; i += 8) {
#pragma unroll
            for (int t = 0; t < 8; ++t) r[t] = __fadd_rn(r[t], __fmul_rn(d[i + t], d[i + t]));
        }
        float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                              __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
#pragma unroll
        for (int i = main_end; i < DS; ++i) res = __fadd_rn(res, __fmul_rn(d[i], d[i]));
        return res;
    }
}

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
// ADC scan.  One persistent CTA walks work units (db slice, query group of G).  The group's LUTs
// live in shared memory INTERLEAVED by query -- lutS[(j*256 + c)*G + g] -- so one 128-bit shared
// load (G=4) serves four queries per code byte.  One thread = one row: it loads the row's m code
// bytes with 128-/32-bit loads, walks the sub-spaces in order (fp32 accumulation order j=0..m-1,
// identical to the oracle) and pushes (score,row) keys that beat the query's running threshold
// into that query's BlockSelect.  Units are ordered slice-major so the CTAs resident at any time
// stream the same code slice out of L2.
constexpr int kAdcThreads = 256;

template <int G>
struct LutVec;
template <> struct LutVec<1> { typedef float type; };
template <> struct LutVec<2> { typedef float2 type; };
template <> struct LutVec<4> { typedef float4 type; };

template <int G>
__device__ __forceinline__ void lut_add(float (&acc)[G], const float* lutS, int entry) {
    typedef typename LutVec<G>::type V;
    const V v = *reinterpret_cast<const V*>(lutS + (size_t)entry * G);
    const float* pv = reinterpret_cast<const float*>(&v);
#pragma unroll
    for (int g = 0; g < G; ++g) acc[g] += pv[g];
}

// MW = m/4 when m is a multiple of 4 and small enough to unroll, 0 = generic byte walk.
template <int G, int MW>
__global__ void __launch_bounds__(kAdcThreads)
adc_scan_kernel(const uint8_t* __restrict__ codes, int64_t N, int m, const float* __restrict__ lut, int Q, int R,
                int C, int64_t slice_rows, int S, int groups, u64* __restrict__ partial) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* sel_buf = reinterpret_cast<u64*>(smem_raw);                 // G * C
    u64* sel_thr = sel_buf + (size_t)G * C;                          // G (padded to 2)
    float* lutS = reinterpret_cast<float*>(sel_thr + ((G + 1) & ~1));// m*256*G
    int* sel_cnt = reinterpret_cast<int*>(lutS + (size_t)m * 256 * G);
    __shared__ unsigned compact_mask;
    const int tid = threadIdx.x;
    const int units = S * groups;

    for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int s = u / groups, grp = u - s * groups;
        const int q0 = grp * G;
        const int nq = min(G, Q - q0);
        __syncthreads();   // previous unit fully drained
        // stage the group's LUTs, interleaved by query
        for (int g = 0; g < G; ++g) {
            if (g < nq) {
                const float* src = lut + (size_t)(q0 + g) * m * 256;
                for (int i = tid; i < m * 256; i += kAdcThreads) lutS[(size_t)i * G + g] = src[i];
            } else {
                for (int i = tid; i < m * 256; i += kAdcThreads) lutS[(size_t)i * G + g] = 0.f;
            }
        }
        for (int i = tid; i < G * C; i += kAdcThreads) sel_buf[i] = VDB_KEY_MAX;
        if (tid < G) { sel_cnt[tid] = 0; sel_thr[tid] = VDB_KEY_MAX; }
        __syncthreads();

        const int64_t r_begin = (int64_t)s * slice_rows;
        const int64_t r_end = min(N, r_begin + slice_rows);
        for (int64_t rb = r_begin; rb < r_end; rb += kAdcThreads) {
            const int64_t r = rb + tid;
            if (r < r_end) {
                float acc[G];
#pragma unroll
                for (int g = 0; g < G; ++g) acc[g] = 0.f;
                if (MW > 0) {
                    uint32_t w[MW > 0 ? MW : 1];
                    if ((MW & 3) == 0) {
                        const uint4* p = reinterpret_cast<const uint4*>(codes + r * m);
#pragma unroll
                        for (int i = 0; i < MW / 4; ++i) {
                            uint4 v = __ldg(p + i);
                            w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
                        }
                    } else {
                        const uint32_t* p = reinterpret_cast<const uint32_t*>(codes + r * m);
#pragma unroll
                        for (int i = 0; i < MW; ++i) w[i] = __ldg(p + i);
                    }
#pragma unroll
                    for (int i = 0; i < MW; ++i) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) {
                            const int c = (w[i] >> (8 * b)) & 0xFF;
                            lut_add<G>(acc, lutS, (4 * i + b) * 256 + c);
                        }
                    }
                } else {
                    const uint8_t* p = codes + r * m;
                    for (int j = 0; j < m; ++j) lut_add<G>(acc, lutS, j * 256 + (int)__ldg(p + j));
                }
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    if (g < nq) {
                        u64 key = pack_key(acc[g], (uint32_t)r);
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

struct AdcPlan {
    int G, C, S, groups, grid;
    int64_t slice_rows;
    size_t smem;
};

static bool plan_adc(int Q, int64_t N, int m, int R, AdcPlan* p) {
    const size_t smem_max = (size_t)max_smem_optin() - 64;
    const int C = select_capacity(R, kAdcThreads);          // >= R + 2*256
    const int gs[3] = {4, 2, 1};
    for (int t = 0; t < 3; ++t) {
        const int G = gs[t];
        if (G > 1 && Q < G && t < 2 && Q <= gs[t + 1]) continue;   // do not waste LUT space on absent queries
        size_t smem = (size_t)G * C * 8 + (size_t)((G + 1) & ~1) * 8 + (size_t)m * 256 * G * 4 + (size_t)G * 4 + 16;
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
static int launch_adc(const AdcPlan& p, const uint8_t* codes, int64_t N, int m, const float* lut, int Q, int R,
                      u64* partial, cudaStream_t st) {
#define VDB_ADC_CASE(MWV)                                                                                       \
    {                                                                                                           \
        auto kern = adc_scan_kernel<G, MWV>;                                                                    \
        VDB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));         \
        kern<<<p.grid, kAdcThreads, p.smem, st>>>(codes, N, m, lut, Q, R, p.C, p.slice_rows, p.S, p.groups, partial); \
    }
    ProfScope prof(kProfAdcScan, st);
    const bool aligned4 = (m % 4 == 0) && ((reinterpret_cast<uintptr_t>(codes) & 15) == 0);
    if (aligned4 && m == 16) VDB_ADC_CASE(4)
    else if (aligned4 && m == 32) VDB_ADC_CASE(8)
    else if (aligned4 && m == 48) VDB_ADC_CASE(12)
    else if (aligned4 && m == 64) VDB_ADC_CASE(16)
    else if (aligned4 && m == 8) VDB_ADC_CASE(2)
    else if (aligned4 && m == 24) VDB_ADC_CASE(6)
    else if (aligned4 && m == 96) VDB_ADC_CASE(24)
    else VDB_ADC_CASE(0)
#undef VDB_ADC_CASE
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

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
    return (size_t)Q * p.S * R * 8 + 256;
}

int vdb_pq_adc_topk(const uint8_t* codes, int64_t N, int m, const float* lut, int Q, int R, const int64_t* ids,
                    float* out_d, int64_t* out_i, void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(Q >= 0 && N >= 0 && m > 0 && R > 0, VDB_ERR_INVALID, "vdb_pq_adc_topk: bad shape");
    VDB_REQUIRE(N < (1ll << 32), VDB_ERR_UNSUPPORTED, "N=%lld exceeds 2^32 rows per call", (long long)N);
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(lut && out_d && out_i && (codes || N == 0), VDB_ERR_INVALID, "vdb_pq_adc_topk: null pointer");
    AdcPlan p;
    VDB_REQUIRE(plan_adc(Q, N, m, R, &p), VDB_ERR_UNSUPPORTED, "m=%d R=%d do not fit shared memory", m, R);
    size_t need = (size_t)Q * p.S * R * 8;
    VDB_REQUIRE(ws && ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    u64* partial = static_cast<u64*>(ws);
    int rc;
    if (p.G == 4) rc = launch_adc<4>(p, codes, N, m, lut, Q, R, partial, st);
    else if (p.G == 2) rc = launch_adc<2>(p, codes, N, m, lut, Q, R, partial, st);
    else rc = launch_adc<1>(p, codes, N, m, lut, Q, R, partial, st);
    if (rc) return rc;
    return finalize_keys(partial, Q, p.S * R, R, ids, 0, 0, out_d, out_i, st);
}

}  // extern "C"
