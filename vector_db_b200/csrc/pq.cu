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

// Same accumulation with the assignment given (uint8 codes from the encode kernels: argmin with the reference's
// arithmetic, first minimum wins) -- the Lloyd round then costs one tensor-core encode plus this pass.
__global__ void __launch_bounds__(256)
kmeans_accumulate_kernel(const float* __restrict__ x, int64_t N, int D, int m, int K, int dsub,
                         const uint8_t* __restrict__ codes, float* __restrict__ sums, float* __restrict__ counts) {
    extern __shared__ __align__(16) float sm[];
    float* ssum = sm;                       // K*dsub
    float* scnt = ssum + (size_t)K * dsub;  // K
    const int j = blockIdx.y;
    for (int i = threadIdx.x; i < K * dsub; i += blockDim.x) ssum[i] = 0.f;
    for (int i = threadIdx.x; i < K; i += blockDim.x) scnt[i] = 0.f;
    __syncthreads();
    for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < N; row += (int64_t)gridDim.x * blockDim.x) {
        const float* xr = x + row * D + (size_t)j * dsub;
        const int bi = codes[row * m + j];
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

struct AdcArgs {
    const uint8_t* codes;
    const uint32_t* mask;       // optional row bitmap (bit cleared = row skipped): tombstones / metadata pre-filter
    int64_t N;                  // logical rows scanned
    int64_t row_stride;         // physical row = logical row * row_stride (> 1: strided sample for the threshold estimate)
    int m;
    const uint32_t* lutq;
    int Q, R, C;
    int64_t slice_rows;
    int S, groups;
    u64* partial;               // select mode: [query slot][S][R]
    const int32_t* qsel;        // optional indirection: only queries qsel[0 .. *qcount) are processed,
    const int32_t* qcount;      //   query slot b = position in qsel (device-side count: no host round trip)
    const uint32_t* tau;        // collect mode: per-query fixed-point threshold
    u64* cbuf;                  //   [Q][ccap] keys with d_int <= tau, unordered
    int32_t* ccnt;              //   [Q] number of keys found (may exceed ccap: overflow, handled by the caller)
    int ccap;
    const uint16_t* lut16;      // coarse collect: 15-bit lower-bound tables, [group][slot][G] in the shared-memory order
    const uint32_t* thr16;      //   per query 0x8000 + coarse threshold (0 = absent query)
};

// d_int of one row for the G queries of the group: 128-bit code loads, rotated conflict-free lookups.
// MB = m/8 when m is a multiple of 8 and small enough to unroll, 0 = generic (any m, byte loads).
// the row's code bytes as 2*MB words (MB > 0 only): 128-bit streaming loads when MB is even
// STREAM: bypass L1 (the full scan reads every code once per CTA).  A lane's 16-byte pieces sit 8*MB bytes apart, so
// each load instruction touches a 32-byte sector per lane for 16 useful bytes; without L1 the IVF-PQ scan (one
// query per CTA: four times the code traffic per lookup of the grouped full scan) moved twice its code bytes out
// of L2 and ran at the L2's pace.
template <int MB, bool STREAM = true>
__device__ __forceinline__ void adc_load_codes(const uint8_t* __restrict__ prow, uint32_t (&w)[MB > 0 ? 2 * MB : 2]) {
    if ((MB & 1) == 0) {
        const uint4* p = reinterpret_cast<const uint4*>(prow);
#pragma unroll
        for (int i = 0; i < MB / 2; ++i) {
            const uint4 v = STREAM ? ld_stream_u4(p + i) : __ldg(p + i);
            w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
        }
    } else {
        const uint2* p = reinterpret_cast<const uint2*>(prow);
#pragma unroll
        for (int i = 0; i < MB; ++i) { const uint2 v = __ldg(p + i); w[2 * i] = v.x; w[2 * i + 1] = v.y; }
    }
}

// P = copies of every LUT row interleaved in shared memory (slot (b, c, copy) at ((b*256 + c)*P + copy)*8 + j');
// the caller passes lutS already advanced to its own copy
template <int G, int MB, int P = 1>
__device__ __forceinline__ void adc_sum_codes(const uint32_t (&w)[MB > 0 ? 2 * MB : 2], const uint32_t* lutS, int rot,
                                              uint32_t (&acc)[G]) {
#pragma unroll
    for (int g = 0; g < G; ++g) acc[g] = 0u;
#pragma unroll
    for (int b = 0; b < MB; ++b) {
        uint32_t lo = w[2 * b], hi = w[2 * b + 1];
        rotr_bytes(lo, hi, rot);                         // byte t of (lo,hi) is now sub-space (t + rot) & 7
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const uint32_t c = ((t < 4 ? lo : hi) >> (8 * (t & 3))) & 0xFFu;
            lut_add<G>(acc, lutS, (uint32_t)(b * 2048 * P) + c * (uint32_t)(8 * P) + (uint32_t)((t + rot) & 7));
        }
    }
}

template <int G, int MB>
__device__ __forceinline__ void adc_row(const uint8_t* __restrict__ prow, int m, int mb, const uint32_t* lutS, int rot,
                                        uint32_t (&acc)[G]) {
    if (MB > 0) {
        uint32_t w[MB > 0 ? 2 * MB : 2];
        adc_load_codes<MB>(prow, w);
        adc_sum_codes<G, MB>(w, lutS, rot, acc);
    } else {
#pragma unroll
        for (int g = 0; g < G; ++g) acc[g] = 0u;
        for (int b = 0; b < mb; ++b) {
            for (int t = 0; t < 8; ++t) {
                const int jj = (t + rot) & 7, j = b * 8 + jj;
                const uint32_t c = j < m ? (uint32_t)__ldg(prow + j) : 0u;   // padded sub-spaces hold zeros
                lut_add<G>(acc, lutS, (uint32_t)(b * 2048) + c * 8u + (uint32_t)jj);
            }
        }
    }
}

// MODE 0: running top-R per (query, slice) in shared-memory BlockSelects -> partial.
// MODE 1: collect every row with d_int <= tau[q] into the query's global buffer (large R: the running
//         selection would dominate; with the threshold known the scan runs at its small-R speed).
template <int G, int MB, int MODE>
__global__ void __launch_bounds__(kAdcThreads)
adc_scan_kernel(const AdcArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int m = a.m, R = a.R, C = MODE == 0 ? a.C : 0, S = a.S;
    const int mb = (m + 7) >> 3;                                     // blocks of 8 sub-spaces (last one zero padded)
    u64* sel_buf = reinterpret_cast<u64*>(smem_raw);                 // G * C
    u64* sel_thr = sel_buf + (size_t)G * C;                          // G (padded to 2)
    uint32_t* lutS = reinterpret_cast<uint32_t*>(sel_thr + ((G + 1) & ~1));   // mb*256*8 slots of G
    int* sel_cnt = reinterpret_cast<int*>(lutS + (size_t)mb * 2048 * G);
    const int tid = threadIdx.x;
    const int rot = tid & 7;
    const int Qe = a.qsel ? min(a.Q, __ldg(a.qcount)) : a.Q;
    const int groups = a.qsel ? (Qe + G - 1) / G : a.groups;
    const int units = S * groups;

    for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int s = u / groups, grp = u - s * groups;
        const int q0 = grp * G;
        const int nq = min(G, Qe - q0);
        int qid[G];
#pragma unroll
        for (int g = 0; g < G; ++g) qid[g] = g < nq ? (a.qsel ? __ldg(a.qsel + q0 + g) : q0 + g) : 0;
        __syncthreads();   // previous unit fully drained
        // stage the group's LUTs into the rotated-slot layout
        // in SLOT order (i = ((j/8)*256 + c)*8 + j%8): consecutive lanes write consecutive G-word slots, one vector store
        // each (the (j, c)-ordered loop of r01 put a whole warp on one bank: 32 wavefronts per store, and its dependent
        // global loads were not overlapped -- 200 us per unit, most of the sample pass at 1M rows)
#pragma unroll 4
        for (int i = tid; i < mb * 2048; i += kAdcThreads) {
            const int j = ((i >> 11) << 3) | (i & 7), c = (i >> 3) & 255;
            typename LutVec<G>::type v;
            uint32_t* pv = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
            for (int g = 0; g < G; ++g) pv[g] = (g < nq && j < m) ? __ldg(a.lutq + ((size_t)qid[g] * m + j) * 256 + c) : 0u;
            *reinterpret_cast<typename LutVec<G>::type*>(lutS + (size_t)i * G) = v;
        }
        if (MODE == 0) {
            for (int i = tid; i < G * C; i += kAdcThreads) sel_buf[i] = VDB_KEY_MAX;
            if (tid < G) { sel_cnt[tid] = 0; sel_thr[tid] = VDB_KEY_MAX; }
        }
        uint32_t tau[G];
#pragma unroll
        for (int g = 0; g < G; ++g) tau[g] = (MODE == 1 && g < nq) ? __ldg(a.tau + qid[g]) : 0u;
        __syncthreads();

        const int64_t r_begin = (int64_t)s * a.slice_rows;
        const int64_t r_end = min(a.N, r_begin + a.slice_rows);
        // the next batch's code bytes are requested before this batch's lookups: MODE 0 has a barrier per batch, so the
        // load latency (a DRAM sector per row on the strided sample pass) is otherwise exposed once per batch
        // (MB >= 3 only: the small tables of m = 16 leave three CTAs per SM, whose warps hide the latency themselves, and
        // the extra registers cost more than the prefetch gains there)
        constexpr bool kPre = MB >= 3;
        uint32_t wn[MB > 0 ? 2 * MB : 2];
        if (kPre && r_begin + tid < r_end) adc_load_codes<MB>(a.codes + (r_begin + tid) * a.row_stride * m, wn);
        for (int64_t rb = r_begin; rb < r_end; rb += kAdcThreads) {
            const int64_t r = rb + tid;
            int nearly_full = 0;
            bool live = r < r_end;
            if (live && a.mask) { const int64_t pr = r * a.row_stride; live = (a.mask[pr >> 5] >> (pr & 31)) & 1u; }
            uint32_t w[MB > 0 ? 2 * MB : 2];
            if (kPre) {
#pragma unroll
                for (int i = 0; i < 2 * MB; ++i) w[i] = wn[i];
                if (r + kAdcThreads < r_end) adc_load_codes<MB>(a.codes + (r + kAdcThreads) * a.row_stride * m, wn);
            }
            if (live) {
                uint32_t acc[G];
                if (kPre) adc_sum_codes<G, MB>(w, lutS, rot, acc);
                else adc_row<G, MB>(a.codes + r * a.row_stride * m, m, mb, lutS, rot, acc);
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    if (g < nq) {
                        const u64 key = ((u64)acc[g] << 32) | (u64)(uint32_t)r;
                        if (MODE == 0) {
                            if (key < *(volatile u64*)(sel_thr + g)) {
                                const int p = atomicAdd(sel_cnt + g, 1);
                                sel_buf[(size_t)g * C + R + p] = key;
                                nearly_full |= (p + 1 + kAdcThreads > C - R);
                            }
                        } else if (acc[g] <= tau[g]) {
                            const int p = atomicAdd(a.ccnt + qid[g], 1);
                            if ((unsigned)p < (unsigned)a.ccap) a.cbuf[(size_t)qid[g] * a.ccap + p] = key;   // a wrapped counter (> 2^31 hits) must not index
                        }
                    }
                }
            }
            if (MODE == 0) {
                // compaction check, one barrier: a pusher that saw its list come within one batch of full
                // raises the flag (every push of this batch is done at the barrier, so the vote is CTA-uniform)
                if (__syncthreads_or(nearly_full)) {      // pushes of the G queries are correlated: sort all lists in one pass
                    block_bitonic_sort_lists(sel_buf, C, G);
                    if (tid < G) { sel_cnt[tid] = 0; sel_thr[tid] = sel_buf[(size_t)tid * C + R - 1]; }
                    __syncthreads();
                }
            }
        }
        if (MODE == 0) {
            __syncthreads();
            block_bitonic_sort_lists(sel_buf, C, G);
            for (int i = tid; i < nq * R; i += kAdcThreads) {
                const int g = i / R, j = i - g * R;
                a.partial[((size_t)(q0 + g) * S + s) * R + j] = sel_buf[(size_t)g * C + j];   // query SLOT (position in qsel)
            }
        }
    }
}

struct AdcFinArgs {
    const u64* keys;            // [slot][L]
    int L, R, C;
    const int32_t* lens;        // optional per-query key count (collect buffers): count < R or > L flags the query
    const int32_t* lens_valid;  // optional: how many of those keys are not VDB_KEY_MAX (coarse collect); < R flags the query
    const int32_t* expo;
    const int64_t* ids;
    float* out_d;
    int64_t* out_i;
    uint32_t* tau_out;          // threshold mode: write d_int of the R-th key instead of the result rows
    const int32_t* qsel;        // optional indirection (slot b -> query qsel[b], b < *qcount)
    const int32_t* qcount;
    int32_t* bad;               // flagged queries are appended here
    int32_t* bad_count;
};

// R best of L fixed-point keys per query -> (float(d_int) * 2^-e, id).  One CTA per query slot.
__global__ void __launch_bounds__(256)
adc_finalize_kernel(const AdcFinArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* buf = reinterpret_cast<u64*>(smem_raw);
    __shared__ int cnt;
    __shared__ u64 thr;
    int q = blockIdx.x;
    if (a.qsel) {
        if ((int)blockIdx.x >= __ldg(a.qcount)) return;
        q = __ldg(a.qsel + blockIdx.x);
    }
    const int R = a.R;
    int L = a.L;
    if (a.lens) {
        const int n = a.lens[q];
        if (n < R || n > L || (a.lens_valid && a.lens_valid[q] < R)) {   // too few rows under the threshold, or the buffer overflowed: exact rescan
            if (threadIdx.x == 0) a.bad[atomicAdd(a.bad_count, 1)] = q;
            return;
        }
        L = n;
    }
    BlockSelect sel;
    sel.init(buf, &cnt, &thr, a.C, R, blockDim.x);
    __syncthreads();
    const u64* src = a.keys + (size_t)(a.qsel ? (int)blockIdx.x : q) * a.L;
    for (int base = 0; base < L; base += blockDim.x) {
        int i = base + threadIdx.x;
        if (i < L) {
            u64 key = src[i];
            if (key != VDB_KEY_MAX) sel.push(key);
        }
        sel.maybe_compact();
    }
    sel.finalize();
    if (a.tau_out) {
        if (threadIdx.x == 0) {
            const u64 key = buf[R - 1];
            a.tau_out[q] = key == VDB_KEY_MAX ? 0xFFFFFFFFu : (uint32_t)(key >> 32);
        }
        return;
    }
    const int e = a.expo[q];
    for (int i = threadIdx.x; i < R; i += blockDim.x) {
        const u64 key = buf[i];
        const size_t o = (size_t)q * R + i;
        if (key == VDB_KEY_MAX) {
            a.out_d[o] = INFINITY;
            a.out_i[o] = -1;
        } else {
            a.out_d[o] = ldexpf((float)(uint32_t)(key >> 32), -e);
            const int64_t idx = (int64_t)(uint32_t)(key & 0xFFFFFFFFull);
            a.out_i[o] = a.ids ? a.ids[idx] : idx;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Coarse collect (r02).  The collect scan above is bound by the shared-memory pipe: one 128-bit lookup per code
// byte serves FOUR queries.  The threshold test does not need 32-bit entries: with
//     lut16[j][c] = min(lutq[j][c] - min_j, budget) >> s,   budget = tau - sum_j min_j,
// the packed sum is a LOWER bound of the row's distance in units of 2^s above sum_j min_j, so
//     d_int <= tau   ==>   sum_j lut16[j][code_j] <= budget >> s,
// and a 128-bit lookup now serves EIGHT queries (two 16-bit fields per word, plain 32-bit adds: s is the
// smallest shift that keeps every possible sum below 2^15, so no carry crosses a field and bit 15 is free
// for the packed compare (0x8000 + thr) - sum).  Rows that pass are a superset of the rows the exact test
// collects; `adc_exactify_kernel` then evaluates d_int of every collected row from the 32-bit tables (a few
// hundred rows per query) and drops the ones above tau, after which the result is what the exact collect
// would have produced -- same keys, same final selection, bit for bit.
template <int MB, int W>
__global__ void __launch_bounds__(kAdcThreads)
adc_coarse_collect_kernel(const AdcArgs a) {
    constexpr int G = 2 * W;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* lutS = reinterpret_cast<uint32_t*>(smem_raw);          // mb*256*8 slots of W words
    const int m = a.m, S = a.S;
    const int mb = (m + 7) >> 3;
    const int tid = threadIdx.x;
    const int rot = tid & 7;
    const int groups = a.groups, units = S * groups;

    for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int s = u / groups, grp = u - s * groups;
        const int q0 = grp * G;
        const int nq = min(G, a.Q - q0);
        __syncthreads();   // previous unit fully drained
        {
            // adc_coarse_lut_kernel wrote the group's table in the shared-memory layout already (rotated slots, the 2W
            // queries' 16-bit entries side by side, zeros for padded sub-spaces and absent queries): staging is a copy
            const uint4* src = reinterpret_cast<const uint4*>(a.lut16 + (size_t)grp * mb * 2048 * G);
            uint4* dst = reinterpret_cast<uint4*>(lutS);
            const int n16 = mb * 2048 * W / 4;
#pragma unroll 8
            for (int i = tid; i < n16; i += kAdcThreads) dst[i] = __ldg(src + i);
        }
        uint32_t thr[W];
#pragma unroll
        for (int w = 0; w < W; ++w) {
            const uint32_t lo = 2 * w < nq ? __ldg(a.thr16 + q0 + 2 * w) : 0u;
            const uint32_t hi = 2 * w + 1 < nq ? __ldg(a.thr16 + q0 + 2 * w + 1) : 0u;
            thr[w] = lo | (hi << 16);
        }
        __syncthreads();

        const int64_t r_begin = (int64_t)s * a.slice_rows;
        const int64_t r_end = min(a.N, r_begin + a.slice_rows);
        constexpr bool kPre = MB >= 3;             // next batch's code bytes in flight during this batch's lookups
        uint32_t wn[MB > 0 ? 2 * MB : 2];
        if (kPre && r_begin + tid < r_end) adc_load_codes<MB>(a.codes + (r_begin + tid) * m, wn);
        for (int64_t rb = r_begin; rb < r_end; rb += kAdcThreads) {
            const int64_t r = rb + tid;
            bool live = r < r_end;
            if (live && a.mask) live = (a.mask[r >> 5] >> (r & 31)) & 1u;
            uint32_t cw[MB > 0 ? 2 * MB : 2];
            if (kPre) {
#pragma unroll
                for (int i = 0; i < 2 * MB; ++i) cw[i] = wn[i];
                if (r + kAdcThreads < r_end) adc_load_codes<MB>(a.codes + (r + kAdcThreads) * m, wn);
            }
            if (live) {
                uint32_t acc[W];
                if (kPre) adc_sum_codes<W, MB>(cw, lutS, rot, acc);
                else adc_row<W, MB>(a.codes + r * m, m, mb, lutS, rot, acc);
                uint32_t hit[W], any = 0u;
#pragma unroll
                for (int w = 0; w < W; ++w) { hit[w] = (thr[w] - acc[w]) & 0x80008000u; any |= hit[w]; }
                if (any) {
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        if (hit[g >> 1] & (g & 1 ? 0x80000000u : 0x8000u)) {
                            const int p = atomicAdd(a.ccnt + q0 + g, 1);
                            if ((unsigned)p < (unsigned)a.ccap) a.cbuf[(size_t)(q0 + g) * a.ccap + p] = (u64)(uint32_t)r;
                        }
                    }
                }
            }
        }
    }
}

// 15-bit lower-bound tables and the packed thresholds (see adc_coarse_collect_kernel).  One CTA per group of G <= 8
// queries; the group's table leaves in the collect kernel's shared-memory order -- slot ((j/8)*256 + c)*8 + j%8, the
// G queries' 16-bit entries side by side, zeros for padded sub-spaces and absent queries -- so staging it is a copy.
__global__ void __launch_bounds__(256)
adc_coarse_lut_kernel(const uint32_t* __restrict__ lutq, int m, int Q, const uint32_t* __restrict__ tau,
                      uint16_t* __restrict__ lut16, uint32_t* __restrict__ thr16, int G) {
    __shared__ uint32_t mn[8][256], rg[8][256];
    __shared__ uint32_t s_budget[8];
    __shared__ int s_shift[8];
    const int grp = blockIdx.x, q0 = grp * G, nq = min(G, Q - q0);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int p = warp; p < nq * m; p += 8) {                 // (query, sub-space) pairs: min and range of the 256 entries
        const int e = p / m, j = p - e * m;
        const uint32_t* src = lutq + ((size_t)(q0 + e) * m + j) * 256;
        uint32_t lo = 0xFFFFFFFFu, hi = 0u;
        for (int c = lane; c < 256; c += 32) { const uint32_t v = __ldg(src + c); lo = min(lo, v); hi = max(hi, v); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if (lane == 0) { mn[e][j] = lo; rg[e][j] = hi - lo; }
    }
    __syncthreads();
    if (warp < nq) {                                         // warp e: budget and shift of query e
        const int e = warp;
        uint32_t off = 0u;                                   // m <= 255 entries <= 2^23: fits
        for (int j = lane; j < m; j += 32) off += mn[e][j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) off += __shfl_xor_sync(0xffffffffu, off, o);
        const uint32_t t = __ldg(tau + q0 + e);
        const uint32_t budget = t >= off ? t - off : 0u;     // t < off cannot happen for a tau taken from a row; 0 keeps a superset
        int sh = 0;
        for (; sh < 32; ++sh) {                              // smallest shift that keeps every possible sum below 2^15
            uint32_t tot = 0u;
            for (int j = lane; j < m; j += 32) tot += min(rg[e][j], budget) >> sh;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            if (tot <= 32767u) break;
        }
        if (lane == 0) {
            s_budget[e] = budget; s_shift[e] = sh;
            const uint32_t tc = sh < 32 ? budget >> sh : 0u;
            thr16[q0 + e] = 0x8000u + min(tc, 32767u);
        }
    }
    __syncthreads();
    const int mb = (m + 7) >> 3;
    uint16_t* dst = lut16 + (size_t)grp * mb * 2048 * G;
    for (int i = threadIdx.x; i < mb * 2048; i += 256) {     // slot order: consecutive threads write consecutive slots
        const int j = ((i >> 11) << 3) | (i & 7), c = (i >> 3) & 255;
        uint32_t w[4] = {0u, 0u, 0u, 0u};
        if (j < m) {
            for (int e = 0; e < nq; ++e) {
                const uint32_t v = min(__ldg(lutq + ((size_t)(q0 + e) * m + j) * 256 + c) - mn[e][j], s_budget[e]);
                const uint32_t x = s_shift[e] < 32 ? v >> s_shift[e] : 0u;
                w[e >> 1] |= x << (16 * (e & 1));
            }
        }
        uint32_t* d32 = reinterpret_cast<uint32_t*>(dst + (size_t)i * G);
        for (int k = 0; k < G / 2; ++k) d32[k] = w[k];
    }
}

// collected rows -> exact keys: d_int from the 32-bit tables; rows above tau become VDB_KEY_MAX.  One CTA per query.
__global__ void __launch_bounds__(256)
adc_exactify_kernel(const uint8_t* __restrict__ codes, int m, const uint32_t* __restrict__ lutq,
                    const uint32_t* __restrict__ tau, u64* __restrict__ cbuf, const int32_t* __restrict__ ccnt, int ccap,
                    int32_t* __restrict__ cvalid) {
    __shared__ int valid;
    const int q = blockIdx.x;
    const int n = ccnt[q];
    if (threadIdx.x == 0) valid = 0;
    __syncthreads();
    if (n <= ccap) {
        const uint32_t t = tau[q];
        const uint32_t* lq = lutq + (size_t)q * m * 256;
        u64* buf = cbuf + (size_t)q * ccap;
        int mine = 0;
        for (int p = threadIdx.x; p < n; p += 256) {
            const uint32_t r = (uint32_t)buf[p];
            const uint8_t* pr = codes + (size_t)r * m;
            uint32_t d = 0u;
#pragma unroll 8
            for (int j = 0; j < m; ++j) d += __ldg(lq + j * 256 + __ldg(pr + j));
            const bool ok = d <= t;
            buf[p] = ok ? (((u64)d << 32) | (u64)r) : VDB_KEY_MAX;
            mine += ok ? 1 : 0;
        }
        if (mine) atomicAdd(&valid, mine);
    }
    __syncthreads();
    if (threadIdx.x == 0) cvalid[q] = valid;
}

// ------------------------------------------------------------------------------------------------
// IVF-PQ: ADC restricted to the probed inverted lists.  Codes are stored LIST-MAJOR
// (code of CSR position p at codes[p*m]), so a list is one contiguous byte run.  grid (kIvfAdcSplits, Q):
// CTA (s, q) evaluates candidates [s*chunk, (s+1)*chunk) of query q's flattened candidate sequence (probe
// rank major, list position minor, as ivf.cu), with the query's fixed-point LUT in shared memory.  Keys are
// (d_int, sequence number).
// LUT layout: one query per CTA means 32-bit lookups, and with the rotated slot order of the full scan alone
// (bank = (c mod 4)*8 + j', 8 rotations for 32 lanes) the four lanes of a rotation class collide whenever their
// codes agree mod 4: ~3 wavefronts per lookup, and the kernel ran at the shared-memory pipe's limit.  P = 4
// keeps FOUR copies of every row, lane l reads copy (l >> 3) & 3: bank = copy*8 + j', conflict-free whatever
// the codes.  4 x 48 KB at m = 48 leaves one CTA per SM, so that variant runs 1024 threads; P = 1 / 256 threads
// is the fallback when the copies do not fit.
constexpr int kIvfAdcSplits = 32;

__device__ __forceinline__ int adc_find_probe(const int32_t* prefix, int n, int seq) {
    int lo = 0, hi = n;                 // invariant: prefix[lo] <= seq < prefix[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (prefix[mid] <= seq) lo = mid; else hi = mid;
    }
    return lo;
}

template <int MB, int P>
__global__ void __launch_bounds__(P == 4 ? 1024 : 256)
ivf_adc_scan_kernel(const uint8_t* __restrict__ codes, int m, const uint32_t* __restrict__ lutq,
                    const int32_t* __restrict__ list_ptr, const int32_t* __restrict__ probes,
                    const int32_t* __restrict__ prefix_g, int n_probe, int R, int C, int chunk_min,
                    u64* __restrict__ partial) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int qi = blockIdx.y, s = blockIdx.x, nt = blockDim.x;
    const int total = __ldg(prefix_g + (size_t)qi * (n_probe + 1) + n_probe);
    const int chunk = max(chunk_min, (total + kIvfAdcSplits - 1) / kIvfAdcSplits);
    if (s * chunk >= total) return;
    const int mb = (m + 7) >> 3;
    u64* buf = reinterpret_cast<u64*>(smem_raw);                                // C
    uint32_t* lutS = reinterpret_cast<uint32_t*>(buf + C);                      // mb * 2048 * P
    int32_t* prefix = reinterpret_cast<int32_t*>(lutS + (size_t)mb * 2048 * P); // n_probe + 1
    int32_t* pbase = prefix + n_probe + 1;                                      // n_probe
    __shared__ int cnt;
    __shared__ u64 thr;
    __shared__ int scratch[kSelectRadixScratch];
    const int tid = threadIdx.x, rot = tid & 7;
    // a warp stages 8 sub-spaces x 4 consecutive codes: 16-byte runs from global, 32 distinct banks per store
    for (int i = tid; i < mb * 2048; i += nt) {
        const int jj = i & 7, c = (i >> 3) & 255, b = i >> 11, j = b * 8 + jj;
        const uint32_t v = j < m ? __ldg(lutq + ((size_t)qi * m + j) * 256 + c) : 0u;
#pragma unroll
        for (int k = 0; k < P; ++k) {
            const int copy = (c + k) & (P - 1);
            lutS[(((b * 256 + c) * P + copy) << 3) + jj] = v;
        }
    }
    for (int p = tid; p <= n_probe; p += nt) prefix[p] = prefix_g[(size_t)qi * (n_probe + 1) + p];
    for (int p = tid; p < n_probe; p += nt) pbase[p] = list_ptr[probes[(size_t)qi * n_probe + p]];
    // a shortlist is large next to the CTA's candidates: select by radix, never sort (the keys leave unordered)
    BlockSelectRadix sel;
    sel.init(buf, &cnt, &thr, scratch, C, R, nt);
    __syncthreads();
    const uint32_t* lutL = lutS + ((tid >> 3) & (P - 1)) * 8;                   // this lane's copy
    const int begin = s * chunk, end = min(total, begin + chunk);
    // a thread's sequence numbers only grow, and lists are long next to the batch: the probe of the previous row is
    // almost always the probe of the next one (one compare instead of a binary search per row)
    int pcur = adc_find_probe(prefix, n_probe, min(begin + tid, total - 1));
    auto row_of = [&](int seq) -> const uint8_t* {
        while (seq >= prefix[pcur + 1]) ++pcur;          // seq < total = prefix[n_probe]
        return codes + ((int64_t)pbase[pcur] + (seq - prefix[pcur])) * m;
    };
    // No barrier per batch: the warps drift apart (their load, lookup and push phases then overlap instead of
    // hitting the same pipe together) and meet only when a pusher raises the selection's flag or the candidates
    // run out; a thread that is done keeps attending the meetings until every thread is.
    // The codes of the NEXT batch are in flight while this one is summed (one CTA per SM at P = 4).
    uint32_t wn[MB > 0 ? 2 * MB : 2];
    if (MB > 0 && begin + tid < end) adc_load_codes<MB, false>(row_of(begin + tid), wn);
    int base = begin;
    bool done = false;
    while (true) {
        while (!done && !sel.flag_raised()) {
            const int seq = base + tid;
            uint32_t acc[1];
            if (MB > 0) {
                uint32_t w[MB > 0 ? 2 * MB : 2];
#pragma unroll
                for (int i = 0; i < (MB > 0 ? 2 * MB : 2); ++i) w[i] = wn[i];
                if (seq + nt < end) adc_load_codes<MB, false>(row_of(seq + nt), wn);
                if (seq < end) adc_sum_codes<1, MB, P>(w, lutL, rot, acc);
            } else if (seq < end) {
                adc_row<1, MB>(row_of(seq), m, mb, lutS, rot, acc);
            }
            if (seq < end) sel.push_flagging(((u64)acc[0] << 32) | (u64)(uint32_t)seq);
            base += nt;
            done = base >= end;
        }
        __syncthreads();
        sel.compact_after_barrier();
        if (__syncthreads_and(done)) break;
    }
    for (int i = sel.kept + tid; i < R; i += nt) buf[i] = VDB_KEY_MAX;
    __syncthreads();
    u64* dst = partial + ((size_t)qi * kIvfAdcSplits + s) * R;
    for (int i = tid; i < R; i += nt) dst[i] = buf[i];
}

// R best (d_int, sequence) keys of the splits -> (float(d_int) * 2^-e, row).  One CTA per query.
__global__ void __launch_bounds__(256)
ivf_adc_merge_kernel(const u64* __restrict__ partial, int L, int R, int C, const int32_t* __restrict__ expo,
                     const int32_t* __restrict__ list_ptr, const int32_t* __restrict__ list_rows,
                     const int32_t* __restrict__ probes, const int32_t* __restrict__ prefix_g, int n_probe, int chunk_min,
                     float* __restrict__ out_d, int32_t* __restrict__ out_rows) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* buf = reinterpret_cast<u64*>(smem_raw);
    __shared__ int cnt;
    __shared__ u64 thr;
    __shared__ int scratch[kSelectRadixScratch];
    const int qi = blockIdx.x;
    const int32_t* prefix = prefix_g + (size_t)qi * (n_probe + 1);
    // the splits that ran (the same chunk rule as the scan); the others never wrote their slots
    const int total = __ldg(prefix + n_probe);
    const int chunk = max(chunk_min, (total + kIvfAdcSplits - 1) / kIvfAdcSplits);
    const int live = min(L, ((total + chunk - 1) / chunk) * R);
    BlockSelectRadix sel;
    sel.init(buf, &cnt, &thr, scratch, C, R, blockDim.x);
    __syncthreads();
    const u64* src = partial + (size_t)qi * L;
    for (int base = 0; base < live; base += blockDim.x) {
        const int i = base + threadIdx.x;
        if (i < live) {
            const u64 key = src[i];
            if (key != VDB_KEY_MAX) sel.push(key);
        }
        sel.maybe_compact();
    }
    const int P = next_pow2(R);                       // <= C
    sel.finalize(P);
    block_bitonic_sort(buf, P);                       // the result rows leave in ascending ADC distance
    const int e = expo[qi];
    for (int i = threadIdx.x; i < R; i += blockDim.x) {
        const u64 key = buf[i];
        const size_t o = (size_t)qi * R + i;
        if (key == VDB_KEY_MAX) {
            out_d[o] = INFINITY;
            out_rows[o] = -1;
        } else {
            const int seq = (int)(uint32_t)(key & 0xFFFFFFFFull);
            const int p = adc_find_probe(prefix, n_probe, seq);
            out_d[o] = ldexpf((float)(uint32_t)(key >> 32), -e);
            out_rows[o] = list_rows[list_ptr[probes[(size_t)qi * n_probe + p]] + (seq - prefix[p])];
        }
    }
}

struct AdcPlan {
    int G, C, S, groups, grid;
    int64_t slice_rows;
    size_t smem;
};

// Slices and grid for a (slice, query group) unit loop.  ~16 MB of codes per slice (L2-resident while the
// CTAs of one slice stream it), enough units to fill the machine, not more than max_S; among the admissible
// slice counts near that, the one whose unit count fills the last wave of CTAs best (768 units on 148 CTAs
// idle 14% of the machine in the sixth wave; 1036 do not).
static void plan_adc_slices(AdcPlan* p, int64_t N, int m, int64_t max_S, size_t smem_max) {
    int per_sm = (int)(smem_max / p->smem);
    if (per_sm > 4) per_sm = 4;
    if (per_sm < 1) per_sm = 1;
    const int64_t cap = (int64_t)sm_count() * per_sm;
    int64_t S = (N * (int64_t)m + (16ll << 20) - 1) / (16ll << 20);
    const int64_t fill = (2ll * sm_count() + p->groups - 1) / p->groups;
    if (S < fill) S = fill;
    int64_t max_by_rows = (N + 4095) / 4096;
    if (max_S > max_by_rows) max_S = max_by_rows;
    if (max_S < 1) max_S = 1;
    if (S > max_S) S = max_S;
    if (S < 1) S = 1;
    int64_t best = S;
    double best_eff = 0.0;
    for (int64_t c = S; c <= max_S && c <= 2 * S + 4; ++c) {
        const int64_t units = c * p->groups, grid = units < cap ? units : cap;
        const double eff = (double)units / (double)(((units + grid - 1) / grid) * grid);
        if (eff > best_eff + 0.01) { best_eff = eff; best = c; }
    }
    S = best;
    p->slice_rows = (N + S - 1) / S;
    if (p->slice_rows < 1) p->slice_rows = 1;
    p->S = (int)((N + p->slice_rows - 1) / p->slice_rows);
    if (p->S < 1) p->S = 1;
    const int64_t units = (int64_t)p->S * p->groups;
    p->grid = (int)(units < cap ? units : cap);
}

static bool plan_adc(int Q, int64_t N, int m, int R, AdcPlan* p) {
    const size_t smem_max = (size_t)max_smem_optin() - 64;
    // pending area: one batch of pushes (kAdcThreads) that must always fit, plus room for ~R more so that a
    // block sort is amortised over many pushes; 1024 up to R = 512 keeps 4 queries per CTA next to an m = 48 LUT
    // (R in (32, 64] took 512 until r02: the threshold sample with Rs = 64 then sorted after almost every batch)
    int C = R <= 32 ? 512 : (R <= 512 ? 1024 : next_pow2(R + 2 * kAdcThreads));
    const int mb = (m + 7) / 8;
    const int gs[3] = {4, 2, 1};
    for (int t = 0; t < 3; ++t) {
        const int G = gs[t];
        if (G > 1 && Q < G && t < 2 && Q <= gs[t + 1]) continue;   // do not waste LUT space on absent queries
        size_t smem = (size_t)G * C * 8 + (size_t)((G + 1) & ~1) * 8 + (size_t)mb * 2048 * G * 4 + (size_t)G * 4 + 16;
        if (smem > smem_max) continue;
        p->G = G; p->C = C; p->smem = smem;
        p->groups = (Q + G - 1) / G;
        int64_t max_by_keys = (1ll << 18) / R;          // S*R partial keys per query stay reasonable
        plan_adc_slices(p, N, m, max_by_keys < 1 ? 1 : max_by_keys, smem_max);
        return true;
    }
    return false;
}

template <int G, int MODE>
static int launch_adc(const AdcPlan& p, const AdcArgs& a, cudaStream_t st) {
#define VDB_ADC_CASE(MBV)                                                                                       \
    {                                                                                                           \
        auto kern = adc_scan_kernel<G, MBV, MODE>;                                                              \
        VDB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));         \
        kern<<<p.grid, kAdcThreads, p.smem, st>>>(a);                                                           \
    }
    // vector code loads need every visited row 16 (8) byte aligned: row pitch row_stride*m keeps the alignment of m
    const int m = a.m;
    const bool aligned16 = (reinterpret_cast<uintptr_t>(a.codes) & 15) == 0;
    const bool aligned8 = (reinterpret_cast<uintptr_t>(a.codes) & 7) == 0;
    if (aligned16 && m == 16) VDB_ADC_CASE(2)
    else if (aligned16 && m == 32) VDB_ADC_CASE(4)
    else if (aligned16 && m == 48) VDB_ADC_CASE(6)
    else if (aligned16 && m == 64) VDB_ADC_CASE(8)
    else if (aligned16 && m == 96) VDB_ADC_CASE(12)
    else if (aligned8 && m == 8) VDB_ADC_CASE(1)
    else if (aligned8 && m == 24) VDB_ADC_CASE(3)
    else VDB_ADC_CASE(0)
#undef VDB_ADC_CASE
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

template <int MODE>
static int launch_adc_g(const AdcPlan& p, const AdcArgs& a, cudaStream_t st) {
    if (p.G == 4) return launch_adc<4, MODE>(p, a, st);
    if (p.G == 2) return launch_adc<2, MODE>(p, a, st);
    return launch_adc<1, MODE>(p, a, st);
}

static int launch_adc_finalize(const AdcFinArgs& f, int slots, cudaStream_t st) {
    const size_t smem = (size_t)f.C * 8;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "R=%d too large for the final merge", f.R);
    VDB_CUDA(cudaFuncSetAttribute(adc_finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    adc_finalize_kernel<<<slots, 256, smem, st>>>(f);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

// Threshold-collect.  The running selection costs a barrier per batch of rows plus block sorts, and for
// large R its C >= R + 384 keys of shared memory per query push queries out of the CTA (the LUTs need the
// space), so it, not the code stream, sets the speed.  Instead: (A) the Rs best of a strided row sample give
// a per-query threshold tau with E[#rows <= tau] = T = Rs * stride; (B) one barrier-free scan collects every
// row with d_int <= tau; (C) the R best of the collected keys are the exact answer whenever
// R <= count <= capacity; (D, E) any other query (probability < 1e-5 on exchangeable rows -- count has
// relative spread 1/sqrt(Rs) and T is chosen >= 4.5 sigma above R; always on adversarial row orders) is
// rescanned by the running-selection kernel through a device-side list, so the result is exact in every
// case and the call never synchronises.
struct AdcCollectPlan {
    bool use;
    int64_t stride, n_sample;
    int cap, Rs;
    AdcPlan sample, collect, rescan;
    size_t off_tau, off_cnt, off_bad, off_badn, off_cbuf, off_partial, total;   // after lutq + expo
    bool coarse;                // 16-bit lower-bound collect (adc_coarse_collect_kernel) available for this shape
    AdcPlan cscan;              //   its unit plan; G = queries per group = 2 * W
    int W;
    size_t off_lut16, off_thr16, off_valid, lut16_bytes;
};

static int g_adc_coarse = 1;    // debug switch (vdb_debug_set_adc_coarse): 0 = always the 32-bit collect

static bool plan_adc_coarse_scan(int Q, int64_t N, int m, AdcPlan* p, int* W_out) {
    const int mb = (m + 7) / 8;
    const size_t smem_max = (size_t)max_smem_optin() - 64;
    const int ws[3] = {4, 2, 1};
    for (int t = 0; t < 3; ++t) {
        const int W = ws[t];
        const size_t smem = (size_t)mb * 2048 * W * 4;
        if (smem > smem_max) continue;
        p->G = 2 * W; p->C = 0; p->smem = smem;
        p->groups = (Q + p->G - 1) / p->G;
        plan_adc_slices(p, N, m, 1ll << 30, smem_max);
        *W_out = W;
        return true;
    }
    return false;
}

static bool plan_adc_collect_scan(int Q, int64_t N, int m, AdcPlan* p) {
    const int mb = (m + 7) / 8;
    const size_t smem_max = (size_t)max_smem_optin() - 64;
    const int gs[3] = {4, 2, 1};
    for (int t = 0; t < 3; ++t) {
        const int G = gs[t];
        const size_t smem = (size_t)((G + 1) & ~1) * 8 + (size_t)mb * 2048 * G * 4 + (size_t)G * 4 + 16;
        if (smem > smem_max) continue;
        p->G = G; p->C = 0; p->smem = smem;
        p->groups = (Q + G - 1) / G;
        plan_adc_slices(p, N, m, 1ll << 30, smem_max);
        return true;
    }
    return false;
}

static size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

static bool plan_adc_collect(int Q, int64_t N, int m, int R, AdcCollectPlan* c) {
    c->use = false;
    double T;
    // relative spread of the collected count: sigma = 1/sqrt(Rs); T (1 - 4.5 sigma) >= R and cap >= T (1 + 6 sigma)
    // r02: the sample pass runs the 32-bit running selection over N / stride rows, 1/16 of the scan at T = 512; since the
    // collect pass went to 16-bit tables (half the time per row) T is 4x larger -- collected rows cost next to nothing
    // (one atomic + 48 table reads in adc_exactify_kernel each), sample rows cost a full lookup pass.  cap also
    // covers the 16-bit tables' looser test (+20..40% rows).
    // R <= 100: T is 20x R or more, so the sample may be small -- Rs = 16 (count / T ~ Gamma(16)/16: P(count > cap / 1.4
    // = 2.9 T) ~ 1e-8 per query, P(count < R) far below that) halves the sample pass
    if (R <= 100) { c->Rs = 16; T = 2048.0; c->cap = 8192; }                                  // sigma 25%
    else if (R <= 512) { c->Rs = 32; T = R < 512 ? 2048.0 : 4.0 * R; c->cap = 3 * (int)T; }  // sigma 18%: T >= 4R, cap / 1.3 = 2.3 T
    else if (R <= 2048) { c->Rs = 64; T = 3.0 * R; c->cap = 8 * R; }                          // sigma 12.5%: cap / 1.3 = 2.05 T (8 sigma)
    else { c->Rs = 128; T = 1.8 * R; c->cap = 5 * R; }       // large R: the final selection over the collected keys dominates, keep them few
    c->stride = (int64_t)(T / c->Rs);
    if (c->stride < 1) c->stride = 1;
    c->n_sample = N / c->stride;
    if (c->n_sample < 8 * c->Rs || (int64_t)(16.0 * T) > N) return true;   // too few rows for a useful threshold
    if (!plan_adc(Q, c->n_sample, m, c->Rs, &c->sample)) return true;
    if (!plan_adc_collect_scan(Q, N, m, &c->collect)) return true;
    if (!plan_adc(Q, N, m, R, &c->rescan)) return false;
    size_t o = 0;
    c->off_tau = o; o += al256((size_t)Q * 4);
    c->off_cnt = o; o += al256((size_t)Q * 4);
    c->off_bad = o; o += al256((size_t)Q * 4);
    c->off_badn = o; o += 256;
    c->off_cbuf = o; o += al256((size_t)Q * c->cap * 8);
    c->off_partial = o;
    const size_t ps = (size_t)Q * c->sample.S * c->Rs * 8, pr = (size_t)Q * c->rescan.S * R * 8;
    o += al256(ps > pr ? ps : pr);
    c->coarse = Q > 4 && plan_adc_coarse_scan(Q, N, m, &c->cscan, &c->W);
    c->lut16_bytes = c->coarse ? (size_t)c->cscan.groups * ((m + 7) / 8) * 2048 * c->cscan.G * 2 : 0;   // [group][slot][G] u16
    c->off_lut16 = o; o += al256(c->lut16_bytes);
    c->off_thr16 = o; o += al256((size_t)Q * 4);
    c->off_valid = o; o += al256((size_t)Q * 4);
    c->total = o;
    c->use = true;
    return true;
}

template <int W>
static int launch_adc_coarse(const AdcPlan& p, const AdcArgs& a, cudaStream_t st) {
#define VDB_ADCC_CASE(MBV)                                                                                      \
    {                                                                                                           \
        auto kern = adc_coarse_collect_kernel<MBV, W>;                                                          \
        VDB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));         \
        kern<<<p.grid, kAdcThreads, p.smem, st>>>(a);                                                           \
    }
    const int m = a.m;
    const bool aligned16 = (reinterpret_cast<uintptr_t>(a.codes) & 15) == 0;
    const bool aligned8 = (reinterpret_cast<uintptr_t>(a.codes) & 7) == 0;
    if (aligned16 && m == 16) VDB_ADCC_CASE(2)
    else if (aligned16 && m == 32) VDB_ADCC_CASE(4)
    else if (aligned16 && m == 48) VDB_ADCC_CASE(6)
    else if (aligned16 && m == 64) VDB_ADCC_CASE(8)
    else if (aligned16 && m == 96) VDB_ADCC_CASE(12)
    else if (aligned8 && m == 8) VDB_ADCC_CASE(1)
    else if (aligned8 && m == 24) VDB_ADCC_CASE(3)
    else VDB_ADCC_CASE(0)
#undef VDB_ADCC_CASE
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

int vdb_pq_kmeans_step_coded(const float* x, int64_t N, int D, int m, int K, const uint8_t* codes, float* codebooks,
                             void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(N > 0 && D > 0 && m > 0 && K > 0 && K <= 256 && D % m == 0, VDB_ERR_INVALID,
                "vdb_pq_kmeans_step_coded: bad shape");
    const int dsub = D / m;
    VDB_REQUIRE(x && codes && codebooks, VDB_ERR_INVALID, "vdb_pq_kmeans_step_coded: null pointer");
    const size_t need = vdb_pq_kmeans_workspace(m, K, dsub);
    VDB_REQUIRE(ws && ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
    const size_t smem = ((size_t)K * dsub + K) * 4;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "K*dsub too large for shared memory");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* sums = static_cast<float*>(ws);
    float* counts = sums + (size_t)m * K * dsub;
    VDB_CUDA(cudaMemsetAsync(ws, 0, need, st));
    VDB_CUDA(cudaFuncSetAttribute(kmeans_accumulate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t rb = (N + 255) / 256;
    int gx = (int)((4ll * sm_count() + m - 1) / m);
    if (gx < 1) gx = 1;
    if (gx > rb) gx = (int)rb;
    kmeans_accumulate_kernel<<<dim3(gx, m), 256, smem, st>>>(x, N, D, m, K, dsub, codes, sums, counts);
    const int total = m * K * dsub;
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

void vdb_debug_set_adc_coarse(int on) { g_adc_coarse = on ? 1 : 0; }

// test hook: the collect pass's 15-bit tables for given 32-bit tables and thresholds (adc_coarse_lut_kernel as the scan
// launches it).  lut16 is [ceil(Q/G)][ceil8(m)/8 * 2048 slots][G] u16, slot = ((j/8)*256 + c)*8 + j%8; thr16 = 0x8000 + thr.
int vdb_debug_adc_coarse_tables(const uint32_t* lutq, int m, int Q, const uint32_t* tau, uint16_t* lut16, uint32_t* thr16,
                                int G, void* stream) {
    VDB_REQUIRE(lutq && tau && lut16 && thr16 && Q > 0 && m > 0 && m <= 255 && (G == 2 || G == 4 || G == 8), VDB_ERR_INVALID,
                "vdb_debug_adc_coarse_tables: bad arguments");
    adc_coarse_lut_kernel<<<(Q + G - 1) / G, 256, 0, static_cast<cudaStream_t>(stream)>>>(lutq, m, Q, tau, lut16, thr16, G);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

size_t vdb_pq_adc_topk_workspace(int Q, int64_t N, int m, int R) {
    if (Q <= 0 || R <= 0 || m <= 0) return 0;
    AdcCollectPlan c;
    if (!plan_adc_collect(Q, N, m, R, &c)) return 0;
    if (c.use) return adc_lutq_bytes(Q, m) + adc_expo_bytes(Q) + c.total + 256;
    AdcPlan p;
    if (!plan_adc(Q, N, m, R, &p)) return 0;
    return adc_lutq_bytes(Q, m) + adc_expo_bytes(Q) + (size_t)Q * p.S * R * 8 + 256;
}

int vdb_pq_adc_topk(const uint8_t* codes, int64_t N, int m, const float* lut, int Q, int R, const int64_t* ids,
                    const uint32_t* row_mask, float* out_d, int64_t* out_i, void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(Q >= 0 && N >= 0 && m > 0 && R > 0, VDB_ERR_INVALID, "vdb_pq_adc_topk: bad shape");
    VDB_REQUIRE(N < (1ll << 32), VDB_ERR_UNSUPPORTED, "N=%lld exceeds 2^32 rows per call", (long long)N);
    VDB_REQUIRE(m <= 255, VDB_ERR_UNSUPPORTED, "m=%d > 255 overflows the 32-bit fixed-point sum", m);
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(lut && out_d && out_i && (codes || N == 0), VDB_ERR_INVALID, "vdb_pq_adc_topk: null pointer");
    AdcCollectPlan c;
    AdcPlan p;
    VDB_REQUIRE(plan_adc_collect(Q, N, m, R, &c) && (c.use || plan_adc(Q, N, m, R, &p)), VDB_ERR_UNSUPPORTED,
                "m=%d R=%d do not fit shared memory", m, R);
    const size_t head = adc_lutq_bytes(Q, m) + adc_expo_bytes(Q);
    const size_t need = head + (c.use ? c.total : (size_t)Q * p.S * R * 8);
    VDB_REQUIRE(ws && ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
    VDB_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15) == 0, VDB_ERR_INVALID, "workspace must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* base = static_cast<char*>(ws);
    uint32_t* lutq = reinterpret_cast<uint32_t*>(base);
    int32_t* expo = reinterpret_cast<int32_t*>(base + adc_lutq_bytes(Q, m));
    adc_quantize_lut_kernel<<<Q, 256, 0, st>>>(lut, m * 256, lutq, expo);
    count_launch(1);

    AdcArgs a{};
    a.codes = codes; a.mask = row_mask; a.N = N; a.row_stride = 1; a.m = m; a.lutq = lutq; a.Q = Q; a.R = R;
    AdcFinArgs f{};
    f.R = R; f.expo = expo; f.ids = ids; f.out_d = out_d; f.out_i = out_i;
    int rc;
    if (!c.use) {
        u64* partial = reinterpret_cast<u64*>(base + head);
        a.C = p.C; a.slice_rows = p.slice_rows; a.S = p.S; a.groups = p.groups; a.partial = partial;
        {
            ProfScope prof(kProfAdcScan, st);
            rc = launch_adc_g<0>(p, a, st);
        }
        if (rc) return rc;
        f.keys = partial; f.L = p.S * R; f.C = select_capacity(R, 256);
        ProfScope prof(kProfFinalize, st);
        return launch_adc_finalize(f, Q, st);
    }

    char* w = base + head;
    uint32_t* tau = reinterpret_cast<uint32_t*>(w + c.off_tau);
    int32_t* ccnt = reinterpret_cast<int32_t*>(w + c.off_cnt);
    int32_t* bad = reinterpret_cast<int32_t*>(w + c.off_bad);
    int32_t* badn = reinterpret_cast<int32_t*>(w + c.off_badn);
    u64* cbuf = reinterpret_cast<u64*>(w + c.off_cbuf);
    u64* partial = reinterpret_cast<u64*>(w + c.off_partial);
    int32_t* cvalid = nullptr;                                             // set by the coarse collect
    VDB_CUDA(cudaMemsetAsync(ccnt, 0, c.off_cbuf - c.off_cnt, st));       // counters, flagged list and its length
    {
        ProfScope prof(kProfAdcScan, st);
        // (A) threshold from a strided sample
        AdcArgs s = a;
        s.N = c.n_sample; s.row_stride = c.stride; s.R = c.Rs; s.C = c.sample.C;
        s.slice_rows = c.sample.slice_rows; s.S = c.sample.S; s.groups = c.sample.groups; s.partial = partial;
        if ((rc = launch_adc_g<0>(c.sample, s, st))) return rc;
        AdcFinArgs fs{};
        fs.keys = partial; fs.L = c.sample.S * c.Rs; fs.R = c.Rs; fs.C = select_capacity(c.Rs, 256);
        fs.tau_out = tau;
        if ((rc = launch_adc_finalize(fs, Q, st))) return rc;
        // (B) collect
        AdcArgs b = a;
        b.tau = tau; b.cbuf = cbuf; b.ccnt = ccnt; b.ccap = c.cap;
        if (c.coarse && g_adc_coarse) {
            uint16_t* lut16 = reinterpret_cast<uint16_t*>(w + c.off_lut16);
            uint32_t* thr16 = reinterpret_cast<uint32_t*>(w + c.off_thr16);
            cvalid = reinterpret_cast<int32_t*>(w + c.off_valid);
            adc_coarse_lut_kernel<<<c.cscan.groups, 256, 0, st>>>(lutq, m, Q, tau, lut16, thr16, c.cscan.G);
            b.C = 0; b.slice_rows = c.cscan.slice_rows; b.S = c.cscan.S; b.groups = c.cscan.groups;
            b.lut16 = lut16; b.thr16 = thr16;
            rc = c.W == 4 ? launch_adc_coarse<4>(c.cscan, b, st)
                          : (c.W == 2 ? launch_adc_coarse<2>(c.cscan, b, st) : launch_adc_coarse<1>(c.cscan, b, st));
            if (rc) return rc;
            adc_exactify_kernel<<<Q, 256, 0, st>>>(codes, m, lutq, tau, cbuf, ccnt, c.cap, cvalid);
            count_launch(2);
            VDB_CUDA(cudaGetLastError());
        } else {
            b.C = 0; b.slice_rows = c.collect.slice_rows; b.S = c.collect.S; b.groups = c.collect.groups;
            if ((rc = launch_adc_g<1>(c.collect, b, st))) return rc;
        }
    }
    ProfScope prof(kProfFinalize, st);
    // (C) exact top-R of the collected keys; queries outside [R, cap] go to the flagged list
    f.keys = cbuf; f.L = c.cap; f.lens = ccnt; f.lens_valid = cvalid; f.C = select_capacity(R, 256); f.bad = bad; f.bad_count = badn;
    if ((rc = launch_adc_finalize(f, Q, st))) return rc;
    // (D, E) flagged queries: running-selection scan + merge, driven by the device-side list
    a.C = c.rescan.C; a.slice_rows = c.rescan.slice_rows; a.S = c.rescan.S; a.groups = c.rescan.groups;
    a.partial = partial; a.qsel = bad; a.qcount = badn;
    if ((rc = launch_adc_g<0>(c.rescan, a, st))) return rc;
    f.keys = partial; f.L = c.rescan.S * R; f.lens = nullptr; f.lens_valid = nullptr; f.bad = nullptr; f.bad_count = nullptr;
    f.qsel = bad; f.qcount = badn;
    return launch_adc_finalize(f, Q, st);
}

size_t vdb_ivf_adc_workspace(int Q, int m, int R) {
    if (Q <= 0 || m <= 0 || R <= 0) return 0;
    return adc_lutq_bytes(Q, m) + adc_expo_bytes(Q) + (size_t)Q * kIvfAdcSplits * R * 8 + 256;
}

int vdb_ivf_adc_scan(const uint8_t* codes, int m, const float* lut, int Q, const int32_t* list_ptr,
                     const int32_t* list_rows, const int32_t* probes, const int32_t* prefix, int n_probe, int R,
                     float* out_d, int32_t* out_rows, void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(Q >= 0 && m > 0 && R > 0 && n_probe >= 1, VDB_ERR_INVALID, "vdb_ivf_adc_scan: bad shape");
    VDB_REQUIRE(m <= 255, VDB_ERR_UNSUPPORTED, "m=%d > 255 overflows the 32-bit fixed-point sum", m);
    VDB_REQUIRE(Q <= 65535 && n_probe <= 8192 && R <= 4096, VDB_ERR_UNSUPPORTED, "Q=%d n_probe=%d R=%d outside the envelope",
                Q, n_probe, R);
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(codes && lut && list_ptr && list_rows && probes && prefix && out_d && out_rows, VDB_ERR_INVALID,
                "vdb_ivf_adc_scan: null pointer");
    const size_t head = adc_lutq_bytes(Q, m) + adc_expo_bytes(Q);
    const size_t need = head + (size_t)Q * kIvfAdcSplits * R * 8;
    VDB_REQUIRE(ws && ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
    VDB_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15) == 0, VDB_ERR_INVALID, "workspace must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* base = static_cast<char*>(ws);
    uint32_t* lutq = reinterpret_cast<uint32_t*>(base);
    int32_t* expo = reinterpret_cast<int32_t*>(base + adc_lutq_bytes(Q, m));
    u64* partial = reinterpret_cast<u64*>(base + head);
    adc_quantize_lut_kernel<<<Q, 256, 0, st>>>(lut, m * 256, lutq, expo);
    const int mb = (m + 7) / 8;
    const bool a16 = (reinterpret_cast<uintptr_t>(codes) & 15) == 0, a8 = (reinterpret_cast<uintptr_t>(codes) & 7) == 0;
    const bool unrolled = (a16 && (m == 16 || m == 32 || m == 48 || m == 64)) || (a8 && m == 8);
    const size_t tail = (size_t)(2 * n_probe + 1) * 4, smem_cap = (size_t)max_smem_optin() - 3072;   // static shared: 2.3 KB
    // four LUT copies + 1024 threads when they fit next to the selection buffer (radix select: any C >= R + 2 bursts)
    const size_t smem4 = (size_t)(R + 2 * 1024) * 8 + (size_t)mb * 2048 * 4 * 4 + tail;
    const bool four = unrolled && smem4 <= smem_cap;
    const int threads = four ? 1024 : 256;
    const int C = four ? R + 2 * 1024 : select_capacity(R, 256);
    const size_t smem = four ? smem4 : (size_t)C * 8 + (size_t)mb * 2048 * 4 + tail;
    VDB_REQUIRE(smem <= smem_cap, VDB_ERR_UNSUPPORTED, "m=%d R=%d n_probe=%d do not fit shared memory", m, R, n_probe);
    // candidates per CTA: the LUT staging (48 KB read, up to 192 KB written) is paid once per CTA
    const int chunk_min = Q >= 1024 ? 65536 : (Q >= 128 ? 16384 : 2048);
#define VDB_IVFADC_LAUNCH(MBV, PV)                                                                              \
    {                                                                                                           \
        auto kern = ivf_adc_scan_kernel<MBV, PV>;                                                               \
        VDB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));           \
        kern<<<dim3(kIvfAdcSplits, Q), threads, smem, st>>>(codes, m, lutq, list_ptr, probes, prefix, n_probe, R, C, \
                                                            chunk_min, partial);                                \
    }
#define VDB_IVFADC_CASE(MBV)                                                                                    \
    {                                                                                                           \
        if (four) VDB_IVFADC_LAUNCH(MBV, 4) else VDB_IVFADC_LAUNCH(MBV, 1)                                      \
    }
    {
        ProfScope prof(kProfAdcScan, st);
        if (a16 && m == 16) VDB_IVFADC_CASE(2)
        else if (a16 && m == 32) VDB_IVFADC_CASE(4)
        else if (a16 && m == 48) VDB_IVFADC_CASE(6)
        else if (a16 && m == 64) VDB_IVFADC_CASE(8)
        else if (a8 && m == 8) VDB_IVFADC_CASE(1)
        else VDB_IVFADC_LAUNCH(0, 1)
    }
#undef VDB_IVFADC_LAUNCH
#undef VDB_IVFADC_CASE
    const int C2 = select_capacity(R, 256);
    VDB_CUDA(cudaFuncSetAttribute(ivf_adc_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C2 * 8));
    ivf_adc_merge_kernel<<<Q, 256, (size_t)C2 * 8, st>>>(partial, kIvfAdcSplits * R, R, C2, expo, list_ptr, list_rows, probes,
                                                         prefix, n_probe, chunk_min, out_d, out_rows);
    count_launch(3);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"
