// Exact top-k on the 5th-gen tensor cores: bf16 coarse scan (tcgen05.mma fed by TMA, accumulators in
// TMEM) with a fused score + per-bin-minimum + per-row top-bins epilogue, followed by an exact fp32
// rerank of every row in the selected bins and a RIGOROUS certificate.  Queries whose certificate
// fails are reported so the caller can re-run them through the fp32 scan (exact_f32.cu).
//
// Replaces: the all-pairs distance + sort of tests/indexing/test_hnsw.py:128-134 /
// src/vector_db/indexing/ivf.py:176-199 (arithmetic: src/vector_db/util/distance.py:3-4).
//
// Why the answer is still exact.  The reference computes sqrt(sum((q-x)^2)) in fp32.  The coarse
// pass ranks by c(x) = |x|^2 - 2 q~.x~ (q~, x~ = bf16 roundings), whose error against the true
// t(x) = |x|^2 - 2 q.x is bounded by eps = 2 * (2^-8 * 1.01 + 2^-13) * |q| * max|x| (Cauchy-Schwarz
// on the two roundings + fp32 accumulation).  The kernel keeps, per query, the NB bins (8 consecutive
// rows) with the smallest bin minimum; every row outside them has c(x) >= tau (the largest kept bin
// minimum), hence t(x) >= tau - eps.  All 8*NB rows inside them are re-scored exactly.  If
// tau - eps > t_k (the k-th exact score) no outside row can enter the top-k: certified.
#include <cuda.h>

#include "common.cuh"
#include "sm100.cuh"

namespace vdb {

using namespace sm100;

constexpr int kTcThreads = 256;
constexpr int kTileM = 128;              // queries per CTA tile
constexpr int kTileN = 256;              // database rows per accumulator tile
constexpr int kBlockK = 64;              // bf16 elements per 128-byte swizzled row
constexpr int kUmmaK = 16;
constexpr int kABytes = kTileM * kBlockK * 2;    // 16 KB per A k-block
constexpr int kBBytes = kTileN * kBlockK * 2;    // 32 KB per B stage
constexpr int kBin = VDB_TC_BIN;
constexpr int kMaxStages = 8;

struct TcSmemLayout {
    int a_off, b_off, heap_off, coef_off, bar_off, total;
};

__host__ __device__ inline TcSmemLayout tc_layout(int num_kb, int stages, int NB) {
    TcSmemLayout L;
    L.a_off = 0;
    L.b_off = L.a_off + num_kb * kABytes;
    L.heap_off = L.b_off + stages * kBBytes;
    L.coef_off = L.heap_off + NB * kTileM * 8;
    L.bar_off = L.coef_off + kTileN * 4;
    L.total = L.bar_off + (2 * kMaxStages + 8) * 8 + 16;
    return L;
}

// max-heap (by key) of NB entries for one row, entries strided by kTileM; replace the root.
__device__ __forceinline__ void heap_replace_root(u64* h, int NB, u64 key) {
    int e = 0;
    while (true) {
        int c = 2 * e + 1;
        if (c >= NB) break;
        u64 kc = h[c * kTileM];
        if (c + 1 < NB) {
            u64 k2 = h[(c + 1) * kTileM];
            if (k2 > kc) { kc = k2; c = c + 1; }
        }
        if (kc <= key) break;
        h[e * kTileM] = kc;
        e = c;
    }
    h[e * kTileM] = key;
}

template <int METRIC>
__global__ void __launch_bounds__(kTcThreads, 1)
exact_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                const float* __restrict__ coef, int Q, int64_t N, int num_kb, int stages, int NB, int m_tiles, int S,
                int tiles_per_slice, int tiles_total, u64* __restrict__ partial) {
    extern __shared__ __align__(1024) unsigned char smem[];   // 128B-swizzled operand tiles need 1024 B alignment
    if ((smem_u32(smem) & 1023u) != 0) __trap();
    const TcSmemLayout L = tc_layout(num_kb, stages, NB);
    unsigned char* sA = smem + L.a_off;
    unsigned char* sB = smem + L.b_off;
    u64* heap = reinterpret_cast<u64*>(smem + L.heap_off);
    float* coefS = reinterpret_cast<float*>(smem + L.coef_off);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    uint64_t* full = bars;                         // [stages]
    uint64_t* empty = bars + kMaxStages;           // [stages]
    uint64_t* tmem_full = bars + 2 * kMaxStages;   // [2]
    uint64_t* tmem_empty = tmem_full + 2;          // [2]
    uint64_t* a_full = tmem_empty + 2;
    uint64_t* a_empty = a_full + 1;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(a_empty + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int units = m_tiles * S;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 128); }
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ================= TMA producer (one lane) =================
        if (lane == 0) {
            int st = 0, ph = 0, ui = 0;
            for (int u = blockIdx.x; u < units; u += gridDim.x, ++ui) {
                const int mt = u % m_tiles, s = u / m_tiles;
                const int t0 = s * tiles_per_slice;
                const int t1 = min(tiles_total, t0 + tiles_per_slice);
                mbar_wait(a_empty, (ui & 1) ^ 1);                 // previous unit's MMAs are done with A
                mbar_expect_tx(a_full, (uint32_t)(num_kb * kABytes));
                for (int kb = 0; kb < num_kb; ++kb)
                    tma_load_2d(sA + kb * kABytes, &map_a, a_full, kb * kBlockK, mt * kTileM, kEvictLast);
                for (int t = t0; t < t1; ++t) {
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
            int st = 0, ph = 0, as = 0, aph = 0, ui = 0;
            for (int u = blockIdx.x; u < units; u += gridDim.x, ++ui) {
                const int s = u / m_tiles;
                const int t0 = s * tiles_per_slice;
                const int t1 = min(tiles_total, t0 + tiles_per_slice);
                mbar_wait(a_full, ui & 1);
                tc_fence_after();
                for (int t = t0; t < t1; ++t) {
                    mbar_wait(&tmem_empty[as], aph ^ 1);          // epilogue drained this accumulator
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + (uint32_t)(as * kTileN);
                    for (int kb = 0; kb < num_kb; ++kb) {
                        mbar_wait(&full[st], ph);
                        tc_fence_after();
                        const uint64_t da = make_desc_k_sw128(smem_u32(sA + kb * kABytes));
                        const uint64_t db = make_desc_k_sw128(smem_u32(sB + st * kBBytes));
#pragma unroll
                        for (int k = 0; k < kBlockK / kUmmaK; ++k)
                            umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
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
        // ================= epilogue: TMEM -> scores -> bin minima -> per-row top-NB bins =================
        const int ew = warp & 3;                       // TMEM lane quarter this warp may touch
        const int row = ew * 32 + lane;                // accumulator row == query within the tile
        const int et = threadIdx.x - 128;              // 0..127
        u64* my_heap = heap + row;
        const u64 empty_key = pack_key(INFINITY, 0xFFFFFFFFu);
        int as = 0, aph = 0;
        for (int u = blockIdx.x; u < units; u += gridDim.x) {
            const int mt = u % m_tiles, s = u / m_tiles;
            const int t0 = s * tiles_per_slice;
            const int t1 = min(tiles_total, t0 + tiles_per_slice);
            for (int e = 0; e < NB; ++e) my_heap[e * kTileM] = empty_key;
            float tau = INFINITY;
            for (int t = t0; t < t1; ++t) {
                float* cs = coefS;
                {
                    const int64_t r0 = (int64_t)t * kTileN + et, r1 = r0 + 128;
                    const float oob = (METRIC == VDB_METRIC_L2) ? INFINITY : 0.f;
                    cs[et] = r0 < N ? __ldg(coef + r0) : oob;
                    cs[et + 128] = r1 < N ? __ldg(coef + r1) : oob;
                }
                named_bar_sync(1, 128);
                mbar_wait(&tmem_full[as], aph);
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * kTileN);
#pragma unroll 1
                for (int ch = 0; ch < kTileN / 32; ++ch) {
                    float v[32];
                    tmem_ld_x32(taddr + (uint32_t)(ch * 32), v);
                    tmem_ld_wait();
#pragma unroll
                    for (int b = 0; b < 32 / kBin; ++b) {
                        float m = INFINITY;
#pragma unroll
                        for (int e = 0; e < kBin; ++e) {
                            const float c = cs[ch * 32 + b * kBin + e];
                            const float sc = (METRIC == VDB_METRIC_L2) ? fmaf(-2.f, v[b * kBin + e], c) : v[b * kBin + e] * c;
                            m = fminf(m, sc);
                        }
                        if (m < tau) {
                            const uint32_t bin = (uint32_t)t * (kTileN / kBin) + (uint32_t)(ch * (32 / kBin) + b);
                            heap_replace_root(my_heap, NB, pack_key(m, bin));
                            tau = key_score(my_heap[0]);
                        }
                    }
                }
                tc_fence_before();
                mbar_arrive(&tmem_empty[as]);
                as ^= 1;
                if (as == 0) aph ^= 1;
                named_bar_sync(1, 128);                            // everyone is done with this tile's coefficients
            }
            const int q = mt * kTileM + row;
            if (q < Q) {
                u64* dst = partial + ((size_t)q * S + s) * NB;
                for (int e = 0; e < NB; ++e) dst[e] = my_heap[e * kTileM];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}


// ------------------------------------------------------------------------------------------
// helper kernels around the coarse pass

__global__ void f32_to_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = __float2bfloat16_rn(x[i]);
}

// cosine epilogue coefficient: -1/|x| (0 for zero rows, which can never be a meaningful neighbour)
__global__ void cosine_coef_kernel(const float* __restrict__ sqnorm, float* __restrict__ coef, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float s = sqnorm[i];
        coef[i] = s > 0.f ? -1.f / sqrtf(s) : 0.f;
    }
}

// Per query: merge the S per-slice bin lists, keep the NB best bins, expand them to candidate rows
// and record tau (see the certificate in the file header).  One CTA per query.
__global__ void __launch_bounds__(256)
select_bins_kernel(const u64* __restrict__ partial, int S, int NB, int C, int64_t N, int32_t* __restrict__ cand,
                   float* __restrict__ tau_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* buf = reinterpret_cast<u64*>(smem_raw);
    __shared__ int some_slice_full;
    const int q = blockIdx.x, total = S * NB;
    const u64 empty_key = pack_key(INFINITY, 0xFFFFFFFFu);
    if (threadIdx.x == 0) some_slice_full = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < C; i += blockDim.x) buf[i] = i < total ? partial[(size_t)q * total + i] : VDB_KEY_MAX;
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        bool full = true;
        for (int e = 0; e < NB; ++e) full &= partial[(size_t)q * total + (size_t)s * NB + e] != empty_key;
        if (full) some_slice_full = 1;
    }
    __syncthreads();
    block_bitonic_sort(buf, C);
    // empty_key sorts after every real key (score +inf, idx max) and before VDB_KEY_MAX padding
    __shared__ int valid_s;
    if (threadIdx.x == 0) {
        int valid = 0;
        while (valid < total && valid <= NB && buf[valid] < empty_key) ++valid;   // only need to know <= NB or more
        valid_s = valid;
        float tau;
        if (valid <= NB && !some_slice_full) tau = INFINITY;          // every bin of every slice is a candidate
        else tau = key_score(buf[NB - 1]);
        tau_out[q] = tau;
    }
    __syncthreads();
    const int valid = valid_s < NB ? valid_s : NB;
    for (int i = threadIdx.x; i < NB * kBin; i += blockDim.x) {
        const int j = i / kBin, e = i - j * kBin;
        int32_t r = -1;
        if (j < valid) {
            int64_t row = (int64_t)key_index(buf[j]) * kBin + e;
            if (row < N) r = (int32_t)row;
        }
        cand[(size_t)q * NB * kBin + i] = r;
    }
}

// One warp per query: decide whether the reranked top-k is provably the exact top-k.
template <int METRIC>
__global__ void certify_kernel(const float* __restrict__ q, int Q, int D, int k, const float* __restrict__ out_d,
                               const int64_t* __restrict__ out_i, const float* __restrict__ tau, float max_norm,
                               int32_t* __restrict__ uncert) {
    const int lane = threadIdx.x & 31;
    const int qi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (qi >= Q) return;
    float a = 0.f;
    for (int t = lane; t < D; t += 32) { float v = q[(size_t)qi * D + t]; a = fmaf(v, v, a); }
    a = warp_sum(a);
    if (lane != 0) return;
    const float qn = sqrtf(a);
    const float c = 0.00406742f;                                   // 2^-8 * 1.01 + 2^-13
    const float t_bound = tau[qi];
    bool ok;
    if (out_i[(size_t)qi * k + k - 1] < 0) {
        ok = (t_bound == INFINITY);                                // fewer than k rows exist: all must have been reranked
    } else if (METRIC == VDB_METRIC_L2) {
        const float dk = out_d[(size_t)qi * k + k - 1];
        const float tk = dk * dk - a;
        const float eps = 2.f * c * qn * max_norm;
        const float slack = 4e-6f * (qn + max_norm) * (qn + max_norm);
        ok = (t_bound - eps - slack) > tk;
    } else {
        const float simk = out_d[(size_t)qi * k + k - 1];
        const float tk = -simk * qn;
        const float eps = c * qn;
        const float slack = 4e-6f * qn;
        ok = (t_bound - eps - slack) > tk;
    }
    if (!ok) {
        int slot = atomicAdd(uncert, 1);
        uncert[1 + slot] = qi;
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

// bf16 row-major [rows][D] matrix, box = 64 (K) x box_rows, 128-byte swizzle, zero fill out of bounds
static int make_map(CUtensorMap* map, const void* base, int64_t rows, int D, int box_rows) {
    EncodeTiledFn fn = encode_tiled_fn();
    VDB_REQUIRE(fn != nullptr, VDB_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)D, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)D * 2};
    cuuint32_t box[2] = {(cuuint32_t)kBlockK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    VDB_REQUIRE(r == CUDA_SUCCESS, VDB_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return VDB_OK;
}

struct TcPlan {
    int num_kb, stages, NB, m_tiles, S, tiles_per_slice, tiles_total, grid;
    size_t smem;
};

static int tc_pick_nb(int k) { return k <= 16 ? 32 : 0; }

static bool plan_tc(int Q, int64_t N, int D, int k, TcPlan* p) {
    if (D % kBlockK != 0 || D <= 0 || Q <= 0 || N <= 0) return false;
    p->NB = tc_pick_nb(k);
    if (p->NB == 0) return false;
    p->num_kb = D / kBlockK;
    const int smem_max = max_smem_optin();
    int stages = kMaxStages;
    // + 1024: the 1024-byte aligned extern array makes the compiler reserve 1 KB of static shared memory
    while (stages >= 2 && tc_layout(p->num_kb, stages, p->NB).total + 1024 > smem_max) --stages;
    if (stages < 2) return false;
    p->stages = stages;
    p->smem = (size_t)tc_layout(p->num_kb, stages, p->NB).total;
    p->m_tiles = (Q + kTileM - 1) / kTileM;
    int64_t tiles_total = (N + kTileN - 1) / kTileN;
    if (tiles_total > (1 << 24)) return false;
    p->tiles_total = (int)tiles_total;
    const int sms = sm_count();
    // number of slices: minimise (waves x tiles per slice), i.e. the length of the critical path
    long long best_cost = -1;
    int best_S = 1;
    for (int S = 1; S <= 64; ++S) {
        int tps = (p->tiles_total + S - 1) / S;
        if (S > 1 && tps < 8) break;
        int S_eff = (p->tiles_total + tps - 1) / tps;
        long long units = (long long)p->m_tiles * S_eff;
        long long grid = units < sms ? units : sms;
        long long waves = (units + grid - 1) / grid;
        long long cost = waves * (tps + 2);           // +2: per-unit A load / output overhead in tile units
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_S = S; }
    }
    p->tiles_per_slice = (p->tiles_total + best_S - 1) / best_S;
    p->S = (p->tiles_total + p->tiles_per_slice - 1) / p->tiles_per_slice;
    long long units = (long long)p->m_tiles * p->S;
    p->grid = (int)(units < sms ? units : sms);
    return true;
}

struct TcWorkspace {
    size_t qbf16, coef, partial, cand, tau, total;
};

static TcWorkspace tc_workspace(const TcPlan& p, int Q, int64_t N, int D) {
    TcWorkspace w;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    w.qbf16 = take((size_t)Q * D * 2);
    w.coef = take((size_t)N * 4);
    w.partial = take((size_t)Q * p.S * p.NB * 8);
    w.cand = take((size_t)Q * p.NB * kBin * 4);
    w.tau = take((size_t)Q * 4);
    w.total = off;
    return w;
}

}  // namespace vdb

using namespace vdb;

extern "C" {

int vdb_exact_topk_tc_supported(int Q, int64_t N, int D, int k) {
    TcPlan p;
    return plan_tc(Q, N, D, k, &p) && N < (1ll << 31) ? 1 : 0;
}

size_t vdb_exact_topk_tc_workspace(int Q, int64_t N, int D, int k) {
    TcPlan p;
    if (!plan_tc(Q, N, D, k, &p)) return 0;
    return tc_workspace(p, Q, N, D).total;
}

int vdb_exact_topk_tc(const float* q, int Q, const float* db, const void* db_bf16, const float* db_sqnorm,
                      float db_max_norm, const int64_t* ids, int64_t N, int D, int k, int metric, float* out_d,
                      int64_t* out_i, int32_t* out_uncertified, void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(metric == VDB_METRIC_L2 || metric == VDB_METRIC_COSINE, VDB_ERR_INVALID, "unknown metric %d", metric);
    VDB_REQUIRE(Q > 0 && N > 0 && D > 0 && k > 0, VDB_ERR_INVALID, "vdb_exact_topk_tc: bad shape");
    VDB_REQUIRE(q && db && db_bf16 && db_sqnorm && out_d && out_i && out_uncertified && ws, VDB_ERR_INVALID,
                "vdb_exact_topk_tc: null pointer");
    VDB_REQUIRE(N < (1ll << 31), VDB_ERR_UNSUPPORTED, "N too large for int32 candidate rows");
    TcPlan p;
    VDB_REQUIRE(plan_tc(Q, N, D, k, &p), VDB_ERR_UNSUPPORTED, "tensor-core scan does not support D=%d k=%d", D, k);
    const TcWorkspace w = tc_workspace(p, Q, N, D);
    VDB_REQUIRE(ws_bytes >= w.total, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, w.total);
    VDB_REQUIRE((reinterpret_cast<uintptr_t>(db_bf16) & 15) == 0 && (reinterpret_cast<uintptr_t>(ws) & 255) == 0,
                VDB_ERR_INVALID, "bf16 mirror / workspace must be 16 / 256-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* base = static_cast<char*>(ws);
    __nv_bfloat16* qb = reinterpret_cast<__nv_bfloat16*>(base + w.qbf16);
    float* coef = reinterpret_cast<float*>(base + w.coef);
    u64* partial = reinterpret_cast<u64*>(base + w.partial);
    int32_t* cand = reinterpret_cast<int32_t*>(base + w.cand);
    float* tau = reinterpret_cast<float*>(base + w.tau);

    f32_to_bf16_kernel<<<sm_count() * 4, 256, 0, st>>>(q, qb, (int64_t)Q * D);
    const float* coef_ptr = db_sqnorm;
    if (metric == VDB_METRIC_COSINE) {
        cosine_coef_kernel<<<sm_count() * 4, 256, 0, st>>>(db_sqnorm, coef, N);
        coef_ptr = coef;
        count_launch(1);
    }
    CUtensorMap map_a, map_b;
    int rc = make_map(&map_a, qb, Q, D, kTileM);
    if (rc) return rc;
    rc = make_map(&map_b, db_bf16, N, D, kTileN);
    if (rc) return rc;
    {
        ProfScope prof(kProfExactTC, st);
        if (metric == VDB_METRIC_L2) {
            VDB_CUDA(cudaFuncSetAttribute(exact_tc_kernel<VDB_METRIC_L2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
            exact_tc_kernel<VDB_METRIC_L2><<<p.grid, kTcThreads, p.smem, st>>>(map_a, map_b, coef_ptr, Q, N, p.num_kb, p.stages, p.NB,
                                                                              p.m_tiles, p.S, p.tiles_per_slice, p.tiles_total, partial);
        } else {
            VDB_CUDA(cudaFuncSetAttribute(exact_tc_kernel<VDB_METRIC_COSINE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
            exact_tc_kernel<VDB_METRIC_COSINE><<<p.grid, kTcThreads, p.smem, st>>>(map_a, map_b, coef_ptr, Q, N, p.num_kb, p.stages, p.NB,
                                                                                  p.m_tiles, p.S, p.tiles_per_slice, p.tiles_total, partial);
        }
    }
    VDB_CUDA(cudaGetLastError());
    const int C = next_pow2(p.S * p.NB < 64 ? 64 : p.S * p.NB);
    VDB_CUDA(cudaFuncSetAttribute(select_bins_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C * 8));
    select_bins_kernel<<<Q, 256, (size_t)C * 8, st>>>(partial, p.S, p.NB, C, N, cand, tau);
    count_launch(3);
    VDB_CUDA(cudaGetLastError());
    rc = vdb_rerank_f32(q, Q, db, ids, N, D, cand, p.NB * kBin, k, metric, out_d, out_i, stream);
    if (rc) return rc;
    VDB_CUDA(cudaMemsetAsync(out_uncertified, 0, 4, st));
    const int warps_per_block = 8;
    const int blocks = (Q + warps_per_block - 1) / warps_per_block;
    if (metric == VDB_METRIC_L2)
        certify_kernel<VDB_METRIC_L2><<<blocks, warps_per_block * 32, 0, st>>>(q, Q, D, k, out_d, out_i, tau, db_max_norm, out_uncertified);
    else
        certify_kernel<VDB_METRIC_COSINE><<<blocks, warps_per_block * 32, 0, st>>>(q, Q, D, k, out_d, out_i, tau, db_max_norm, out_uncertified);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"
