// IVF (inverted file) search on B200 -- SURVEY.md 8(f) rank 3.
//
// Reference: src/vector_db/indexing/ivf.py.  The index is K centroids + K inverted lists of node ids;
// search (ivf.py:143-199) ranks the centroids by np.linalg.norm(centroids - query, axis=1), takes the
// n_probe nearest (np.argsort), computes euclidean_vector_distance to every member of those lists,
// sorts the candidates by distance with a STABLE sort (candidate enumeration order = probe rank, then
// list position, breaks ties) and returns the first top_k.
//
// Here: lists are CSR (list_ptr[K+1], list_rows[] = row indices into the HBM-resident fp32 matrix, in
// the reference's list order) and queries are processed in batches.
//   ivf_centroid_dist_kernel  distances to the centroids in numpy's exact operation order (fp32
//                             subtract, fp32 square, pairwise_sum over the row, sqrt) so that the probe
//                             set -- index work -- is bit-identical to the reference's
//   ivf_probe_kernel          n_probe smallest (distance, centroid) per query + candidate prefix sums
//   ivf_scan_kernel           one CTA per (query, split of the query's flattened candidate sequence):
//                             warp per row, 128-bit loads, the exact scan's L2 arithmetic (so that
//                             n_probe == K reproduces vdb_exact_topk_f32 bit for bit), BlockSelect on
//                             (distance, sequence number) keys = the reference's stable order
//   ivf_merge_kernel          k best of the splits, sequence number -> row -> external id
// HBM bound: algorithmic bytes per query = (sum of probed list lengths) * 4 D  (+ 4 per row index).
#include "common.cuh"

namespace vdb {

constexpr int kIvfThreads = 256;
constexpr int kIvfWarps = kIvfThreads / 32;
constexpr int kIvfRowsPerWarp = 4;                         // rows in flight per warp iteration
constexpr int kIvfCheckEvery = 4;                          // warp iterations between two compaction checks (one barrier)
constexpr int kIvfBatch = kIvfWarps * kIvfRowsPerWarp;     // candidates per CTA iteration
constexpr int kIvfBurst = kIvfBatch * kIvfCheckEvery;      // pushes between two compaction checks
constexpr int kIvfSplits = 32;                             // CTAs per query (those past the query's candidates exit)
constexpr int kIvfMaxProbe = 8192;                         // probe prefix table lives in shared memory

// numpy pairwise_sum (numpy/_core/src/umath/loops_utils.h.src) over v[t] = rn(d[t]^2), d[t] = rn(c[t] - q[t]):
// n < 8 plain loop; n <= 128 eight strided accumulators combined as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) then
// the tail; larger n split at n/2 rounded down to a multiple of 8, recursively.  The recursion depends on n
// only, so it is flattened once per CTA into its leaf blocks (start, length <= 128, all starts multiples of
// 8) in visiting order plus, per leaf, the number of pairwise additions that follow it in post-order.
constexpr int kIvfCdThreads = 128;      // centroids per CTA
constexpr int kIvfCdQ = 4;              // queries per thread: one centroid element feeds four distances
constexpr int kIvfCdChunk = 32;         // columns per shared-memory tile
constexpr int kIvfCdStride = kIvfCdChunk + 4;   // tile row stride: 16-byte aligned rows, 128-bit reads of 8 consecutive rows hit 32 banks
constexpr int kIvfCdMaxLeaves = 512;

struct IvfLeaves {
    int n;
    short start[kIvfCdMaxLeaves];       // in units of 8 columns
    short len[kIvfCdMaxLeaves];
    unsigned char comb[kIvfCdMaxLeaves];
};

__device__ void ivf_flatten_leaves(IvfLeaves* L, int n) {
    int off[32], len[32], state[32];
    int sp = 1;
    off[0] = 0; len[0] = n; state[0] = 0;
    L->n = 0;
    while (sp > 0) {
        const int t = sp - 1;
        if (len[t] <= 128) {
            L->start[L->n] = (short)(off[t] >> 3);
            L->len[L->n] = (short)len[t];
            L->comb[L->n] = 0;
            ++L->n;
            --sp;
        } else {
            int n2 = len[t] / 2;
            n2 -= n2 % 8;
            if (state[t] == 0) { state[t] = 1; off[sp] = off[t]; len[sp] = n2; state[sp] = 0; ++sp; }
            else if (state[t] == 1) { state[t] = 2; off[sp] = off[t] + n2; len[sp] = len[t] - n2; state[sp] = 0; ++sp; }
            else { ++L->comb[L->n - 1]; --sp; }
        }
    }
}

// out[q][c] = np.linalg.norm(centroids[c] - query[q])  (ivf.py:109, :170).  CTA = 128 centroids x kIvfCdQ queries;
// centroid columns go through a [128][36] shared tile (coalesced loads, conflict-free per-thread 128-bit row reads); the
// eight columns of a block are two 128-bit loads per operand row instead of eight scalar ones (the kernel was bound by its
// shared-memory load instructions: 40 per 96 flops).
__global__ void __launch_bounds__(kIvfCdThreads)
ivf_centroid_dist_kernel(const float* __restrict__ q, int Q, int D, const float* __restrict__ cent, int K,
                         float* __restrict__ out, const int32_t* __restrict__ flags) {
    extern __shared__ __align__(16) float ivf_smem[];
    float* qs = ivf_smem;                                              // kIvfCdQ * D
    float* tile = qs + (size_t)kIvfCdQ * D;                            // 128 * 33
    __shared__ IvfLeaves leaves;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c0 = blockIdx.x * kIvfCdThreads, q0 = blockIdx.y * kIvfCdQ;
    const int nq = min(kIvfCdQ, Q - q0);
    if (flags) {                                                       // only the queries a cheaper pass could not settle
        bool any = false;
        for (int g = 0; g < nq; ++g) any = any || flags[q0 + g] != 0;
        if (!any) return;
    }
    for (int i = tid; i < kIvfCdQ * D; i += kIvfCdThreads) {
        const int g = i / D;
        qs[i] = g < nq ? q[(size_t)(q0 + g) * D + (i - g * D)] : 0.f;
    }
    if (tid == 0) ivf_flatten_leaves(&leaves, D);
    int chunk_base = -kIvfCdChunk;                                     // no tile loaded yet
    auto need = [&](int t) {                                          // CTA-uniform: make column t resident
        if (t >= chunk_base + kIvfCdChunk) {
            __syncthreads();
            chunk_base = t & ~(kIvfCdChunk - 1);
            for (int r = warp; r < kIvfCdThreads; r += kIvfCdThreads / 32) {
                const int c = c0 + r, col = chunk_base + lane;
                tile[r * kIvfCdStride + lane] = (c < K && col < D) ? __ldg(cent + (size_t)c * D + col) : 0.f;
            }
            __syncthreads();
        }
    };
    __syncthreads();
    const float* mine = tile + tid * kIvfCdStride;
    const bool vec_ok = (D & 3) == 0;                                  // 128-bit shared loads of the query rows need 16-byte aligned rows
    float val[kIvfCdQ][12];                                            // post-order value stack (depth <= log2(D/64)+1)
    int nv = 0;
    for (int l = 0; l < leaves.n; ++l) {
        const int s = (int)leaves.start[l] << 3, n = leaves.len[l];
        float res[kIvfCdQ];
        if (n < 8) {
#pragma unroll
            for (int g = 0; g < kIvfCdQ; ++g) res[g] = 0.f;
            for (int t = 0; t < n; ++t) {
                need(s + t);
                const float cv = mine[s + t - chunk_base];
#pragma unroll
                for (int g = 0; g < kIvfCdQ; ++g) {
                    const float d = __fsub_rn(cv, qs[g * D + s + t]);
                    res[g] = __fadd_rn(res[g], __fmul_rn(d, d));
                }
            }
        } else {
            float r[kIvfCdQ][8];
            const int main_end = n - (n % 8);
            for (int i = 0; i < main_end; i += 8) {
                need(s + i);
                const float* cp = mine + (s + i - chunk_base);
                float cv8[8];
                {
                    const float4 a = *reinterpret_cast<const float4*>(cp), b = *reinterpret_cast<const float4*>(cp + 4);
                    cv8[0] = a.x; cv8[1] = a.y; cv8[2] = a.z; cv8[3] = a.w; cv8[4] = b.x; cv8[5] = b.y; cv8[6] = b.z; cv8[7] = b.w;
                }
#pragma unroll
                for (int g = 0; g < kIvfCdQ; ++g) {
                    float qv[8];
                    const float* qp = qs + g * D + s + i;
                    if (vec_ok) {
                        const float4 a = *reinterpret_cast<const float4*>(qp), b = *reinterpret_cast<const float4*>(qp + 4);
                        qv[0] = a.x; qv[1] = a.y; qv[2] = a.z; qv[3] = a.w; qv[4] = b.x; qv[5] = b.y; qv[6] = b.z; qv[7] = b.w;
                    } else {
#pragma unroll
                        for (int t = 0; t < 8; ++t) qv[t] = qp[t];
                    }
#pragma unroll
                    for (int t = 0; t < 8; ++t) {
                        const float d = __fsub_rn(cv8[t], qv[t]);
                        const float v = __fmul_rn(d, d);
                        r[g][t] = i == 0 ? v : __fadd_rn(r[g][t], v);
                    }
                }
            }
#pragma unroll
            for (int g = 0; g < kIvfCdQ; ++g)
                res[g] = __fadd_rn(__fadd_rn(__fadd_rn(r[g][0], r[g][1]), __fadd_rn(r[g][2], r[g][3])),
                                   __fadd_rn(__fadd_rn(r[g][4], r[g][5]), __fadd_rn(r[g][6], r[g][7])));
            for (int i = main_end; i < n; ++i) {
                need(s + i);
                const float cv = mine[s + i - chunk_base];
#pragma unroll
                for (int g = 0; g < kIvfCdQ; ++g) {
                    const float d = __fsub_rn(cv, qs[g * D + s + i]);
                    res[g] = __fadd_rn(res[g], __fmul_rn(d, d));
                }
            }
        }
#pragma unroll
        for (int g = 0; g < kIvfCdQ; ++g) val[g][nv] = res[g];
        ++nv;
        for (int k = 0; k < (int)leaves.comb[l]; ++k) {
#pragma unroll
            for (int g = 0; g < kIvfCdQ; ++g) val[g][nv - 2] = __fadd_rn(val[g][nv - 2], val[g][nv - 1]);
            --nv;
        }
    }
    const int c = c0 + tid;
    if (c < K) {
#pragma unroll
        for (int g = 0; g < kIvfCdQ; ++g)
            if (g < nq) out[(size_t)(q0 + g) * K + c] = __fsqrt_rn(val[g][0]);
    }
}

// np.argsort(centroid_distances)[:n_probe] (ivf.py:171), ties by centroid index; plus the prefix sums of the
// probed lists' lengths: prefix[q][p] = candidates enumerated before probe p, prefix[q][n_probe] = total.
__global__ void __launch_bounds__(256)
ivf_probe_kernel(const float* __restrict__ cdist, int K, int n_probe, int C, const int32_t* __restrict__ list_ptr,
                 int32_t* __restrict__ probes, int32_t* __restrict__ prefix) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* buf = reinterpret_cast<u64*>(smem_raw);
    __shared__ int cnt;
    __shared__ u64 thr;
    BlockSelect sel;
    sel.init(buf, &cnt, &thr, C, n_probe, blockDim.x);
    __syncthreads();
    const float* src = cdist + (size_t)blockIdx.x * K;
    for (int base = 0; base < K; base += blockDim.x) {
        const int c = base + threadIdx.x;
        if (c < K) sel.push(pack_key(src[c], (uint32_t)c));
        sel.maybe_compact();
    }
    sel.finalize();
    int32_t* pr = probes + (size_t)blockIdx.x * n_probe;
    for (int p = threadIdx.x; p < n_probe; p += blockDim.x) pr[p] = (int32_t)key_index(buf[p]);
    if (prefix && threadIdx.x == 0) {
        int32_t* px = prefix + (size_t)blockIdx.x * (n_probe + 1);
        int run = 0;
        for (int p = 0; p < n_probe; ++p) {
            px[p] = run;
            const int c = (int)key_index(buf[p]);
            run += list_ptr[c + 1] - list_ptr[c];
        }
        px[n_probe] = run;
    }
}

// The same selection for a BATCH with few probes (n_probe <= 32, K <= 4096): one WARP per query, the K keys in registers,
// n_probe rounds of "smallest remaining (distance, centroid) key" -- the CTA-per-query BlockSelect above costs 0.36 ms for
// 4096 queries x 1024 centroids, a quarter of a list-major batch search.
constexpr int kIvfProbeWarpMaxK = 4096;
template <int kPer>                                            // keys per lane: K <= 32 kPer
__global__ void __launch_bounds__(256)
ivf_probe_warp_kernel(const float* __restrict__ cdist, int Q, int K, int n_probe, const int32_t* __restrict__ list_ptr,
                      int32_t* __restrict__ probes, int32_t* __restrict__ prefix, const int32_t* __restrict__ flags) {
    const int q = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (q >= Q || (flags && flags[q] == 0)) return;
    uint32_t dk[kPer];                                         // ordered distance bits; the centroid index is t * 32 + lane
    const float* src = cdist + (size_t)q * K;
#pragma unroll
    for (int t = 0; t < kPer; ++t) {
        const int c = t * 32 + lane;
        dk[t] = (c < K && t * 32 < K) ? f32_to_ordered(src[c]) : 0xFFFFFFFFu;
    }
    int run = 0;
    for (int p = 0; p < n_probe; ++p) {
        u64 best = VDB_KEY_MAX;
#pragma unroll
        for (int t = 0; t < kPer; ++t) {
            const u64 k = ((u64)dk[t] << 32) | (uint32_t)(t * 32 + lane);
            best = (dk[t] != 0xFFFFFFFFu && k < best) ? k : best;
        }
        u64 wbest = best;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(0xffffffffu, wbest, o);
            wbest = other < wbest ? other : wbest;
        }
        const int c = (int)(uint32_t)(wbest & 0xFFFFFFFFull);
        if ((c & 31) == lane) {                                // the owner retires it
#pragma unroll
            for (int t = 0; t < kPer; ++t) dk[t] = (t == (c >> 5)) ? 0xFFFFFFFFu : dk[t];
        }
        if (lane == 0) {
            probes[(size_t)q * n_probe + p] = c;
            if (prefix) prefix[(size_t)q * (n_probe + 1) + p] = run;
        }
        if (prefix) run += __ldg(list_ptr + c + 1) - __ldg(list_ptr + c);
    }
    if (prefix && lane == 0) prefix[(size_t)q * (n_probe + 1) + n_probe] = run;
}

// Probe selection through the tensor cores (batches, n_probe + 4 <= 16): vdb_exact_topk_tc gives the kc nearest centroids of
// every query EXACTLY in the rerank's arithmetic r(.) (sequential fp32 direct differences); the probe set, however, must be
// numpy's -- np.argsort of np.linalg.norm(centroids - q, axis=1), pairwise summation (ivf.py:170-171).  One thread per
// (query, candidate) recomputes the candidate's distance n(.) in numpy's order; both are fp32 evaluations of the same
// number, |n(c) - r(c)| <= eps r(c) with eps = D 2^-23 (generous).  Every centroid outside the candidates has r >= r_max,
// hence n >= r_max (1 - eps): if the n_probe-th smallest n among the candidates is below that, numpy's n_probe nearest are
// all candidates and their order by (n, index) is numpy's.  Otherwise the query is flagged and the full kernels decide.
__device__ float ivf_np_dist_one(const float* __restrict__ c, const float* __restrict__ q, const IvfLeaves& L) {
    float val[12];
    int nv = 0;
    for (int l = 0; l < L.n; ++l) {
        const int s = (int)L.start[l] << 3, n = L.len[l];
        float res;
        if (n < 8) {
            res = 0.f;
            for (int t = 0; t < n; ++t) { const float d = __fsub_rn(c[s + t], q[s + t]); res = __fadd_rn(res, __fmul_rn(d, d)); }
        } else {
            float r[8];
            const int main_end = n - (n % 8);
            for (int i = 0; i < main_end; i += 8) {
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    const float d = __fsub_rn(c[s + i + t], q[s + i + t]);
                    const float v = __fmul_rn(d, d);
                    r[t] = i == 0 ? v : __fadd_rn(r[t], v);
                }
            }
            res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])), __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
            for (int i = main_end; i < n; ++i) { const float d = __fsub_rn(c[s + i], q[s + i]); res = __fadd_rn(res, __fmul_rn(d, d)); }
        }
        val[nv++] = res;
        for (int k = 0; k < (int)L.comb[l]; ++k) { val[nv - 2] = __fadd_rn(val[nv - 2], val[nv - 1]); --nv; }
    }
    return __fsqrt_rn(val[0]);
}

__global__ void __launch_bounds__(256)
ivf_probe_refine_kernel(const float* __restrict__ q, int Q, int D, const float* __restrict__ cent, int K,
                        const float* __restrict__ cand_d, const int64_t* __restrict__ cand_i, int kc, int n_probe,
                        const int32_t* __restrict__ list_ptr, int32_t* __restrict__ probes, int32_t* __restrict__ prefix,
                        int32_t* __restrict__ flags) {
    __shared__ IvfLeaves leaves;
    if (threadIdx.x == 0) ivf_flatten_leaves(&leaves, D);
    __syncthreads();
    const int qi = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (qi >= Q) return;
    int c = -1;
    float r = INFINITY, n = INFINITY;
    if (lane < kc) {
        const int64_t ci = cand_i[(size_t)qi * kc + lane];
        if (ci >= 0 && ci < K) {
            c = (int)ci;
            r = cand_d[(size_t)qi * kc + lane];
            n = ivf_np_dist_one(cent + (size_t)c * D, q + (size_t)qi * D, leaves);
        }
    }
    const u64 key = c >= 0 ? pack_key(n, (uint32_t)c) : VDB_KEY_MAX;
    int rank = 0;
    float r_max = 0.f;
    int n_valid = 0;
    for (int j = 0; j < 32; ++j) {
        const u64 other = __shfl_sync(0xffffffffu, key, j);
        const float rj = __shfl_sync(0xffffffffu, r, j);
        rank += (other < key) ? 1 : 0;
        if (other != VDB_KEY_MAX) { r_max = fmaxf(r_max, rj); ++n_valid; }
    }
    // the n_probe-th smallest numpy-order distance among the candidates
    const unsigned at = __ballot_sync(0xffffffffu, c >= 0 && rank == n_probe - 1);
    const float n_p = at ? __shfl_sync(0xffffffffu, n, __ffs(at) - 1) : INFINITY;
    const float eps = (float)D * 1.1920929e-7f;
    const bool ok = n_valid >= n_probe && (n_valid >= K || n_p < r_max * (1.f - eps));
    if (lane == 0) flags[qi] = ok ? 0 : 1;
    if (!ok) return;
    int run = 0;
    for (int p = 0; p < n_probe; ++p) {
        const unsigned own = __ballot_sync(0xffffffffu, c >= 0 && rank == p);
        const int cp = __shfl_sync(0xffffffffu, c, __ffs(own) - 1);
        if (lane == 0) {
            probes[(size_t)qi * n_probe + p] = cp;
            if (prefix) prefix[(size_t)qi * (n_probe + 1) + p] = run;
        }
        if (prefix) run += __ldg(list_ptr + cp + 1) - __ldg(list_ptr + cp);
    }
    if (prefix && lane == 0) prefix[(size_t)qi * (n_probe + 1) + n_probe] = run;
}

__device__ __forceinline__ bool ivf_bit(const uint32_t* bm, int r) { return (bm[r >> 5] >> (r & 31)) & 1u; }

// largest p with prefix[p] <= seq  (prefix is non-decreasing, prefix[0] = 0, seq < prefix[n])
__device__ __forceinline__ int ivf_find_probe(const int32_t* prefix, int n, int seq) {
    int lo = 0, hi = n;                 // invariant: prefix[lo] <= seq < prefix[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (prefix[mid] <= seq) lo = mid; else hi = mid;
    }
    return lo;
}

// grid (kIvfSplits, Q).  CTA (s, q) evaluates candidates [s*chunk, (s+1)*chunk) of query q's flattened sequence
// (probe rank major, list position minor -- the enumeration order of ivf.py:177-191); chunk =
// max(chunk_min, ceil(total_q / kIvfSplits)), so a query with few candidates uses few CTAs and one with a huge
// list is still spread over 32 (list sizes after k-means are very uneven).  partial is pre-filled with VDB_KEY_MAX.
__global__ void __launch_bounds__(kIvfThreads)
ivf_scan_kernel(const float* __restrict__ q, const float* __restrict__ db, int D, const int32_t* __restrict__ list_ptr,
                const int32_t* __restrict__ list_rows, const int32_t* __restrict__ probes,
                const int32_t* __restrict__ prefix_g, int n_probe, const uint32_t* __restrict__ alive, int k, int C,
                int chunk_min, u64* __restrict__ partial) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    {
        const int total_q = __ldg(prefix_g + (size_t)blockIdx.y * (n_probe + 1) + n_probe);
        const int chunk_q = max(chunk_min, (total_q + kIvfSplits - 1) / kIvfSplits);
        if ((int)blockIdx.x * chunk_q >= total_q) return;               // nothing for this split
    }
    u64* buf = reinterpret_cast<u64*>(smem_raw);                       // C
    float* qs = reinterpret_cast<float*>(buf + C);                     // D (padded to 4)
    int32_t* prefix = reinterpret_cast<int32_t*>(qs + ((D + 3) & ~3)); // n_probe + 1
    int32_t* pbase = prefix + n_probe + 1;                             // n_probe: list_ptr[probe]
    __shared__ int cnt;
    __shared__ u64 thr;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int qi = blockIdx.y, s = blockIdx.x;
    for (int t = tid; t < D; t += kIvfThreads) qs[t] = q[(size_t)qi * D + t];
    for (int p = tid; p <= n_probe; p += kIvfThreads) prefix[p] = prefix_g[(size_t)qi * (n_probe + 1) + p];
    for (int p = tid; p < n_probe; p += kIvfThreads) pbase[p] = list_ptr[probes[(size_t)qi * n_probe + p]];
    BlockSelect sel;
    sel.init(buf, &cnt, &thr, C, k, kIvfBurst);
    __syncthreads();
    const int total = prefix[n_probe];
    const int chunk = max(chunk_min, (total + kIvfSplits - 1) / kIvfSplits);
    const int begin = min(total, s * chunk), end = min(total, begin + chunk);
    const bool vec4 = (D & 3) == 0;
    const int nchunk = D >> 2;

    int batch = 0;
    for (int base = begin; base < end; base += kIvfBatch) {
        // lane j < kIvfRowsPerWarp decodes the warp's j-th candidate
        int my_seq = base + warp * kIvfRowsPerWarp + lane, my_row = -1;
        if (lane < kIvfRowsPerWarp && my_seq < end) {
            const int p = ivf_find_probe(prefix, n_probe, my_seq);
            const int r = __ldg(list_rows + pbase[p] + (my_seq - prefix[p]));
            if (r >= 0 && (alive == nullptr || ivf_bit(alive, r))) my_row = r;      // deleted from storage: skipped (ivf.py:188-190)
        }
        int rows[kIvfRowsPerWarp];
#pragma unroll
        for (int j = 0; j < kIvfRowsPerWarp; ++j) rows[j] = __shfl_sync(0xffffffffu, my_row, j);
        float acc[kIvfRowsPerWarp];
#pragma unroll
        for (int j = 0; j < kIvfRowsPerWarp; ++j) acc[j] = 0.f;
        if (vec4) {
            for (int c = lane; c < nchunk; c += 32) {
                float4 x[kIvfRowsPerWarp];
#pragma unroll
                for (int j = 0; j < kIvfRowsPerWarp; ++j)
                    x[j] = rows[j] >= 0 ? __ldg(reinterpret_cast<const float4*>(db + (size_t)rows[j] * D) + c)
                                        : make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 qv = *reinterpret_cast<const float4*>(qs + 4 * c);
#pragma unroll
                for (int j = 0; j < kIvfRowsPerWarp; ++j) {
                    const float d0 = qv.x - x[j].x, d1 = qv.y - x[j].y, d2 = qv.z - x[j].z, d3 = qv.w - x[j].w;
                    acc[j] = fmaf(d0, d0, acc[j]); acc[j] = fmaf(d1, d1, acc[j]);
                    acc[j] = fmaf(d2, d2, acc[j]); acc[j] = fmaf(d3, d3, acc[j]);
                }
            }
        } else {
            for (int t = lane; t < D; t += 32) {
                const float qv = qs[t];
#pragma unroll
                for (int j = 0; j < kIvfRowsPerWarp; ++j) {
                    const float d = qv - (rows[j] >= 0 ? __ldg(db + (size_t)rows[j] * D + t) : 0.f);
                    acc[j] = fmaf(d, d, acc[j]);
                }
            }
        }
        float mine = 0.f;
#pragma unroll
        for (int j = 0; j < kIvfRowsPerWarp; ++j) {
            const float v = warp_sum(acc[j]);
            if (lane == j) mine = v;
        }
        if (my_row >= 0) sel.push(pack_key(sqrtf(mine), (uint32_t)my_seq));
        if (++batch == kIvfCheckEvery) { batch = 0; sel.maybe_compact(); }
    }
    sel.finalize();
    u64* dst = partial + ((size_t)qi * kIvfSplits + s) * k;
    for (int i = tid; i < k; i += kIvfThreads) dst[i] = buf[i];
}

// k best (distance, sequence) keys of the S splits; sequence -> row -> external id.  One CTA per query.
__global__ void __launch_bounds__(256)
ivf_merge_kernel(const u64* __restrict__ partial, int L, int k, int C, const int32_t* __restrict__ list_ptr,
                 const int32_t* __restrict__ list_rows, const int32_t* __restrict__ probes,
                 const int32_t* __restrict__ prefix_g, int n_probe, const int64_t* __restrict__ ids,
                 float* __restrict__ out_d, int64_t* __restrict__ out_i, int64_t* __restrict__ out_scanned) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* buf = reinterpret_cast<u64*>(smem_raw);
    __shared__ int cnt;
    __shared__ u64 thr;
    const int qi = blockIdx.x;
    BlockSelect sel;
    sel.init(buf, &cnt, &thr, C, k, blockDim.x);
    __syncthreads();
    const u64* src = partial + (size_t)qi * L;
    for (int base = 0; base < L; base += blockDim.x) {
        const int i = base + threadIdx.x;
        if (i < L) {
            const u64 key = src[i];
            if (key != VDB_KEY_MAX) sel.push(key);
        }
        sel.maybe_compact();
    }
    sel.finalize();
    const int32_t* prefix = prefix_g + (size_t)qi * (n_probe + 1);
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
        const u64 key = buf[i];
        const size_t o = (size_t)qi * k + i;
        if (key == VDB_KEY_MAX) {
            out_d[o] = INFINITY;
            out_i[o] = -1;
        } else {
            const int seq = (int)key_index(key);
            const int p = ivf_find_probe(prefix, n_probe, seq);
            const int r = list_rows[list_ptr[probes[(size_t)qi * n_probe + p]] + (seq - prefix[p])];
            out_d[o] = key_score(key);
            out_i[o] = ids ? ids[r] : (int64_t)r;
        }
    }
    if (out_scanned && threadIdx.x == 0) out_scanned[qi] = prefix[n_probe];
}

// K-means mean update over CSR member lists (rows sorted by label, ascending row inside a label -- a
// deterministic summation order).  Empty cluster keeps its centroid (scipy kmeans2 missing='warn').
__global__ void __launch_bounds__(128)
ivf_centroid_mean_kernel(const float* __restrict__ x, int D, const int32_t* __restrict__ list_ptr,
                         const int32_t* __restrict__ list_rows, float* __restrict__ cent) {
    const int c = blockIdx.x;
    const int b = list_ptr[c], e = list_ptr[c + 1];
    if (e == b) return;
    const float inv = 1.0f / (float)(e - b);
    for (int t = threadIdx.x; t < D; t += blockDim.x) {
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        int i = b;
        for (; i + 4 <= e; i += 4) {
            s0 += __ldg(x + (size_t)list_rows[i] * D + t);
            s1 += __ldg(x + (size_t)list_rows[i + 1] * D + t);
            s2 += __ldg(x + (size_t)list_rows[i + 2] * D + t);
            s3 += __ldg(x + (size_t)list_rows[i + 3] * D + t);
        }
        for (; i < e; ++i) s0 += __ldg(x + (size_t)list_rows[i] * D + t);
        cent[(size_t)c * D + t] = ((s0 + s1) + (s2 + s3)) * inv;
    }
}

static size_t ivf_al(size_t v) { return (v + 255) & ~(size_t)255; }
static int ivf_chunk_min(int Q) { return Q >= 1024 ? 16384 : (Q >= 128 ? 2048 : 256); }

// ---------------------------------------------------------------------------------------------------------------
// List-major batch search (vdb_ivf_topk_tc): the grouping of the (query, probed list) pairs by list and the final ordering,
// on the device -- as torch sorts / scatters around the scan they cost 1.9 ms of a 2.4 ms step with two size read-backs.
//   ivf_group_count_kernel    pairs per (non-empty) list
//   ivf_group_plan_kernel     one CTA: 128-row operand tiles per list -> first operand row of every list, the unit table
//                             (pair tile, first row tile, end row tile) and the totals {operand rows, units}
//   ivf_group_scatter_kernel  every pair takes the next operand row of its list (order inside a list is immaterial)
//   ivf_tc_finish_kernel      one warp per query: list-major positions -> (distance, enumeration sequence) order of
//                             ivf.py:177-193, external ids, and a flag when equal distances straddle the cut
__global__ void ivf_group_count_kernel(const int32_t* __restrict__ probes, int P, const int32_t* __restrict__ tile_ptr, int K,
                                       int32_t* __restrict__ n_l) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const int l = probes[p];
    if ((unsigned)l < (unsigned)K && tile_ptr[l + 1] > tile_ptr[l]) atomicAdd(n_l + l, 1);     // a list id out of range probes nothing
}

__global__ void __launch_bounds__(1024)
ivf_group_plan_kernel(const int32_t* __restrict__ n_l, const int32_t* __restrict__ tile_ptr, int K, int32_t* __restrict__ arow0,
                      int4* __restrict__ utab, int32_t* __restrict__ meta) {
    __shared__ int warp_sums[32];
    __shared__ int base_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) base_s = 0;
    __syncthreads();
    for (int l0 = 0; l0 < K; l0 += 1024) {
        const int l = l0 + tid;
        const int mt = l < K ? (n_l[l] + 127) >> 7 : 0;          // operand tiles of the list
        int incl = mt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) warp_sums[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += v;
            }
            warp_sums[lane] = w;
        }
        __syncthreads();
        const int base = base_s;
        const int excl = base + (warp ? warp_sums[warp - 1] : 0) + incl - mt;      // first unit of the list
        if (l < K) {
            arow0[l] = excl * 128;
            for (int i = 0; i < mt; ++i) utab[excl + i] = make_int4(excl + i, tile_ptr[l], tile_ptr[l + 1], 0);
        }
        __syncthreads();
        if (tid == 1023) base_s = base + warp_sums[31];
        __syncthreads();
    }
    if (tid == 0) { meta[0] = base_s * 128; meta[1] = base_s; }
}

__global__ void ivf_group_scatter_kernel(const int32_t* __restrict__ probes, int P, int n_probe, int K,
                                         const int32_t* __restrict__ tile_ptr, const int32_t* __restrict__ arow0,
                                         int32_t* __restrict__ cursor, int32_t* __restrict__ row_query,
                                         int32_t* __restrict__ row_list, int32_t* __restrict__ pair_rows) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const int l = probes[p];
    int arow = -1;
    if ((unsigned)l < (unsigned)K && tile_ptr[l + 1] > tile_ptr[l]) {
        arow = arow0[l] + atomicAdd(cursor + l, 1);
        row_query[arow] = p / n_probe;
        row_list[arow] = l;
    }
    pair_rows[p] = arow;
}

// k2 <= 32 results per query (list-major positions, ascending distance); lane i owns result i
__global__ void __launch_bounds__(256)
ivf_tc_finish_kernel(const float* __restrict__ d, const int64_t* __restrict__ pos, int Q, int k2, int top_k,
                     const int32_t* __restrict__ probes, const int32_t* __restrict__ prefix, int n_probe,
                     const int32_t* __restrict__ lm_list, const int64_t* __restrict__ lm_start,
                     const int64_t* __restrict__ lm_ids, float* __restrict__ out_d, int64_t* __restrict__ out_ids,
                     int32_t* __restrict__ tie) {
    const int q = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (q >= Q) return;
    float dist = INFINITY;
    int64_t id = -1;
    u64 key = VDB_KEY_MAX;
    if (lane < k2) {
        const int64_t ps = pos[(size_t)q * k2 + lane];
        if (ps >= 0) {
            dist = d[(size_t)q * k2 + lane];
            const int l = lm_list[ps];
            int rank = 0;
            for (int p = 0; p < n_probe; ++p)
                if (probes[(size_t)q * n_probe + p] == l) { rank = p; break; }
            const uint32_t seq = (uint32_t)(prefix[(size_t)q * (n_probe + 1) + rank] + (int)(ps - lm_start[l]));
            key = ((u64)__float_as_uint(dist) << 32) | seq;               // dist >= 0: its bits order like the value
            id = lm_ids[ps];
        }
    }
    int r = 0;                                                              // rank of my key (keys are distinct)
    for (int j = 0; j < 32; ++j) {
        const u64 other = __shfl_sync(0xffffffffu, key, j);
        r += (other < key) ? 1 : 0;
    }
    const bool valid = key != VDB_KEY_MAX;
    if (valid && r < top_k) { out_d[(size_t)q * top_k + r] = dist; out_ids[(size_t)q * top_k + r] = id; }
    // slots past the valid results
    const int n_valid = __popc(__ballot_sync(0xffffffffu, valid));
    if (lane >= n_valid && lane < top_k) { out_d[(size_t)q * top_k + lane] = INFINITY; out_ids[(size_t)q * top_k + lane] = -1; }
    // equal distances on both sides of the cut: which of them the reference keeps depends on rows this call did not fetch
    const unsigned at_cut = __ballot_sync(0xffffffffu, valid && r == top_k - 1);
    const unsigned past = __ballot_sync(0xffffffffu, valid && r == top_k);
    if (at_cut && past) {
        const float da = __shfl_sync(0xffffffffu, dist, __ffs(at_cut) - 1), db = __shfl_sync(0xffffffffu, dist, __ffs(past) - 1);
        if (lane == 0) tie[q] = (da == db) ? 1 : 0;
    } else if (lane == 0) {
        tie[q] = 0;
    }
}

}  // namespace vdb

using namespace vdb;

extern "C" {

int vdb_ivf_centroid_dist(const float* q, int Q, int D, const float* centroids, int K, float* out, void* stream) {
    VDB_REQUIRE(Q >= 0 && D > 0 && K > 0, VDB_ERR_INVALID, "vdb_ivf_centroid_dist: bad shape");
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(q && centroids && out, VDB_ERR_INVALID, "vdb_ivf_centroid_dist: null pointer");
    VDB_REQUIRE(Q <= 65535 * kIvfCdQ, VDB_ERR_UNSUPPORTED, "Q=%d too large per call", Q);
    VDB_REQUIRE(D <= 8192, VDB_ERR_UNSUPPORTED, "D=%d > 8192", D);
    const size_t smem = ((size_t)kIvfCdQ * D + (size_t)kIvfCdThreads * kIvfCdStride) * 4;
    VDB_CUDA(cudaFuncSetAttribute(ivf_centroid_dist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((K + kIvfCdThreads - 1) / kIvfCdThreads, (Q + kIvfCdQ - 1) / kIvfCdQ);
    ivf_centroid_dist_kernel<<<grid, kIvfCdThreads, smem, static_cast<cudaStream_t>(stream)>>>(q, Q, D, centroids, K, out, nullptr);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

int vdb_ivf_probe(const float* cdist, int Q, int K, int n_probe, const int32_t* list_ptr, int32_t* probes,
                  int32_t* prefix, void* stream) {
    VDB_REQUIRE(Q >= 0 && K > 0, VDB_ERR_INVALID, "vdb_ivf_probe: bad shape");
    VDB_REQUIRE(n_probe >= 1 && n_probe <= K, VDB_ERR_INVALID, "n_probe must be between 1 and %d", K);
    VDB_REQUIRE(n_probe <= kIvfMaxProbe, VDB_ERR_UNSUPPORTED, "n_probe=%d > %d", n_probe, kIvfMaxProbe);
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(cdist && probes && (!prefix || list_ptr), VDB_ERR_INVALID, "vdb_ivf_probe: null pointer");
    if (n_probe <= 32 && K <= kIvfProbeWarpMaxK && K >= n_probe && Q >= 64) {
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        if (K <= 1024) ivf_probe_warp_kernel<32><<<(Q + 7) / 8, 256, 0, st>>>(cdist, Q, K, n_probe, list_ptr, probes, prefix, nullptr);
        else if (K <= 2048) ivf_probe_warp_kernel<64><<<(Q + 7) / 8, 256, 0, st>>>(cdist, Q, K, n_probe, list_ptr, probes, prefix, nullptr);
        else ivf_probe_warp_kernel<128><<<(Q + 7) / 8, 256, 0, st>>>(cdist, Q, K, n_probe, list_ptr, probes, prefix, nullptr);
        count_launch(1);
        VDB_CUDA(cudaGetLastError());
        return VDB_OK;
    }
    const int C = select_capacity(n_probe, 256);
    const size_t smem = (size_t)C * 8;
    VDB_CUDA(cudaFuncSetAttribute(ivf_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ivf_probe_kernel<<<Q, 256, smem, static_cast<cudaStream_t>(stream)>>>(cdist, K, n_probe, C, list_ptr, probes, prefix);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

size_t vdb_ivf_scan_workspace(int Q, int k) {
    if (Q <= 0 || k <= 0) return 0;
    return ivf_al((size_t)Q * kIvfSplits * k * 8) + 256;
}

int vdb_ivf_scan(const float* q, int Q, const float* db, int D, const int32_t* list_ptr, const int32_t* list_rows,
                 const int32_t* probes, const int32_t* prefix, int n_probe, const uint32_t* alive, const int64_t* ids,
                 int k, float* out_d, int64_t* out_i, int64_t* out_scanned, void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(Q >= 0 && D > 0 && k > 0 && n_probe >= 1, VDB_ERR_INVALID, "vdb_ivf_scan: bad shape");
    VDB_REQUIRE(n_probe <= kIvfMaxProbe, VDB_ERR_UNSUPPORTED, "n_probe=%d > %d", n_probe, kIvfMaxProbe);
    VDB_REQUIRE(k <= 2048, VDB_ERR_UNSUPPORTED, "k=%d > 2048", k);
    VDB_REQUIRE(Q <= 65535, VDB_ERR_UNSUPPORTED, "Q=%d > 65535 per call", Q);
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(q && db && list_ptr && list_rows && probes && prefix && out_d && out_i, VDB_ERR_INVALID,
                "vdb_ivf_scan: null pointer");
    const int S = kIvfSplits;
    const size_t need = (size_t)Q * S * k * 8;
    VDB_REQUIRE(ws && ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    u64* partial = static_cast<u64*>(ws);
    const int C = select_capacity(k, kIvfBurst);
    const size_t smem = (size_t)C * 8 + (size_t)((D + 3) & ~3) * 4 + (size_t)(2 * n_probe + 1) * 4;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "D=%d k=%d n_probe=%d do not fit shared memory",
                D, k, n_probe);
    VDB_CUDA(cudaFuncSetAttribute(ivf_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    VDB_CUDA(cudaMemsetAsync(partial, 0xFF, need, st));          // VDB_KEY_MAX for the splits that exit early
    {
        ProfScope prof(kProfIvfScan, st);
        ivf_scan_kernel<<<dim3(S, Q), kIvfThreads, smem, st>>>(q, db, D, list_ptr, list_rows, probes, prefix, n_probe,
                                                              alive, k, C, ivf_chunk_min(Q), partial);
    }
    const int C2 = select_capacity(k, 256);
    VDB_CUDA(cudaFuncSetAttribute(ivf_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(C2 * 8)));
    ivf_merge_kernel<<<Q, 256, (size_t)C2 * 8, st>>>(partial, S * k, k, C2, list_ptr, list_rows, probes, prefix, n_probe,
                                                     ids, out_d, out_i, out_scanned);
    count_launch(2);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

int vdb_ivf_centroid_mean(const float* x, int D, const int32_t* list_ptr, const int32_t* list_rows, int K,
                          float* centroids, void* stream) {
    VDB_REQUIRE(D > 0 && K > 0, VDB_ERR_INVALID, "vdb_ivf_centroid_mean: bad shape");
    VDB_REQUIRE(x && list_ptr && list_rows && centroids, VDB_ERR_INVALID, "vdb_ivf_centroid_mean: null pointer");
    ivf_centroid_mean_kernel<<<K, 128, 0, static_cast<cudaStream_t>(stream)>>>(x, D, list_ptr, list_rows, centroids);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

// Probes of a batch from the kc nearest centroids per query found by vdb_exact_topk_tc (cand_d / cand_i, L2, ids = NULL):
// numpy-order re-check of the candidates (ivf_probe_refine_kernel), then the full numpy-order kernels for the flagged
// queries only.  cdist_scratch: Q*K floats (written for flagged queries only); flags: Q ints.
int vdb_ivf_probe_refine(const float* q, int Q, int D, const float* centroids, int K, const float* cand_d, const int64_t* cand_i,
                         int kc, int n_probe, const int32_t* list_ptr, int32_t* probes, int32_t* prefix, float* cdist_scratch,
                         int32_t* flags, void* stream) {
    VDB_REQUIRE(Q > 0 && D > 0 && K > 0 && kc > 0 && kc <= 32 && n_probe >= 1 && n_probe <= kc && n_probe <= K, VDB_ERR_INVALID,
                "vdb_ivf_probe_refine: bad shape (kc=%d n_probe=%d K=%d)", kc, n_probe, K);
    VDB_REQUIRE(q && centroids && cand_d && cand_i && probes && cdist_scratch && flags && (!prefix || list_ptr), VDB_ERR_INVALID,
                "vdb_ivf_probe_refine: null pointer");
    VDB_REQUIRE(D <= 8192 && K <= kIvfProbeWarpMaxK && Q <= 65535 * kIvfCdQ, VDB_ERR_UNSUPPORTED, "vdb_ivf_probe_refine: D=%d K=%d Q=%d", D, K, Q);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ivf_probe_refine_kernel<<<(Q + 7) / 8, 256, 0, st>>>(q, Q, D, centroids, K, cand_d, cand_i, kc, n_probe, list_ptr, probes, prefix, flags);
    const size_t smem = ((size_t)kIvfCdQ * D + (size_t)kIvfCdThreads * kIvfCdStride) * 4;
    VDB_CUDA(cudaFuncSetAttribute(ivf_centroid_dist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((K + kIvfCdThreads - 1) / kIvfCdThreads, (Q + kIvfCdQ - 1) / kIvfCdQ);
    ivf_centroid_dist_kernel<<<grid, kIvfCdThreads, smem, st>>>(q, Q, D, centroids, K, cdist_scratch, flags);
    if (K <= 1024) ivf_probe_warp_kernel<32><<<(Q + 7) / 8, 256, 0, st>>>(cdist_scratch, Q, K, n_probe, list_ptr, probes, prefix, flags);
    else if (K <= 2048) ivf_probe_warp_kernel<64><<<(Q + 7) / 8, 256, 0, st>>>(cdist_scratch, Q, K, n_probe, list_ptr, probes, prefix, flags);
    else ivf_probe_warp_kernel<128><<<(Q + 7) / 8, 256, 0, st>>>(cdist_scratch, Q, K, n_probe, list_ptr, probes, prefix, flags);
    count_launch(3);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

size_t vdb_ivf_group_workspace(int K) { return ivf_al((size_t)K * 4) * 3 + 256; }

int vdb_ivf_group_pairs(const int32_t* probes, int Q, int n_probe, const int32_t* list_tile_ptr, int K, int64_t Qa_max,
                        int32_t* row_query, int32_t* row_list, int32_t* pair_rows, int32_t* unit_table, int32_t* meta,
                        void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(Q > 0 && n_probe > 0 && K > 0 && Qa_max > 0 && Qa_max % 128 == 0, VDB_ERR_INVALID, "vdb_ivf_group_pairs: bad shape");
    VDB_REQUIRE(probes && list_tile_ptr && row_query && row_list && pair_rows && unit_table && meta && ws, VDB_ERR_INVALID,
                "vdb_ivf_group_pairs: null pointer");
    VDB_REQUIRE(ws_bytes >= vdb_ivf_group_workspace(K), VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, vdb_ivf_group_workspace(K));
    const int64_t P64 = (int64_t)Q * n_probe;
    VDB_REQUIRE(P64 < (1ll << 30), VDB_ERR_UNSUPPORTED, "too many (query, probe) pairs");
    // the worst case: every pair's list keeps a partly filled last tile
    const int64_t need = ((P64 + 127) / 128 + (P64 < K ? P64 : K)) * 128;
    VDB_REQUIRE(Qa_max >= need, VDB_ERR_INVALID, "Qa_max=%lld < %lld operand rows in the worst case", (long long)Qa_max, (long long)need);
    const int P = (int)P64;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* base = static_cast<char*>(ws);
    int32_t* n_l = reinterpret_cast<int32_t*>(base);
    int32_t* arow0 = reinterpret_cast<int32_t*>(base + ivf_al((size_t)K * 4));
    int32_t* cursor = reinterpret_cast<int32_t*>(base + 2 * ivf_al((size_t)K * 4));
    VDB_CUDA(cudaMemsetAsync(n_l, 0, (size_t)K * 4, st));
    VDB_CUDA(cudaMemsetAsync(cursor, 0, (size_t)K * 4, st));
    VDB_CUDA(cudaMemsetAsync(row_query, 0xFF, (size_t)Qa_max * 4, st));      // -1: padding of the operand tiles
    ivf_group_count_kernel<<<(P + 255) / 256, 256, 0, st>>>(probes, P, list_tile_ptr, K, n_l);
    ivf_group_plan_kernel<<<1, 1024, 0, st>>>(n_l, list_tile_ptr, K, arow0, reinterpret_cast<int4*>(unit_table), meta);
    ivf_group_scatter_kernel<<<(P + 255) / 256, 256, 0, st>>>(probes, P, n_probe, K, list_tile_ptr, arow0, cursor, row_query, row_list, pair_rows);
    count_launch(3);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

int vdb_ivf_tc_finish(const float* d, const int64_t* pos, int Q, int k2, int top_k, const int32_t* probes, const int32_t* prefix,
                      int n_probe, const int32_t* lm_list, const int64_t* lm_start, const int64_t* lm_ids, float* out_d,
                      int64_t* out_ids, int32_t* tie, void* stream) {
    VDB_REQUIRE(Q > 0 && k2 > 0 && k2 <= 32 && top_k > 0 && top_k < k2 + 1 && n_probe > 0, VDB_ERR_INVALID, "vdb_ivf_tc_finish: bad shape");
    VDB_REQUIRE(d && pos && probes && prefix && lm_list && lm_start && lm_ids && out_d && out_ids && tie, VDB_ERR_INVALID,
                "vdb_ivf_tc_finish: null pointer");
    ivf_tc_finish_kernel<<<(Q + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(d, pos, Q, k2, top_k, probes, prefix, n_probe, lm_list,
                                                                                   lm_start, lm_ids, out_d, out_ids, tie);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"
