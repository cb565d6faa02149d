// Exact fp32 scan (direct-difference form), partial-list finalisation, exact rerank and the
// multi-GPU merge.  These are the "always exact" kernels: the tensor-core scan (exact_tc.cu) uses
// vdb_rerank_f32 for its final answer and vdb_exact_topk_f32 for queries it cannot certify.
//
// Reference arithmetic replaced: src/vector_db/util/distance.py:3-4 (sqrt(sum((q-x)^2)), fp32),
// the (dist,id) sort of tests/indexing/test_hnsw.py:128-134 / src/vector_db/indexing/ivf.py:176-199.
#include <mutex>

#include "common.cuh"

namespace vdb {

constexpr int kScanThreads = 256;
constexpr int kScanWarps = kScanThreads / 32;
constexpr int kCheckEvery = 16;                       // row batches between compaction checks
constexpr int kScanBurst = kScanWarps * kCheckEvery;  // max pushes per query between checks

// Subset mode (device-side list of queries, e.g. the tensor-core scan's uncertified ones): the number of
// queries is only known on the device, so the slicing is derived there -- few queries get many row slices
// (all CTAs busy), many queries get few.  Units = groups * slices <= grid + groups, which bounds `partial`.
struct SubsetPlan { int groups, slices; int64_t slice_rows; };
__host__ __device__ inline SubsetPlan subset_plan(int count, int tq, int grid, int64_t N) {
    SubsetPlan p;
    p.groups = (count + tq - 1) / tq;
    if (p.groups < 1) p.groups = 1;
    int64_t slices = (grid + p.groups - 1) / p.groups;
    int64_t max_slices = (N + 2047) / 2048;
    if (max_slices < 1) max_slices = 1;
    if (slices > max_slices) slices = max_slices;
    if (slices < 1) slices = 1;
    p.slice_rows = (N + slices - 1) / slices;
    p.slice_rows = (p.slice_rows + 7) / 8 * 8;          // multiple of the warps per CTA
    if (p.slice_rows < 8) p.slice_rows = 8;
    p.slices = (int)((N + p.slice_rows - 1) / p.slice_rows);
    if (p.slices < 1) p.slices = 1;
    return p;
}

static int subset_grid() { return 4 * sm_count(); }
// queries-per-CTA tile of the subset scan as a function of the (device-side) count: a 16-query tile spends 16x the
// shared-memory reads on a single uncertified query, so small counts get narrow tiles and more row slices
__host__ __device__ inline int subset_tq(int count) { return count <= 2 ? 1 : (count <= 16 ? 4 : 16); }

// One CTA = TQ queries (resident in shared memory) x one slice of database rows.  Each warp owns a
// row at a time: lanes stride the row with 128-bit loads, accumulate TQ partial sums against the
// shared-memory queries, shuffle-reduce, and lane g pushes query g's (score,row) key into that
// query's BlockSelect.  Output: the k best keys of every (query, slice) in `partial`.
template <int TQ, int METRIC>
__global__ void __launch_bounds__(kScanThreads)
exact_scan_f32_kernel(const float* __restrict__ q, int Q, const float* __restrict__ db,
                      const uint32_t* __restrict__ row_mask, int64_t N, int D, int k, int C,
                      int64_t slice_rows, u64* __restrict__ partial, int S, const int32_t* __restrict__ qsel,
                      const int32_t* __restrict__ qcount) {
    // subset mode: only the queries qsel[0 .. *qcount) are scanned (count read on the device, so the caller
    // never synchronises); the grid is 1-D and walks (group, slice) units of the device-side plan; query SLOT
    // b = position in qsel indexes `partial`
    int n_units = 1;
    if (qsel) {
        Q = min(Q, __ldg(qcount));
        if (Q <= 0 || subset_tq(Q) != TQ) return;           // one launch per tile width; the count picks which one runs
        const SubsetPlan sp = subset_plan(Q, TQ, (int)gridDim.x, N);
        S = sp.slices;
        slice_rows = sp.slice_rows;
        n_units = sp.groups * sp.slices;
    }
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* sel_buf = reinterpret_cast<u64*>(smem_raw);                      // TQ * C
    u64* sel_thr = sel_buf + (size_t)TQ * C;                              // TQ (padded to even)
    float* qs = reinterpret_cast<float*>(sel_thr + ((TQ + 1) & ~1));      // TQ * D, 16-byte aligned
    int* sel_cnt = reinterpret_cast<int*>(qs + (size_t)TQ * D);           // TQ
    float* qnorm = reinterpret_cast<float*>(sel_cnt + TQ);                // TQ
    __shared__ unsigned compact_mask;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int unit = qsel ? (int)blockIdx.x : 0; unit < n_units; unit += qsel ? (int)gridDim.x : 1) {
    const int q0 = (qsel ? unit / S : (int)blockIdx.x) * TQ;
    const int nq = min(TQ, Q - q0);
    const int s = qsel ? unit % S : (int)blockIdx.y;
    const int64_t r_begin = (int64_t)s * slice_rows;
    const int64_t r_end = min(N, r_begin + slice_rows);
    __syncthreads();                                  // previous unit drained before its shared state is reset

    for (int i = tid; i < TQ * D; i += kScanThreads) {
        int g = i / D;
        qs[i] = (g < nq) ? q[(size_t)(qsel ? __ldg(qsel + q0 + g) : q0 + g) * D + (i - g * D)] : 0.f;
    }
    BlockSelect sel[TQ];
#pragma unroll
    for (int g = 0; g < TQ; ++g) sel[g].init(sel_buf + (size_t)g * C, sel_cnt + g, sel_thr + g, C, k, kScanBurst);
    __syncthreads();
    if (METRIC == VDB_METRIC_COSINE) {
        for (int g = warp; g < TQ; g += kScanWarps) {
            float a = 0.f;
            for (int t = lane; t < D; t += 32) a = fmaf(qs[g * D + t], qs[g * D + t], a);
            a = warp_sum(a);
            if (lane == 0) qnorm[g] = sqrtf(a);
        }
        __syncthreads();
    }

    const bool vec4 = (D & 3) == 0;
    const int nchunk = D >> 2;
    int batch = 0;
    for (int64_t rb = r_begin; rb < r_end; rb += kScanWarps) {
        const int64_t r = rb + warp;
        bool live = r < r_end;
        if (live && row_mask) live = (row_mask[r >> 5] >> (r & 31)) & 1u;
        if (live) {
            float acc[TQ];
#pragma unroll
            for (int g = 0; g < TQ; ++g) acc[g] = 0.f;
            float xx = 0.f;
            const float* row = db + r * D;
            if (vec4) {
                const float4* row4 = reinterpret_cast<const float4*>(row);
                for (int c = lane; c < nchunk; c += 32) {
                    const float4 x = ld_stream_f4(row4 + c);
                    if (METRIC == VDB_METRIC_COSINE) {
                        xx = fmaf(x.x, x.x, xx); xx = fmaf(x.y, x.y, xx);
                        xx = fmaf(x.z, x.z, xx); xx = fmaf(x.w, x.w, xx);
                    }
#pragma unroll
                    for (int g = 0; g < TQ; ++g) {
                        const float4 qv = *reinterpret_cast<const float4*>(qs + g * D + 4 * c);
                        if (METRIC == VDB_METRIC_L2) {
                            float d0 = qv.x - x.x, d1 = qv.y - x.y, d2 = qv.z - x.z, d3 = qv.w - x.w;
                            acc[g] = fmaf(d0, d0, acc[g]); acc[g] = fmaf(d1, d1, acc[g]);
                            acc[g] = fmaf(d2, d2, acc[g]); acc[g] = fmaf(d3, d3, acc[g]);
                        } else {
                            acc[g] = fmaf(qv.x, x.x, acc[g]); acc[g] = fmaf(qv.y, x.y, acc[g]);
                            acc[g] = fmaf(qv.z, x.z, acc[g]); acc[g] = fmaf(qv.w, x.w, acc[g]);
                        }
                    }
                }
            } else {
                for (int t = lane; t < D; t += 32) {
                    const float x = row[t];
                    if (METRIC == VDB_METRIC_COSINE) xx = fmaf(x, x, xx);
#pragma unroll
                    for (int g = 0; g < TQ; ++g) {
                        const float qv = qs[g * D + t];
                        if (METRIC == VDB_METRIC_L2) { float d = qv - x; acc[g] = fmaf(d, d, acc[g]); }
                        else acc[g] = fmaf(qv, x, acc[g]);
                    }
                }
            }
            float mine = 0.f;
#pragma unroll
            for (int g = 0; g < TQ; ++g) {
                float v = warp_sum(acc[g]);
                if (lane == g) mine = v;
            }
            if (METRIC == VDB_METRIC_COSINE) xx = warp_sum(xx);
            if (lane < nq) {
                float score;
                if (METRIC == VDB_METRIC_L2) score = sqrtf(mine);
                else score = -(mine / (sqrtf(xx) * qnorm[lane]));
                // select the BlockSelect of query `lane` without dynamic indexing of the struct array
                u64 key = pack_key(score, (uint32_t)r);
                u64 thr = *(volatile u64*)(sel_thr + lane);
                if (key < thr) {
                    int p = atomicAdd(sel_cnt + lane, 1);
                    sel_buf[(size_t)lane * C + k + p] = key;
                }
            }
        }
        if (++batch == kCheckEvery) {
            batch = 0;
            __syncthreads();
            if (tid == 0) {
                unsigned m = 0;
                for (int g = 0; g < TQ; ++g)
                    if (sel_cnt[g] + kScanBurst > C - k) m |= 1u << g;
                compact_mask = m;
            }
            __syncthreads();
            const unsigned m = compact_mask;
            if (m) {
#pragma unroll
                for (int g = 0; g < TQ; ++g)
                    if (m & (1u << g)) sel[g].compact_after_barrier();
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int g = 0; g < TQ; ++g) {
        if (g < nq) {   // CTA-uniform
            sel[g].compact_after_barrier();
            u64* dst = partial + ((size_t)(q0 + g) * S + s) * k;
            for (int i = tid; i < k; i += kScanThreads) dst[i] = sel_buf[(size_t)g * C + i];
        }
    }
    }   // unit loop
}

// Select the k best of L keys per query and translate to (value, id).  One CTA per query.
// transform: 0 = score as is, 1 = negate (cosine similarities were keyed on -sim).
__global__ void __launch_bounds__(256)
finalize_keys_kernel(const u64* __restrict__ keys, int L, int k, int C, const int64_t* __restrict__ ids,
                     int64_t idx_offset, int negate, float* __restrict__ out_d, int64_t* __restrict__ out_i,
                     const int32_t* __restrict__ qsel, const int32_t* __restrict__ qcount, int sub_tq, int sub_grid,
                     int64_t sub_N) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int qo = blockIdx.x;                              // output row; subset mode: slot b -> query qsel[b]
    if (qsel) {
        const int count = __ldg(qcount);
        if ((int)blockIdx.x >= count) return;
        qo = __ldg(qsel + blockIdx.x);
        L = subset_plan(count, subset_tq(count), sub_grid, sub_N).slices * k;      // the scan's device-side slicing
    }
    u64* buf = reinterpret_cast<u64*>(smem_raw);
    __shared__ int cnt;
    __shared__ u64 thr;
    BlockSelect sel;
    sel.init(buf, &cnt, &thr, C, k, blockDim.x);
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
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
        u64 key = buf[i];
        size_t o = (size_t)qo * k + i;
        if (key == VDB_KEY_MAX) {
            out_d[o] = negate ? -INFINITY : INFINITY;
            out_i[o] = -1;
        } else {
            float sc = key_score(key);
            out_d[o] = negate ? -sc : sc;
            int64_t idx = (int64_t)key_index(key);
            out_i[o] = ids ? ids[idx] : idx + idx_offset;
        }
    }
}

// Exact rerank: one CTA per query, one warp per candidate row.
template <int METRIC>
__global__ void __launch_bounds__(256)
rerank_f32_kernel(const float* __restrict__ q, const float* __restrict__ db, const int64_t* __restrict__ ids,
                  int64_t N, int D, const int32_t* __restrict__ cand, const int32_t* __restrict__ cand_cnt, int R_max,
                  int k, int C_max, float* __restrict__ out_d, int64_t* __restrict__ out_i) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* buf = reinterpret_cast<u64*>(smem_raw);          // C_max
    float* qs = reinterpret_cast<float*>(buf + C_max);    // D
    __shared__ float qn_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
    const int qi = blockIdx.x;
    // cand_cnt: only the first cand_cnt[qi] entries of the query's candidate row are meaningful -- the sort
    // then runs over the next power of two of that count instead of R_max
    int R = R_max, C = C_max;
    if (cand_cnt) {
        R = min(R_max, max(0, cand_cnt[qi]));
        C = 32;
        while (C < R) C <<= 1;
    }
    for (int t = tid; t < D; t += blockDim.x) qs[t] = q[(size_t)qi * D + t];
    for (int i = tid; i < C; i += blockDim.x) buf[i] = VDB_KEY_MAX;
    __syncthreads();
    if (METRIC == VDB_METRIC_COSINE) {
        if (warp == 0) {
            float a = 0.f;
            for (int t = lane; t < D; t += 32) a = fmaf(qs[t], qs[t], a);
            a = warp_sum(a);
            if (lane == 0) qn_s = sqrtf(a);
        }
        __syncthreads();
    }
    const bool vec4 = (D & 3) == 0;
    if (vec4) {
        // four candidate rows in flight per warp: their loads are independent, the per-row arithmetic (lane-strided
        // fma chain, then the shuffle tree) is the single-row order -- one row at a time the kernel sat on HBM latency
        constexpr int RW = 4;
        const int nchunk = D >> 2;
        for (int base = warp * RW; base < R; base += nwarp * RW) {
            int my_r = -1;
            if (lane < RW && base + lane < R) {
                my_r = cand[(size_t)qi * R_max + base + lane];
                if ((int64_t)my_r >= N) my_r = -1;
            }
            int rows[RW];
#pragma unroll
            for (int j = 0; j < RW; ++j) rows[j] = __shfl_sync(0xffffffffu, my_r, j);
            if (rows[0] < 0 && rows[1] < 0 && rows[2] < 0 && rows[3] < 0) continue;
            float acc[RW], xx[RW];
#pragma unroll
            for (int j = 0; j < RW; ++j) { acc[j] = 0.f; xx[j] = 0.f; }
            for (int t = lane; t < nchunk; t += 32) {
                float4 x[RW];
#pragma unroll
                for (int j = 0; j < RW; ++j)
                    x[j] = rows[j] >= 0 ? __ldg(reinterpret_cast<const float4*>(db + (size_t)rows[j] * D) + t)
                                        : make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 qv = *reinterpret_cast<const float4*>(qs + 4 * t);
#pragma unroll
                for (int j = 0; j < RW; ++j) {
                    if (METRIC == VDB_METRIC_L2) {
                        const float d0 = qv.x - x[j].x, d1 = qv.y - x[j].y, d2 = qv.z - x[j].z, d3 = qv.w - x[j].w;
                        acc[j] = fmaf(d0, d0, acc[j]); acc[j] = fmaf(d1, d1, acc[j]);
                        acc[j] = fmaf(d2, d2, acc[j]); acc[j] = fmaf(d3, d3, acc[j]);
                    } else {
                        acc[j] = fmaf(qv.x, x[j].x, acc[j]); acc[j] = fmaf(qv.y, x[j].y, acc[j]);
                        acc[j] = fmaf(qv.z, x[j].z, acc[j]); acc[j] = fmaf(qv.w, x[j].w, acc[j]);
                        xx[j] = fmaf(x[j].x, x[j].x, xx[j]); xx[j] = fmaf(x[j].y, x[j].y, xx[j]);
                        xx[j] = fmaf(x[j].z, x[j].z, xx[j]); xx[j] = fmaf(x[j].w, x[j].w, xx[j]);
                    }
                }
            }
            float a_mine = 0.f, x_mine = 0.f;
#pragma unroll
            for (int j = 0; j < RW; ++j) {
                const float a = warp_sum(acc[j]);
                const float b = (METRIC == VDB_METRIC_COSINE) ? warp_sum(xx[j]) : 0.f;
                if (lane == j) { a_mine = a; x_mine = b; }
            }
            if (my_r >= 0) {
                const float score = (METRIC == VDB_METRIC_L2) ? sqrtf(a_mine) : -(a_mine / (sqrtf(x_mine) * qn_s));
                buf[base + lane] = pack_key(score, (uint32_t)my_r);
            }
        }
    } else {
        for (int c = warp; c < R; c += nwarp) {
            const int32_t r = cand[(size_t)qi * R_max + c];
            if (r < 0 || (int64_t)r >= N) continue;
            const float* row = db + (size_t)r * D;
            float acc = 0.f, xx = 0.f;
            for (int t = lane; t < D; t += 32) {
                const float x = __ldg(row + t);
                if (METRIC == VDB_METRIC_L2) { float d = qs[t] - x; acc = fmaf(d, d, acc); }
                else { acc = fmaf(qs[t], x, acc); xx = fmaf(x, x, xx); }
            }
            acc = warp_sum(acc);
            if (METRIC == VDB_METRIC_COSINE) xx = warp_sum(xx);
            if (lane == 0) {
                float score = (METRIC == VDB_METRIC_L2) ? sqrtf(acc) : -(acc / (sqrtf(xx) * qn_s));
                buf[c] = pack_key(score, (uint32_t)r);
            }
        }
    }
    __syncthreads();
    block_bitonic_sort(buf, C);
    if (tid == 0) {
        int n = 0;
        u64 prev = VDB_KEY_MAX;
        for (int i = 0; i < C && n < k; ++i) {
            u64 key = buf[i];
            if (key == VDB_KEY_MAX) break;
            if (key == prev) continue;    // duplicate candidate row
            prev = key;
            float sc = key_score(key);
            out_d[(size_t)qi * k + n] = (METRIC == VDB_METRIC_L2) ? sc : -sc;
            int64_t idx = (int64_t)key_index(key);
            out_i[(size_t)qi * k + n] = ids ? ids[idx] : idx;
            ++n;
        }
        for (; n < k; ++n) {
            out_d[(size_t)qi * k + n] = (METRIC == VDB_METRIC_L2) ? INFINITY : -INFINITY;
            out_i[(size_t)qi * k + n] = -1;
        }
    }
}

// Multi-GPU merge: L sorted lists per query -> k best by (value, list, position).
__global__ void __launch_bounds__(256)
merge_lists_kernel(const float* __restrict__ d, const int64_t* __restrict__ idx, int L, int Q, int k, int C,
                   int descending, float* __restrict__ out_d, int64_t* __restrict__ out_i) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* buf = reinterpret_cast<u64*>(smem_raw);
    const int qi = blockIdx.x;
    const int total = L * k;
    for (int i = threadIdx.x; i < C; i += blockDim.x) {
        u64 key = VDB_KEY_MAX;
        if (i < total) {
            int l = i / k, p = i - l * k;
            size_t src = ((size_t)l * Q + qi) * k + p;
            if (idx[src] >= 0) key = pack_key(descending ? -d[src] : d[src], (uint32_t)i);
        }
        buf[i] = key;
    }
    __syncthreads();
    block_bitonic_sort(buf, C);
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
        u64 key = buf[i];
        size_t o = (size_t)qi * k + i;
        if (key == VDB_KEY_MAX) {
            out_d[o] = descending ? -INFINITY : INFINITY;
            out_i[o] = -1;
        } else {
            int pos = (int)key_index(key);
            int l = pos / k, p = pos - l * k;
            size_t src = ((size_t)l * Q + qi) * k + p;
            out_d[o] = d[src];
            out_i[o] = idx[src];
        }
    }
}

// ------------------------------------------------------------------------------ host helpers
struct ScanPlan {
    int tq, groups, slices, C;
    int64_t slice_rows;
    size_t smem;
};

static bool plan_scan(int Q, int64_t N, int D, int k, ScanPlan* p) {
    const int smem_max = max_smem_optin() - 1024;
    const int C = select_capacity(k, kScanBurst);
    const int tqs[3] = {16, 4, 1};
    for (int t = 0; t < 3; ++t) {
        int tq = tqs[t];
        size_t smem = (size_t)tq * C * 8 + (size_t)((tq + 1) & ~1) * 8 + (size_t)tq * D * 4 + (size_t)tq * 8 + 16;
        if (smem <= (size_t)smem_max) {
            // prefer the widest tile that fits; narrow tiles only when Q is small
            if (tq == 16 && Q <= 4) continue;
            if (tq == 4 && Q == 1) continue;
            p->tq = tq;
            p->C = C;
            p->smem = smem;
            p->groups = (Q + tq - 1) / tq;
            int want = (4 * sm_count() + p->groups - 1) / p->groups;
            int64_t max_slices = (N + 2047) / 2048;
            if (max_slices < 1) max_slices = 1;
            int slices = want < 1 ? 1 : want;
            if (slices > max_slices) slices = (int)max_slices;
            if (slices > 65535) slices = 65535;
            p->slice_rows = (N + slices - 1) / slices;
            p->slice_rows = (p->slice_rows + kScanWarps - 1) / kScanWarps * kScanWarps;
            if (p->slice_rows < kScanWarps) p->slice_rows = kScanWarps;
            p->slices = (int)((N + p->slice_rows - 1) / p->slice_rows);
            if (p->slices < 1) p->slices = 1;
            return true;
        }
    }
    return false;
}

int finalize_keys(const u64* keys, int Q, int L, int k, const int64_t* ids, int64_t idx_offset, int negate,
                  float* out_d, int64_t* out_i, cudaStream_t st, const int32_t* qsel, const int32_t* qcount, int sub_tq,
                  int sub_grid, int64_t sub_N) {
    const int C = select_capacity(k, 256);
    size_t smem = (size_t)C * 8;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "top-k too large for finalize (k=%d)", k);
    VDB_CUDA(cudaFuncSetAttribute(finalize_keys_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope prof(kProfFinalize, st);
    finalize_keys_kernel<<<Q, 256, smem, st>>>(keys, L, k, C, ids, idx_offset, negate, out_d, out_i, qsel, qcount, sub_tq,
                                               sub_grid, sub_N);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

template <int TQ, int METRIC>
static int launch_scan(const ScanPlan& p, const float* q, int Q, const float* db, const uint32_t* mask, int64_t N,
                       int D, int k, u64* partial, cudaStream_t st, const int32_t* qsel = nullptr,
                       const int32_t* qcount = nullptr) {
    auto kern = exact_scan_f32_kernel<TQ, METRIC>;
    VDB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
    dim3 grid(p.groups, p.slices);
    if (qsel) grid = dim3(subset_grid(), 1);
    ProfScope prof(kProfExactF32, st);
    kern<<<grid, kScanThreads, p.smem, st>>>(q, Q, db, mask, N, D, k, p.C, p.slice_rows, partial, p.slices, qsel, qcount);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

size_t exact_f32_subset_workspace(int Q, int64_t N, int D, int k) {
    ScanPlan p;
    if (Q <= 0 || !plan_scan(Q, N, D, k, &p)) return 0;
    // units = groups * slices <= grid + groups for every possible count (subset_plan), TQ <= 16 slots per group
    return ((size_t)16 * subset_grid() + (size_t)Q + 16) * k * 8;
}

// Exact fp32 scan of the queries qsel[0 .. *qcount) only (device-side list), results written to their rows of
// out_d / out_i; the grid covers the worst case (all Q) and CTAs past the count exit at once.
int exact_f32_subset(const float* q, int Q, const float* db, const int64_t* ids, int64_t N, int D, int k, int metric,
                     const int32_t* qsel, const int32_t* qcount, float* out_d, int64_t* out_i, void* ws, size_t ws_bytes,
                     cudaStream_t st) {
    ScanPlan p;
    VDB_REQUIRE(plan_scan(Q, N, D, k, &p), VDB_ERR_UNSUPPORTED, "D=%d k=%d does not fit shared memory", D, k);
    const size_t need = exact_f32_subset_workspace(Q, N, D, k);
    VDB_REQUIRE(ws && ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
    u64* partial = static_cast<u64*>(ws);
    const int tqs[3] = {1, 4, 16};
    for (int t = 0; t < 3; ++t) {
        ScanPlan pt = p;
        pt.tq = tqs[t];
        pt.smem = (size_t)pt.tq * p.C * 8 + (size_t)((pt.tq + 1) & ~1) * 8 + (size_t)pt.tq * D * 4 + (size_t)pt.tq * 8 + 16;
        VDB_REQUIRE(pt.smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED,
                    "subset scan: the %d-query tile does not fit shared memory (D=%d k=%d)", pt.tq, D, k);
        int rc;
#define VDB_LAUNCH(TQ)                                                                                                     \
    rc = (metric == VDB_METRIC_L2) ? launch_scan<TQ, VDB_METRIC_L2>(pt, q, Q, db, nullptr, N, D, k, partial, st, qsel, qcount) \
                                   : launch_scan<TQ, VDB_METRIC_COSINE>(pt, q, Q, db, nullptr, N, D, k, partial, st, qsel, qcount)
        if (pt.tq == 16) { VDB_LAUNCH(16); }
        else if (pt.tq == 4) { VDB_LAUNCH(4); }
        else { VDB_LAUNCH(1); }
#undef VDB_LAUNCH
        if (rc) return rc;
    }
    return finalize_keys(partial, Q, p.slices * k, k, ids, 0, metric == VDB_METRIC_COSINE, out_d, out_i, st, qsel, qcount,
                         0, subset_grid(), N);
}

}  // namespace vdb

using namespace vdb;

extern "C" {

size_t vdb_exact_topk_f32_workspace(int Q, int64_t N, int k) {
    if (Q <= 0 || k <= 0) return 0;
    // worst-case slice count: plan_scan never exceeds max(4*SMs, 1) slices per group... bound generously
    int64_t max_slices = (N + 2047) / 2048;
    if (max_slices < 1) max_slices = 1;
    int64_t s = 4 * (int64_t)sm_count() + 1;
    if (s > max_slices) s = max_slices;
    return (size_t)Q * (size_t)s * (size_t)k * 8 + 256;
}

int vdb_exact_topk_f32(const float* q, int Q, const float* db, const int64_t* ids, const uint32_t* row_mask,
                       int64_t N, int D, int k, int metric, float* out_d, int64_t* out_i, void* ws, size_t ws_bytes,
                       void* stream) {
    VDB_REQUIRE(Q >= 0 && N >= 0 && D > 0 && k > 0, VDB_ERR_INVALID, "vdb_exact_topk_f32: bad shape Q=%d N=%lld D=%d k=%d",
                Q, (long long)N, D, k);
    VDB_REQUIRE(metric == VDB_METRIC_L2 || metric == VDB_METRIC_COSINE, VDB_ERR_INVALID, "unknown metric %d", metric);
    VDB_REQUIRE(N < (1ll << 32), VDB_ERR_UNSUPPORTED, "N=%lld exceeds 2^32 rows per call", (long long)N);
    VDB_REQUIRE(k <= 2048, VDB_ERR_UNSUPPORTED, "k=%d > 2048", k);
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(q && out_d && out_i && (db || N == 0), VDB_ERR_INVALID, "vdb_exact_topk_f32: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ScanPlan p;
    VDB_REQUIRE(plan_scan(Q, N, D, k, &p), VDB_ERR_UNSUPPORTED, "D=%d k=%d does not fit shared memory", D, k);
    size_t need = (size_t)Q * p.slices * k * 8;
    VDB_REQUIRE(ws && ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
    u64* partial = static_cast<u64*>(ws);
    int rc;
#define VDB_LAUNCH(TQ)                                                                                   \
    rc = (metric == VDB_METRIC_L2) ? launch_scan<TQ, VDB_METRIC_L2>(p, q, Q, db, row_mask, N, D, k, partial, st) \
                                   : launch_scan<TQ, VDB_METRIC_COSINE>(p, q, Q, db, row_mask, N, D, k, partial, st)
    if (p.tq == 16) { VDB_LAUNCH(16); }
    else if (p.tq == 4) { VDB_LAUNCH(4); }
    else { VDB_LAUNCH(1); }
#undef VDB_LAUNCH
    if (rc) return rc;
    return finalize_keys(partial, Q, p.slices * k, k, ids, 0, metric == VDB_METRIC_COSINE, out_d, out_i, st);
}

int vdb_rerank_f32(const float* q, int Q, const float* db, const int64_t* ids, int64_t N, int D, const int32_t* cand,
                   int R, int k, int metric, float* out_d, int64_t* out_i, void* stream) {
    return vdb::rerank_f32_counted(q, Q, db, ids, N, D, cand, nullptr, R, k, metric, out_d, out_i, stream);
}

}  // extern "C"

namespace vdb {
int rerank_f32_counted(const float* q, int Q, const float* db, const int64_t* ids, int64_t N, int D, const int32_t* cand,
                       const int32_t* cand_cnt, int R, int k, int metric, float* out_d, int64_t* out_i, void* stream) {
    VDB_REQUIRE(Q >= 0 && D > 0 && R > 0 && k > 0, VDB_ERR_INVALID, "vdb_rerank_f32: bad shape");
    VDB_REQUIRE(metric == VDB_METRIC_L2 || metric == VDB_METRIC_COSINE, VDB_ERR_INVALID, "unknown metric %d", metric);
    VDB_REQUIRE(R <= 16384, VDB_ERR_UNSUPPORTED, "R=%d > 16384", R);
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(q && db && cand && out_d && out_i, VDB_ERR_INVALID, "vdb_rerank_f32: null pointer");
    const int C = next_pow2(R < 32 ? 32 : R);
    size_t smem = (size_t)C * 8 + (size_t)D * 4;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "R=%d D=%d exceed shared memory", R, D);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ProfScope prof(kProfRerank, st);
    if (metric == VDB_METRIC_L2) {
        VDB_CUDA(cudaFuncSetAttribute(rerank_f32_kernel<VDB_METRIC_L2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        rerank_f32_kernel<VDB_METRIC_L2><<<Q, 256, smem, st>>>(q, db, ids, N, D, cand, cand_cnt, R, k, C, out_d, out_i);
    } else {
        VDB_CUDA(cudaFuncSetAttribute(rerank_f32_kernel<VDB_METRIC_COSINE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        rerank_f32_kernel<VDB_METRIC_COSINE><<<Q, 256, smem, st>>>(q, db, ids, N, D, cand, cand_cnt, R, k, C, out_d, out_i);
    }
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}
}  // namespace vdb

extern "C" {

int vdb_merge_topk(const float* d, const int64_t* i, int L, int Q, int k, int descending, float* out_d, int64_t* out_i,
                   void* stream) {
    VDB_REQUIRE(L > 0 && Q >= 0 && k > 0, VDB_ERR_INVALID, "vdb_merge_topk: bad shape");
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(d && i && out_d && out_i, VDB_ERR_INVALID, "vdb_merge_topk: null pointer");
    VDB_REQUIRE((int64_t)L * k <= 16384, VDB_ERR_UNSUPPORTED, "L*k=%lld > 16384", (long long)L * k);
    const int C = next_pow2(L * k < 32 ? 32 : L * k);
    size_t smem = (size_t)C * 8;
    VDB_CUDA(cudaFuncSetAttribute(merge_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    merge_lists_kernel<<<Q, 256, smem, static_cast<cudaStream_t>(stream)>>>(d, i, L, Q, k, C, descending, out_d, out_i);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// Allocation-free variant of vdb_exact_topk_f32 (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, tight serving loops).  The scan's scratch --
// the per-(query, slice) partial top-k keys -- lives in this file-scope device array instead of a caller workspace.
// Every GPU that loads the library holds these 8 MiB until the process exits.
#define VDB_STATIC_SCRATCH_BYTES (8u << 20)
static __device__ __align__(256) unsigned char g_static_scratch[VDB_STATIC_SCRATCH_BYTES];

namespace vdb {

// In HOST code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is asked of
// the runtime once per device, with that device current, and cached.
static int static_scratch_ptr(void** out) {
    constexpr int kMaxDevices = 64;
    static void* ptrs[kMaxDevices] = {nullptr};
    static std::mutex mu;
    int dev = -1;
    VDB_CUDA(cudaGetDevice(&dev));
    VDB_REQUIRE(dev >= 0 && dev < kMaxDevices, VDB_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(mu);
    if (ptrs[dev] == nullptr) {
        void* p = nullptr;
        VDB_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        VDB_REQUIRE(p != nullptr, VDB_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address on device %d", dev);
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return VDB_OK;
}

}  // namespace vdb

extern "C" {

size_t vdb_exact_topk_f32_static_max_bytes(void) { return (size_t)VDB_STATIC_SCRATCH_BYTES; }

int vdb_exact_topk_f32_static(const float* q, int Q, const float* db, const int64_t* ids, const uint32_t* row_mask,
                              int64_t N, int D, int k, int metric, float* out_d, int64_t* out_i, void* stream) {
    VDB_REQUIRE(Q >= 0 && N >= 0 && D > 0 && k > 0, VDB_ERR_INVALID, "vdb_exact_topk_f32_static: bad shape Q=%d N=%lld D=%d k=%d",
                Q, (long long)N, D, k);
    VDB_REQUIRE(metric == VDB_METRIC_L2 || metric == VDB_METRIC_COSINE, VDB_ERR_INVALID, "unknown metric %d", metric);
    VDB_REQUIRE(N < (1ll << 32), VDB_ERR_UNSUPPORTED, "N=%lld exceeds 2^32 rows per call", (long long)N);
    VDB_REQUIRE(k <= 2048, VDB_ERR_UNSUPPORTED, "k=%d > 2048", k);
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(q && out_d && out_i && (db || N == 0), VDB_ERR_INVALID, "vdb_exact_topk_f32_static: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ScanPlan p;
    VDB_REQUIRE(plan_scan(Q, N, D, k, &p), VDB_ERR_UNSUPPORTED, "D=%d k=%d does not fit shared memory", D, k);
    const size_t need = (size_t)Q * p.slices * k * 8;
    VDB_REQUIRE(need <= (size_t)VDB_STATIC_SCRATCH_BYTES, VDB_ERR_UNSUPPORTED,
                "vdb_exact_topk_f32_static: Q=%d k=%d needs %zu bytes of scratch (Q x %d slices x k x 8), the built-in "
                "scratch holds %u; use vdb_exact_topk_f32 with a caller workspace", Q, k, need, p.slices,
                (unsigned)VDB_STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    int rc = static_scratch_ptr(&scratch);
    if (rc) return rc;
    u64* partial = static_cast<u64*>(scratch);
    // the array keeps the previous call's keys: start every slot this call reads at "empty" (all ones)
    VDB_CUDA(cudaMemsetAsync(partial, 0xFF, need, st));
#define VDB_LAUNCH(TQ)                                                                                   \
    rc = (metric == VDB_METRIC_L2) ? launch_scan<TQ, VDB_METRIC_L2>(p, q, Q, db, row_mask, N, D, k, partial, st) \
                                   : launch_scan<TQ, VDB_METRIC_COSINE>(p, q, Q, db, row_mask, N, D, k, partial, st)
    if (p.tq == 16) { VDB_LAUNCH(16); }
    else if (p.tq == 4) { VDB_LAUNCH(4); }
    else { VDB_LAUNCH(1); }
#undef VDB_LAUNCH
    if (rc) return rc;
    return finalize_keys(partial, Q, p.slices * k, k, ids, 0, metric == VDB_METRIC_COSINE, out_d, out_i, st);
}

}  // extern "C"
