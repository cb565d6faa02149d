// Batched HNSW traversal (a4/a5/a6) and sequential-semantics construction (a7).
//
// Reference logic replaced: src/vector_db/indexing/hnsw.py
//   :79-134  _search_layer   best-first beam; stop when len(found) >= ef and d > worst(found);
//                            filter_ids gates RESULTS only, navigation ignores it
//   :289-328 _greedy_search_level   strict '<' hill climb, scans all neighbours before moving
//   :330-374 search          descent on levels max_level..1, beam on level 0, sort (dist,id), top-k
//   :143-225 insert_node     (hnsw_insert_kernel below)
//
// One CTA (128 threads) per query.  The frontier (reference: heapq `candidates`) and the result
// list (`found_nodes`) are SORTED ARRAYS of packed (dist,row) keys in shared memory; each expansion
// step gathers the unvisited neighbours, evaluates their distances one warp per row (direct
// difference, fp32, sqrt -- util/distance.py:3-4), rank-sorts the new keys and rank-merges them
// into both lists, so there is no serial heap work.  Dropping frontier entries that are already
// worse than a full result list is exact: the reference would pop them only to hit its break.
#include "common.cuh"

namespace vdb {

constexpr int kHnswThreads = 128;
constexpr int kHnswWarps = kHnswThreads / 32;
constexpr uint32_t kEmpty = 0xFFFFFFFFu;

struct Graph {
    const float* db;
    int64_t N;
    int D;
    const int32_t* adj0;
    int deg0;
    const int32_t* adj_up;
    int deg_up;
    const int32_t* up_ptr;
    const uint32_t* alive;
    const int32_t* levels;     // per-row top level (nullable): a list above it does not exist
};

struct BeamMem {
    float* qs;        // D          query (shared)
    u64* cand[2];     // CC each    frontier, ascending
    u64* found[2];    // EF each    results, ascending
    u64* newk;        // NM         unsorted new keys
    u64* news;        // NM         sorted new keys
    int* todo;        // NM         rows to evaluate
    int* ppass;       // NM + 1     #filter-passing keys before sorted position i
    uint32_t* hash;   // H          visited set (shared or global)
    int* scal;        // [0] todo counter, [1] overflow flag
    int CC, EF, NM, H, logH;
};

__device__ __forceinline__ bool bit_set(const uint32_t* bm, int r) { return (bm[r >> 5] >> (r & 31)) & 1u; }

__device__ __forceinline__ const int32_t* nbr_list(const Graph& g, int node, int level, int* deg) {
    if (level == 0) { *deg = g.deg0; return g.adj0 + (size_t)node * g.deg0; }
    // The reference creates neighbors[l] only for l <= the top level at insert time (hnsw.py:179-191), so the entry
    // point can lack its top lists; a level above the rows reserved for `node` reads as an empty list
    // (neighbors.get(level, []), hnsw.py:113) instead of running into the next node's rows.
    if (g.levels != nullptr && level > g.levels[node]) { *deg = 0; return g.adj_up; }
    *deg = g.deg_up;
    return g.adj_up + ((size_t)g.up_ptr[node] + level - 1) * g.deg_up;
}

// all lanes return sqrt(sum((q-x)^2))
__device__ __forceinline__ float warp_l2(const float* __restrict__ qs, const float* __restrict__ row, int D, int lane) {
    float acc = 0.f;
    if ((D & 3) == 0) {
        const float4* r4 = reinterpret_cast<const float4*>(row);
        const float4* q4 = reinterpret_cast<const float4*>(qs);
        for (int c = lane; c < (D >> 2); c += 32) {
            const float4 x = ld_gather_f4(r4 + c);
            const float4 qv = q4[c];
            float d0 = qv.x - x.x, d1 = qv.y - x.y, d2 = qv.z - x.z, d3 = qv.w - x.w;
            acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc); acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
        }
    } else {
        for (int t = lane; t < D; t += 32) { float d = qs[t] - __ldg(row + t); acc = fmaf(d, d, acc); }
    }
    return sqrtf(warp_sum(acc));
}

// two rows at once: both rows' loads are issued before either reduction, doubling the bytes in flight
// per warp (the traversal is latency bound: ncu shows the FADDs of this loop waiting on the gathers)
__device__ __forceinline__ void warp_l2_x2(const float* __restrict__ qs, const float* __restrict__ row_a,
                                           const float* __restrict__ row_b, int D, int lane, float* da, float* db) {
    float acc_a = 0.f, acc_b = 0.f;
    if ((D & 3) == 0) {
        const float4* a4 = reinterpret_cast<const float4*>(row_a);
        const float4* b4 = reinterpret_cast<const float4*>(row_b);
        const float4* q4 = reinterpret_cast<const float4*>(qs);
        for (int c = lane; c < (D >> 2); c += 32) {
            const float4 xa = ld_gather_f4(a4 + c);             // read once: no L1 allocation (keeps adjacency / stack lines)
            const float4 xb = ld_gather_f4(b4 + c);
            const float4 qv = q4[c];
            float d0 = qv.x - xa.x, d1 = qv.y - xa.y, d2 = qv.z - xa.z, d3 = qv.w - xa.w;
            acc_a = fmaf(d0, d0, acc_a); acc_a = fmaf(d1, d1, acc_a); acc_a = fmaf(d2, d2, acc_a); acc_a = fmaf(d3, d3, acc_a);
            d0 = qv.x - xb.x; d1 = qv.y - xb.y; d2 = qv.z - xb.z; d3 = qv.w - xb.w;
            acc_b = fmaf(d0, d0, acc_b); acc_b = fmaf(d1, d1, acc_b); acc_b = fmaf(d2, d2, acc_b); acc_b = fmaf(d3, d3, acc_b);
        }
    } else {
        for (int t = lane; t < D; t += 32) {
            const float q = qs[t];
            const float ea = q - __ldg(row_a + t), eb = q - __ldg(row_b + t);
            acc_a = fmaf(ea, ea, acc_a);
            acc_b = fmaf(eb, eb, acc_b);
        }
    }
    *da = sqrtf(warp_sum(acc_a));
    *db = sqrtf(warp_sum(acc_b));
}

__device__ __forceinline__ bool visited_insert(uint32_t* tab, int H, int logH, uint32_t key) {
    uint32_t h = (key * 2654435761u) >> (32 - logH);
    for (int probe = 0; probe < H; ++probe) {
        uint32_t prev = atomicCAS(tab + h, kEmpty, key);
        if (prev == kEmpty) return true;
        if (prev == key) return false;
        h = (h + 1) & (uint32_t)(H - 1);
    }
    return false;   // table full: treated as visited; the caller's load-factor check flags overflow first
}

// number of entries of sorted[0..n) that are < key
__device__ __forceinline__ int lower_bound_u64(const u64* sorted, int n, u64 key) {
    int lo = 0, hi = n;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (sorted[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// The traversal code below is written for a GROUP of NT threads working on one query with its own BeamMem: NT = 128 = the
// whole CTA in every kernel of this file (NT = 32, one warp per query, was measured and dropped -- see hnsw_search_kernel).
// group_sync<NT>() is the group's barrier, tid / warp are indices within the group.
template <int NT>
__device__ __forceinline__ void group_sync() {
    if (NT == 32) __syncwarp(); else __syncthreads();
}

// group-wide: dist of query to `node`, returned to every thread (one warp computes, broadcast via smem)
template <int NT>
__device__ __forceinline__ float cta_dist_one(const Graph& g, BeamMem& m, int node) {
    const int lane = threadIdx.x & 31, warp = (threadIdx.x % NT) >> 5;
    if (warp == 0) {
        float d = warp_l2(m.qs, g.db + (size_t)node * g.D, g.D, lane);
        if (lane == 0) m.newk[0] = pack_key(d, 0);
    }
    group_sync<NT>();
    float d = key_score(m.newk[0]);
    group_sync<NT>();
    return d;
}

// a5: strict-improvement hill climb on one upper level.  Returns the local minimum; *evals += work.
template <int NT>
__device__ int greedy_level(const Graph& g, BeamMem& m, int cur, int level, unsigned* evals) {
    constexpr int kNW = NT / 32;
    const int tid = threadIdx.x % NT, lane = tid & 31, warp = tid >> 5;
    float cur_d = cta_dist_one<NT>(g, m, cur);
    *evals += 1;
    while (true) {
        int deg;
        const int32_t* nb = nbr_list(g, cur, level, &deg);
        if (tid == 0) m.scal[0] = 0;
        group_sync<NT>();
        // keep list order: position p stays the tie-break ("first strictly smaller wins")
        for (int p = tid; p < deg; p += NT) {
            int r = __ldcg(nb + p);
            bool ok = r >= 0 && (g.alive == nullptr || bit_set(g.alive, r));
            m.todo[p] = ok ? r : -1;
            if (ok) atomicAdd(m.scal, 1);
        }
        group_sync<NT>();
        *evals += (unsigned)m.scal[0];
        for (int p = warp; p < deg; p += kNW) {
            int r = m.todo[p];
            u64 key = VDB_KEY_MAX;
            if (r >= 0) {
                float d = warp_l2(m.qs, g.db + (size_t)r * g.D, g.D, lane);
                key = pack_key(d, (uint32_t)p);
            }
            if (lane == 0) m.newk[p] = key;
        }
        group_sync<NT>();
        // min over (d, position)
        u64 best = VDB_KEY_MAX;
        for (int p = 0; p < deg; ++p) { u64 kx = m.newk[p]; best = kx < best ? kx : best; }
        int next = cur;
        if (best != VDB_KEY_MAX && key_score(best) < cur_d) { cur_d = key_score(best); next = m.todo[key_index(best)]; }
        group_sync<NT>();
        if (next == cur) return cur;
        cur = next;
    }
}

// a4: best-first beam on one level.  On return found[*which_found][0..return value) holds the
// (dist,row) keys ascending.  `filter` gates results only.  stats[0] += evals, stats[1] += expansions.
template <int NT>
__device__ int beam_search(const Graph& g, BeamMem& m, int entry, int level, int ef, const uint32_t* filter,
                           int* which_found, unsigned* evals, unsigned* expansions) {
    constexpr int kNW = NT / 32;
    const int tid = threadIdx.x % NT, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < m.H; i += NT) m.hash[i] = kEmpty;
    group_sync<NT>();
    const float d0 = cta_dist_one<NT>(g, m, entry);
    *evals += 1;
    int cc = 0, cf = 0;           // current buffers
    int ncand = 1, nfound = 0;
    if (tid == 0) {
        m.cand[0][0] = pack_key(d0, (uint32_t)entry);
        visited_insert(m.hash, m.H, m.logH, (uint32_t)entry);
        m.scal[1] = 0;
        m.scal[2] = 0;
    }
    if (filter == nullptr || bit_set(filter, entry)) {
        if (tid == 0) m.found[0][0] = pack_key(d0, (uint32_t)entry);
        nfound = 1;
    }
    unsigned nvisited = 1;
    group_sync<NT>();

    while (ncand > 0) {
        const u64 top = m.cand[cc][0];
        const float top_d = key_score(top);
        if (nfound >= ef && top_d > key_score(m.found[cf][nfound - 1])) break;
        const int node = (int)key_index(top);
        *expansions += 1;
        int deg;
        const int32_t* nb = nbr_list(g, node, level, &deg);
        if (tid == 0) m.scal[0] = 0;
        // the node expanded NEXT is most often the current runner-up: pull its adjacency row towards the SM while this
        // expansion's distances are evaluated (the row's load heads the next expansion's critical path)
        if (tid == 32 && ncand > 1) {
            int d2;
            const int32_t* nx = nbr_list(g, (int)key_index(m.cand[cc][1]), level, &d2);
            if (d2 > 0) {
                asm volatile("prefetch.global.L2 [%0];" ::"l"(nx));
                if (d2 > 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + 32));
            }
        }
        group_sync<NT>();
        for (int p = tid; p < deg; p += NT) {
            int r = __ldcg(nb + p);
            if (r >= 0 && visited_insert(m.hash, m.H, m.logH, (uint32_t)r)) {
                if (g.alive == nullptr || bit_set(g.alive, r)) {     // tombstones: visited, never evaluated
                    int slot = atomicAdd(m.scal, 1);
                    m.todo[slot] = r;
                } else {
                    atomicAdd(m.scal + 2, 1);
                }
            }
        }
        group_sync<NT>();
        const int ntodo = m.scal[0];
        *evals += (unsigned)ntodo;
        nvisited += (unsigned)ntodo;
        for (int i = 2 * warp; i < ntodo; i += 2 * kNW) {
            const int r = m.todo[i];
            if (i + 1 < ntodo) {
                const int r2 = m.todo[i + 1];
                float d, d2;
                warp_l2_x2(m.qs, g.db + (size_t)r * g.D, g.db + (size_t)r2 * g.D, g.D, lane, &d, &d2);
                if (lane == 0) { m.newk[i] = pack_key(d, (uint32_t)r); m.newk[i + 1] = pack_key(d2, (uint32_t)r2); }
            } else {
                float d = warp_l2(m.qs, g.db + (size_t)r * g.D, g.D, lane);
                if (lane == 0) m.newk[i] = pack_key(d, (uint32_t)r);
            }
        }
        group_sync<NT>();
        // rank sort of the new keys; ppass[i] = number of filter-passing keys before sorted slot i
        for (int i = tid; i < ntodo; i += NT) {
            const u64 mine = m.newk[i];
            int rank = 0, prank = 0;
            for (int j = 0; j < ntodo; ++j) {
                const u64 other = m.newk[j];
                if (other < mine) {
                    ++rank;
                    if (filter == nullptr || bit_set(filter, (int)key_index(other))) ++prank;
                }
            }
            m.news[rank] = mine;
            m.ppass[rank] = prank;
        }
        group_sync<NT>();
        int npass = ntodo;
        if (filter != nullptr) {
            npass = 0;
            if (ntodo > 0) {
                const u64 last = m.news[ntodo - 1];
                npass = m.ppass[ntodo - 1] + (bit_set(filter, (int)key_index(last)) ? 1 : 0);
            }
        }
        if (tid == 0) m.ppass[ntodo] = npass;
        group_sync<NT>();

        // ---- merge into found (keep the ef smallest)
        const u64* fsrc = m.found[cf];
        u64* fdst = m.found[cf ^ 1];
        for (int i = tid; i < nfound; i += NT) {
            const u64 key = fsrc[i];
            const int lb = lower_bound_u64(m.news, ntodo, key);
            const int pos = i + m.ppass[lb];
            if (pos < ef) fdst[pos] = key;
        }
        for (int i = tid; i < ntodo; i += NT) {
            const u64 key = m.news[i];
            if (filter == nullptr || bit_set(filter, (int)key_index(key))) {
                const int pos = m.ppass[i] + lower_bound_u64(fsrc, nfound, key);
                if (pos < ef) fdst[pos] = key;
            }
        }
        int nfound_new = nfound + npass;
        if (nfound_new > ef) nfound_new = ef;
        group_sync<NT>();
        cf ^= 1;
        nfound = nfound_new;

        // ---- merge into the frontier: pop the head, drop what can never be expanded
        const bool full = nfound >= ef;
        const float worst = full ? key_score(m.found[cf][nfound - 1]) : INFINITY;
        const u64* csrc = m.cand[cc] + 1;          // popped head removed
        const int nold = ncand - 1;
        u64* cdst = m.cand[cc ^ 1];
        // both inputs are ascending, so "d <= worst" keeps a prefix of each
        int keep_old = nold, keep_new = ntodo;
        if (full) {
            const u64 limit = pack_key(worst, 0xFFFFFFFFu);   // largest key with score == worst
            keep_old = lower_bound_u64(csrc, nold, limit);
            if (keep_old < nold && csrc[keep_old] == limit) ++keep_old;
            keep_new = lower_bound_u64(m.news, ntodo, limit);
            if (keep_new < ntodo && m.news[keep_new] == limit) ++keep_new;
        }
        for (int i = tid; i < keep_old; i += NT) {
            const u64 key = csrc[i];
            const int pos = i + lower_bound_u64(m.news, keep_new, key);
            if (pos < m.CC) cdst[pos] = key;
        }
        for (int i = tid; i < keep_new; i += NT) {
            const u64 key = m.news[i];
            const int pos = i + lower_bound_u64(csrc, keep_old, key);
            if (pos < m.CC) cdst[pos] = key;
        }
        int ncand_new = keep_old + keep_new;
        if (ncand_new > m.CC) {
            // Only possible with a filter (without one the frontier never outgrows ef): the reference's heap is
            // unbounded, so entries dropped here could have been expanded -- flag the query, the host finishes it
            // with the exact masked scan over the allowed rows (ops.hnsw_search).
            ncand_new = m.CC;
            if (tid == 0) m.scal[1] |= 2;
        }
        group_sync<NT>();
        cc ^= 1;
        ncand = ncand_new;
        if (nvisited + (unsigned)m.scal[2] > (unsigned)(m.H - (m.H >> 2))) {   // visited table above 75% load
            if (tid == 0) m.scal[1] |= 1;
            break;
        }
    }
    group_sync<NT>();
    *which_found = cf;
    return nfound;
}

__device__ __forceinline__ void carve(BeamMem& m, unsigned char* smem, int D, int CC, int EF, int NM, int H,
                                      uint32_t* global_hash) {
    u64* p = reinterpret_cast<u64*>(smem);
    m.cand[0] = p; p += CC;
    m.cand[1] = p; p += CC;
    m.found[0] = p; p += EF;
    m.found[1] = p; p += EF;
    m.newk = p; p += NM;
    m.news = p; p += NM;
    float* f = reinterpret_cast<float*>(p);
    m.qs = f; f += (D + 3) & ~3;
    int* ip = reinterpret_cast<int*>(f);
    m.todo = ip; ip += NM;
    m.ppass = ip; ip += NM + 1;
    m.scal = ip; ip += 4;
    m.hash = global_hash ? global_hash : reinterpret_cast<uint32_t*>(ip);
    m.CC = CC; m.EF = EF; m.NM = NM; m.H = H;
    int l = 0;
    while ((1 << l) < H) ++l;
    m.logH = l;
}

__host__ __device__ __forceinline__ size_t beam_smem_bytes_dev(int D, int CC, int EF, int NM, int H_in_smem) {
    size_t b = (size_t)(2 * CC + 2 * EF + 2 * NM) * 8 + (size_t)((D + 3) & ~3) * 4 + (size_t)(NM + NM + 1 + 4) * 4 +
               (size_t)H_in_smem * 4 + 16;
    return (b + 15) & ~(size_t)15;
}
static size_t beam_smem_bytes(int D, int CC, int EF, int NM, int H_in_smem) {
    return beam_smem_bytes_dev(D, CC, EF, NM, H_in_smem);
}
// 12 resident CTAs per SM (40 registers): the traversal is bound by the row gathers it keeps in flight, and the number of
// resident warps is what sets that -- 8 CTAs (56 registers): 40.2 ms per 10k queries at cfg 4, 9: 37.9, 10: 35.8, 12: 34.1,
// 14-16 (32 registers): 33.7-34.1.  (One WARP per query, no CTA barrier at all, measured 66 ms: the per-query sort / merge
// phases done by 32 lanes cost what the barriers saved.)
__global__ void __launch_bounds__(kHnswThreads, 12)
hnsw_search_kernel(Graph g, const float* __restrict__ q, int Q, int32_t entry_row, int max_level, int ef, int k,
                   const uint32_t* __restrict__ filter, int CC, int NM, int H, uint32_t* global_hash,
                   float* __restrict__ out_d, int32_t* __restrict__ out_rows, uint32_t* __restrict__ out_stats,
                   int32_t* __restrict__ out_status, int* __restrict__ overflow_flag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BeamMem m;
    carve(m, smem_raw, g.D, CC, ef, NM, H, global_hash ? global_hash + (size_t)blockIdx.x * H : nullptr);
    const int tid = threadIdx.x;
    for (int qi = blockIdx.x; qi < Q; qi += gridDim.x) {
        __syncthreads();
        for (int t = tid; t < g.D; t += kHnswThreads) m.qs[t] = q[(size_t)qi * g.D + t];
        __syncthreads();
        unsigned evals = 0, expansions = 0;
        int nfound = 0, wf = 0;
        if (entry_row >= 0) {
            int cur = entry_row;
            for (int level = max_level; level >= 1; --level) cur = greedy_level<kHnswThreads>(g, m, cur, level, &evals);
            nfound = beam_search<kHnswThreads>(g, m, cur, 0, ef, filter, &wf, &evals, &expansions);
            if (m.scal[1] && tid == 0) atomicOr(overflow_flag, m.scal[1]);
            if (out_status && tid == 0) out_status[qi] = m.scal[1];      // bit 0 visited table full, bit 1 frontier cut
        } else if (out_status && tid == 0) {
            out_status[qi] = 0;
        }
        for (int i = tid; i < k; i += kHnswThreads) {
            size_t o = (size_t)qi * k + i;
            if (i < nfound) {
                u64 key = m.found[wf][i];
                out_d[o] = key_score(key);
                out_rows[o] = (int32_t)key_index(key);
            } else {
                out_d[o] = INFINITY;
                out_rows[o] = -1;
            }
        }
        if (out_stats && tid == 0) { out_stats[2 * qi] = evals; out_stats[2 * qi + 1] = expansions; }
    }
}

}  // namespace vdb

namespace vdb {

// ------------------------------------------------------------------------------------------
// a7: sequential-semantics insert (hnsw.py:143-225) by ONE CTA.  Rows first..first+count-1 are
// inserted in order, each against the graph left by the previous one, so the result equals the
// reference's node-by-node build (same level draws, same M-closest links, same closest-cap
// truncation, same list order: append order, or (dist,id) order after a truncation).
__device__ __forceinline__ int32_t* nbr_list_rw(const Graph& g, int node, int level, int* deg) {
    return const_cast<int32_t*>(nbr_list(g, node, level, deg));
}

// NT threads: the insert is ONE CTA of serial steps (sequential semantics), so the only parallelism is inside a step --
// rows evaluated per round (one warp each), rank sorts, list copies.
template <int NT>
__global__ void __launch_bounds__(NT)
hnsw_insert_kernel(Graph g, const int32_t* __restrict__ levels, int32_t first, int32_t count, int M, int efc,
                   int32_t* __restrict__ state, int CC, int NM, int H, uint32_t* global_hash,
                   int* __restrict__ overflow_flag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BeamMem m;
    carve(m, smem_raw, g.D, CC, efc, NM, H, global_hash);
    // second vector buffer (the neighbour whose list is being truncated) sits after the beam memory
    float* qs2 = reinterpret_cast<float*>(smem_raw + beam_smem_bytes_dev(g.D, CC, efc, NM, global_hash ? 0 : H));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long total_evals = ((unsigned long long)(uint32_t)state[3] << 32) | (uint32_t)state[2];

    for (int it = 0; it < count; ++it) {
        const int node = first + it;
        const int lvl = levels[node];
        __syncthreads();
        for (int t = tid; t < g.D; t += NT) m.qs[t] = g.db[(size_t)node * g.D + t];
        const int entry = ((volatile int32_t*)state)[0], top = ((volatile int32_t*)state)[1];
        __syncthreads();
        if (entry < 0) {                       // first node: becomes the entry point, no links
            if (tid == 0) { state[0] = node; state[1] = lvl; __threadfence(); }
            __syncthreads();
            continue;
        }
        unsigned evals = 0, expansions = 0;
        int cur = entry;
        for (int level = top; level > lvl; --level) cur = greedy_level<NT>(g, m, cur, level, &evals);
        for (int level = (lvl < top ? lvl : top); level >= 0; --level) {
            int wf = 0;
            const int nf = beam_search<NT>(g, m, cur, level, efc, nullptr, &wf, &evals, &expansions);
            if (m.scal[1] && tid == 0) atomicOr(overflow_flag, 1);
            const u64* found = m.found[wf];
            const int nsel = nf < M ? nf : M;
            int deg;
            int32_t* mine = nbr_list_rw(g, node, level, &deg);
            for (int t = tid; t < deg; t += NT) mine[t] = t < nsel ? (int32_t)key_index(found[t]) : -1;
            __threadfence();
            __syncthreads();
            for (int sidx = 0; sidx < nsel; ++sidx) {
                const int nb = (int)key_index(found[sidx]);
                int cap;
                int32_t* lst = nbr_list_rw(g, nb, level, &cap);
                const int len = __syncthreads_count(tid < cap && __ldcg(lst + tid) >= 0);
                if (len < cap) {
                    if (tid == 0) { lst[len] = node; __threadfence(); }
                    __syncthreads();
                } else {
                    // truncate to the `cap` closest of (old links + new node), measured from nb
                    for (int t = tid; t < g.D; t += NT) qs2[t] = g.db[(size_t)nb * g.D + t];
                    __syncthreads();
                    if (tid == 0) m.scal[0] = 0;
                    for (int c = warp; c <= cap; c += (NT / 32)) {
                        const int other = c < cap ? __ldcg(lst + c) : node;
                        u64 key = VDB_KEY_MAX;
                        if (other >= 0 && (g.alive == nullptr || bit_set(g.alive, other))) {
                            float d = warp_l2(qs2, g.db + (size_t)other * g.D, g.D, lane);
                            key = pack_key(d, (uint32_t)other);
                        }
                        if (lane == 0) m.newk[c] = key;
                    }
                    __syncthreads();
                    const int n = cap + 1;
                    int valid = 0;
                    for (int c = 0; c < n; ++c) valid += (m.newk[c] != VDB_KEY_MAX) ? 1 : 0;
                    evals += (unsigned)valid;
                    for (int c = tid; c < n; c += NT) {
                        const u64 me = m.newk[c];
                        if (me != VDB_KEY_MAX) {
                            int rank = 0;
                            for (int j = 0; j < n; ++j) rank += (m.newk[j] < me) ? 1 : 0;
                            m.news[rank] = me;
                        }
                    }
                    __syncthreads();
                    const int keep = valid < cap ? valid : cap;
                    for (int t = tid; t < cap; t += NT) lst[t] = t < keep ? (int32_t)key_index(m.news[t]) : -1;
                    __threadfence();
                    __syncthreads();
                }
            }
            cur = (int)key_index(found[0]);
            __syncthreads();
        }
        total_evals += evals;
        if (tid == 0 && lvl > top) { state[0] = node; state[1] = lvl; __threadfence(); }
        __syncthreads();
    }
    if (tid == 0) { state[2] = (int32_t)(uint32_t)(total_evals & 0xFFFFFFFFull); state[3] = (int32_t)(uint32_t)(total_evals >> 32); }
}

// ------------------------------------------------------------------------------------------
// Batched construction (SURVEY.md 8f rank 1 -- the reference builds node by node at ~50 inserts/s,
// hnsw.py:143-225, which cannot produce the 1M x 768 graph of BASELINE configs[3]).  Levels are
// built top-down; within a level, nodes are inserted in waves: every node of a wave runs the
// reference's search on the graph built so far (greedy descent on the finished upper levels, then
// the ef_construction beam on this level, hnsw.py:174-182), keeps its M closest as links
// (hnsw.py:188) and is then reverse-linked into its neighbours' lists, which are cut to the closest
// `cap` entries (hnsw.py:199-213).  Nodes of one wave do not see each other, so the graph is a
// wave-parallel approximation of the sequential build, NOT the reference's graph.
//
// One CTA per wave node.  out_rows/out_d [count][M]: the M closest found, ascending, -1/inf padded.
__global__ void __launch_bounds__(kHnswThreads)
hnsw_build_search_kernel(Graph g, const int32_t* __restrict__ nodes, int count, int level, int32_t entry_row,
                         int max_level, int32_t seed_entry, int efc, int M, int CC, int NM, int H, uint32_t* global_hash,
                         int32_t* __restrict__ out_rows, float* __restrict__ out_d, int* __restrict__ overflow_flag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BeamMem m;
    carve(m, smem_raw, g.D, CC, efc, NM, H, global_hash ? global_hash + (size_t)blockIdx.x * H : nullptr);
    const int tid = threadIdx.x;
    for (int it = blockIdx.x; it < count; it += gridDim.x) {
        const int node = nodes[it];
        __syncthreads();
        for (int t = tid; t < g.D; t += kHnswThreads) m.qs[t] = g.db[(size_t)node * g.D + t];
        __syncthreads();
        unsigned evals = 0, expansions = 0;
        int cur = entry_row;
        for (int l = max_level; l > level; --l) cur = greedy_level<kHnswThreads>(g, m, cur, l, &evals);
        // `cur` may not be linked on this level yet (it is the node itself when the node also lives
        // on the level above): enter through its closest upper-level neighbour that is, else the seed
        int deg;
        const int32_t* mine = nbr_list(g, cur, level, &deg);
        if (cur == node || deg == 0 || __ldcg(mine) < 0) {
            int pick = seed_entry;
            if (level < max_level) {
                int dup;
                const int32_t* up = nbr_list(g, cur, level + 1, &dup);
                for (int p = 0; p < dup; ++p) {
                    const int r = __ldcg(up + p);
                    if (r < 0) break;
                    int d2;
                    const int32_t* rl = nbr_list(g, r, level, &d2);
                    if (r != node && d2 > 0 && __ldcg(rl) >= 0) { pick = r; break; }
                }
            }
            cur = pick;
        }
        int wf = 0;
        const int nf = beam_search<kHnswThreads>(g, m, cur, level, efc, nullptr, &wf, &evals, &expansions);
        if (m.scal[1] && tid == 0) atomicOr(overflow_flag, 1);
        const u64* found = m.found[wf];
        // the node itself can be reached (it is linked on the upper levels): skip it
        if (tid == 0) {
            int w = 0;
            for (int i = 0; i < nf && w < M; ++i) {
                const int r = (int)key_index(found[i]);
                if (r == node) continue;
                out_rows[(size_t)it * M + w] = r;
                out_d[(size_t)it * M + w] = key_score(found[i]);
                ++w;
            }
            for (; w < M; ++w) { out_rows[(size_t)it * M + w] = -1; out_d[(size_t)it * M + w] = INFINITY; }
        }
    }
}

// Reverse links of one wave.  Segment t = all new (source, dist) pairs whose target list is
// list_idx[t], sorted by distance; one warp merges them with the list's current entries (kept with
// their distances in adjd) and writes back the closest `deg`.
constexpr int kMergeMax = 256;      // keys a warp sorts: deg (<= 128) existing + at most 128 incoming
__global__ void __launch_bounds__(128)
hnsw_reverse_merge_kernel(int32_t* __restrict__ adj, float* __restrict__ adjd, int deg, const int32_t* __restrict__ list_idx,
                          const int32_t* __restrict__ seg_start, int segments, const int32_t* __restrict__ in_rows,
                          const float* __restrict__ in_d) {
    __shared__ u64 keys_s[4][kMergeMax];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = blockIdx.x * 4 + warp;
    if (t >= segments) return;
    u64* keys = keys_s[warp];
    const size_t base = (size_t)list_idx[t] * deg;
    const int s0 = seg_start[t], s1 = seg_start[t + 1];
    const int n_in = min(s1 - s0, kMergeMax - deg);
    for (int i = lane; i < kMergeMax; i += 32) {
        u64 key = VDB_KEY_MAX;
        if (i < deg) {
            const int r = adj[base + i];
            if (r >= 0) key = pack_key(adjd[base + i], (uint32_t)r);
        } else if (i - deg < n_in) {
            key = pack_key(in_d[s0 + i - deg], (uint32_t)in_rows[s0 + i - deg]);
        }
        keys[i] = key;
    }
    __syncwarp();
    for (int k = 2; k <= kMergeMax; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = lane; i < kMergeMax; i += 32) {
                const int p = i ^ j;
                if (p > i) {
                    const u64 a = keys[i], b = keys[p];
                    const bool up = ((i & k) == 0);
                    if ((a > b) == up) { keys[i] = b; keys[p] = a; }
                }
            }
            __syncwarp();
        }
    }
    for (int i = lane; i < deg; i += 32) {
        const u64 key = keys[i];
        // duplicates (same row twice) cannot occur: a wave node is new to every list it is merged into
        adj[base + i] = key == VDB_KEY_MAX ? -1 : (int32_t)key_index(key);
        adjd[base + i] = key == VDB_KEY_MAX ? INFINITY : key_score(key);
    }
}

static int g_hnsw_insert_threads = 256;          // threads of the single insert CTA (debug setter below)
static int g_hnsw_ctas_per_sm = 12;              // resident CTAs per SM the search grid is sized for (debug setter below)
static int g_hnsw_smem_hash_limit = 48 * 1024;   // visited hash stays in shared memory only below this CTA footprint
static int pick_hash_slots(int ef, int deg) {
    long long want = ((long long)(ef + 32) * deg * 5) / 4;
    int h = 1024;
    while (h < want && h < (1 << 22)) h <<= 1;
    return h;
}

}  // namespace vdb

using namespace vdb;

extern "C" {

void vdb_debug_set_hnsw_smem_hash_limit(int bytes) { g_hnsw_smem_hash_limit = bytes; }
void vdb_debug_set_hnsw_insert_threads(int n) { g_hnsw_insert_threads = (n == 256 || n == 512) ? n : (n == 128 ? 128 : 256); }
void vdb_debug_set_hnsw_ctas_per_sm(int n) { g_hnsw_ctas_per_sm = n < 1 ? 1 : (n > 16 ? 16 : n); }

size_t vdb_hnsw_search_workspace(int Q, int ef, int deg0) {
    (void)Q;
    size_t h = (size_t)pick_hash_slots(ef, deg0) * 4;   // forced-retry variant may ask for 4x
    return 256 + (size_t)sm_count() * 32 * h * 4;          // up to 32 traversals (warps or CTAs) resident per SM
}

int vdb_hnsw_search(const float* q, int Q, const float* db, int64_t N, int D, const int32_t* adj0, int deg0,
                    const int32_t* adj_up, int deg_up, const int32_t* up_ptr, int32_t entry_row, int max_level, int ef,
                    int k, const uint32_t* alive, const uint32_t* filter, const int32_t* levels, int hash_slots,
                    float* out_d, int32_t* out_rows, uint32_t* out_stats, int32_t* out_status, void* ws, size_t ws_bytes,
                    void* stream) {
    VDB_REQUIRE(Q >= 0 && D > 0 && ef > 0 && k > 0 && deg0 > 0, VDB_ERR_INVALID, "vdb_hnsw_search: bad shape");
    VDB_REQUIRE(ef <= 1024, VDB_ERR_UNSUPPORTED, "ef=%d > 1024", ef);
    VDB_REQUIRE(deg0 <= 128 && deg_up <= 128, VDB_ERR_UNSUPPORTED, "degree > 128");
    VDB_REQUIRE(N < (1ll << 31), VDB_ERR_UNSUPPORTED, "N too large for int32 rows");
    VDB_REQUIRE(hash_slots == 0 || (hash_slots & (hash_slots - 1)) == 0, VDB_ERR_INVALID, "hash_slots must be a power of two");
    if (Q == 0) return VDB_OK;
    VDB_REQUIRE(q && out_d && out_rows && ws && ws_bytes >= 256, VDB_ERR_INVALID, "vdb_hnsw_search: null pointer");
    VDB_REQUIRE(entry_row < 0 || (db && adj0), VDB_ERR_INVALID, "vdb_hnsw_search: null graph");
    VDB_REQUIRE(max_level <= 0 || (adj_up && up_ptr), VDB_ERR_INVALID, "vdb_hnsw_search: upper levels missing");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Graph g{db, N, D, adj0, deg0, adj_up, deg_up, up_ptr, alive, levels};
    const int NM = (deg0 > deg_up ? deg0 : deg_up) + 1;
    // filtered: the reference's heap is unbounded; 4 ef (>= 512) entries cover a filter that passes ~1 row in 8
    // before the beam fills, anything beyond is flagged (out_status bit 1) and finished by the exact masked scan
    const int CC = filter ? (4 * ef > 512 ? 4 * ef : 512) : ef;
    int H = hash_slots ? hash_slots : pick_hash_slots(ef, deg0);
    size_t smem_with = beam_smem_bytes(D, CC, ef, NM, H);
    size_t smem_without = beam_smem_bytes(D, CC, ef, NM, 0);
    int* flag = static_cast<int*>(ws);
    // a big shared-memory hash caps the CTAs per SM, and the traversal needs occupancy to hide gather latency
    const bool use_global = hash_slots != 0 || smem_with > (size_t)g_hnsw_smem_hash_limit;
    size_t smem = use_global ? smem_without : smem_with;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "D=%d ef=%d exceed shared memory", D, ef);
    VDB_CUDA(cudaFuncSetAttribute(hnsw_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    VDB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hnsw_search_kernel, kHnswThreads, smem));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > g_hnsw_ctas_per_sm) per_sm = g_hnsw_ctas_per_sm;
    int grid = sm_count() * per_sm;
    if (grid > Q) grid = Q;
    uint32_t* ghash = nullptr;
    if (use_global) {
        size_t need = 256 + (size_t)grid * H * 4;
        VDB_REQUIRE(ws_bytes >= need, VDB_ERR_WORKSPACE, "workspace %zu < %zu", ws_bytes, need);
        ghash = reinterpret_cast<uint32_t*>(static_cast<char*>(ws) + 256);
    }
    VDB_CUDA(cudaMemsetAsync(flag, 0, 4, st));
    ProfScope prof(kProfHnswSearch, st);
    hnsw_search_kernel<<<grid, kHnswThreads, smem, st>>>(g, q, Q, entry_row, max_level, ef, k, filter, CC, NM, H, ghash,
                                                         out_d, out_rows, out_stats, out_status, flag);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

size_t vdb_hnsw_build_search_workspace(int count, int efc, int deg0) {
    (void)count;
    return 256 + (size_t)sm_count() * 8 * (size_t)pick_hash_slots(efc, deg0) * 4;
}

int vdb_hnsw_build_search(const float* db, int64_t N, int D, const int32_t* adj0, int deg0, const int32_t* adj_up,
                          int deg_up, const int32_t* up_ptr, const int32_t* nodes, int count, int level,
                          int32_t entry_row, int max_level, int32_t seed_entry, int efc, int M, int32_t* out_rows,
                          float* out_d, void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(D > 0 && M > 0 && efc >= M && count >= 0 && level >= 0 && level <= max_level, VDB_ERR_INVALID,
                "vdb_hnsw_build_search: bad arguments");
    VDB_REQUIRE(efc <= 1024 && deg0 <= 128 && deg_up <= 128, VDB_ERR_UNSUPPORTED, "efc > 1024 or degree > 128");
    VDB_REQUIRE(N < (1ll << 31), VDB_ERR_UNSUPPORTED, "N too large for int32 rows");
    if (count == 0) return VDB_OK;
    VDB_REQUIRE(db && adj0 && adj_up && up_ptr && nodes && out_rows && out_d && ws && ws_bytes >= 256, VDB_ERR_INVALID,
                "vdb_hnsw_build_search: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Graph g{db, N, D, adj0, deg0, adj_up, deg_up, up_ptr, nullptr, nullptr};
    const int NM = (deg0 > deg_up ? deg0 : deg_up) + 1;
    const int CC = efc;
    const int H = pick_hash_slots(efc, level == 0 ? deg0 : deg_up);
    size_t smem_with = beam_smem_bytes(D, CC, efc, NM, H);
    size_t smem_without = beam_smem_bytes(D, CC, efc, NM, 0);
    const bool use_global = smem_with > (size_t)110 * 1024;
    size_t smem = use_global ? smem_without : smem_with;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "D=%d efc=%d exceed shared memory", D, efc);
    VDB_CUDA(cudaFuncSetAttribute(hnsw_build_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    VDB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hnsw_build_search_kernel, kHnswThreads, smem));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 8) per_sm = 8;
    int grid = sm_count() * per_sm;
    if (grid > count) grid = count;
    int* flag = static_cast<int*>(ws);
    uint32_t* ghash = nullptr;
    if (use_global) {
        VDB_REQUIRE(ws_bytes >= 256 + (size_t)grid * H * 4, VDB_ERR_WORKSPACE, "workspace too small");
        ghash = reinterpret_cast<uint32_t*>(static_cast<char*>(ws) + 256);
    }
    // the overflow flag is STICKY over the waves of a build: the caller zeroes it once and reads it once
    ProfScope prof(kProfHnswInsert, st);
    hnsw_build_search_kernel<<<grid, kHnswThreads, smem, st>>>(g, nodes, count, level, entry_row, max_level, seed_entry, efc,
                                                              M, CC, NM, H, ghash, out_rows, out_d, flag);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

int vdb_hnsw_reverse_merge(int32_t* adj, float* adjd, int deg, const int32_t* list_idx, const int32_t* seg_start,
                           int segments, const int32_t* in_rows, const float* in_d, void* stream) {
    VDB_REQUIRE(deg > 0 && deg <= 128 && segments >= 0, VDB_ERR_INVALID, "vdb_hnsw_reverse_merge: bad arguments");
    if (segments == 0) return VDB_OK;
    VDB_REQUIRE(adj && adjd && list_idx && seg_start && in_rows && in_d, VDB_ERR_INVALID, "vdb_hnsw_reverse_merge: null pointer");
    hnsw_reverse_merge_kernel<<<(segments + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(
        adj, adjd, deg, list_idx, seg_start, segments, in_rows, in_d);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

size_t vdb_hnsw_insert_workspace(int efc, int deg0) { return 256 + (size_t)pick_hash_slots(efc, deg0) * 4 * 4; }

int vdb_hnsw_insert(const float* db, int64_t N, int D, int32_t* adj0, int deg0, int32_t* adj_up, int deg_up,
                    const int32_t* up_ptr, const int32_t* levels, int32_t first, int32_t count, int M, int efc,
                    int32_t* state, const uint32_t* alive, void* ws, size_t ws_bytes, void* stream) {
    VDB_REQUIRE(D > 0 && M > 0 && efc > 0 && count >= 0 && first >= 0, VDB_ERR_INVALID, "vdb_hnsw_insert: bad shape");
    VDB_REQUIRE(deg0 >= M && deg_up >= M, VDB_ERR_INVALID, "adjacency narrower than M");
    VDB_REQUIRE(deg0 <= 127 && deg_up <= 127, VDB_ERR_UNSUPPORTED, "degree > 127");
    VDB_REQUIRE(efc <= 1024, VDB_ERR_UNSUPPORTED, "ef_construction=%d > 1024", efc);
    VDB_REQUIRE((int64_t)first + count <= N && N < (1ll << 31), VDB_ERR_INVALID, "rows out of range");
    if (count == 0) return VDB_OK;
    VDB_REQUIRE(db && adj0 && adj_up && up_ptr && levels && state && ws && ws_bytes >= 256, VDB_ERR_INVALID,
                "vdb_hnsw_insert: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Graph g{db, N, D, adj0, deg0, adj_up, deg_up, up_ptr, alive, levels};
    const int NM = (deg0 > deg_up ? deg0 : deg_up) + 1;
    const int CC = efc;
    const int H = pick_hash_slots(efc, deg0);
    const size_t qs2 = (size_t)((D + 3) & ~3) * 4;
    size_t smem_with = beam_smem_bytes(D, CC, efc, NM, H) + qs2;
    size_t smem_without = beam_smem_bytes(D, CC, efc, NM, 0) + qs2;
    const bool use_global = smem_with > (size_t)max_smem_optin() - 1024;
    size_t smem = use_global ? smem_without : smem_with;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin() - 1024, VDB_ERR_UNSUPPORTED, "D=%d efc=%d exceed shared memory", D, efc);
    int* flag = static_cast<int*>(ws);
    uint32_t* ghash = nullptr;
    if (use_global) {
        VDB_REQUIRE(ws_bytes >= 256 + (size_t)H * 4, VDB_ERR_WORKSPACE, "workspace too small");
        ghash = reinterpret_cast<uint32_t*>(static_cast<char*>(ws) + 256);
    }
    const int nt = g_hnsw_insert_threads;
    // the overflow flag (first word of ws) is STICKY over the chunks of a build: the caller zeroes it once before
    // the first chunk and reads it once after the last (a per-call memset erased an overflow of an earlier chunk)
    ProfScope prof(kProfHnswInsert, st);
    if (nt == 512) {
        VDB_CUDA(cudaFuncSetAttribute(hnsw_insert_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        hnsw_insert_kernel<512><<<1, 512, smem, st>>>(g, levels, first, count, M, efc, state, CC, NM, H, ghash, flag);
    } else if (nt == 256) {
        VDB_CUDA(cudaFuncSetAttribute(hnsw_insert_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        hnsw_insert_kernel<256><<<1, 256, smem, st>>>(g, levels, first, count, M, efc, state, CC, NM, H, ghash, flag);
    } else {
        VDB_CUDA(cudaFuncSetAttribute(hnsw_insert_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        hnsw_insert_kernel<128><<<1, 128, smem, st>>>(g, levels, first, count, M, efc, state, CC, NM, H, ghash, flag);
    }
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"
