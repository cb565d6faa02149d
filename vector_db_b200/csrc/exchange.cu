// Shard exchange + merge in ONE kernel over NVLink peer memory.
//
// Every rank answers the same query batch over its own rows; the per-shard top-k lists then have to meet
// and be merged.  The baseline (vector_db_b200/sharded.py: gather_topk + merge_shards) packs the lists,
// calls one NCCL all-gather, unpacks and launches vdb_merge_topk: seven small launches and a collective
// for 120-500 KB of payload -- all latency.  Here each rank owns one peer-mapped exchange buffer (allocated
// once with torch's symmetric memory, or plain device memory when the "ranks" share a GPU in the tests):
//
//   header   uint32 arrived[2][kXMaxWorld]      arrived[p][r] = CTAs of rank r that finished pushing (parity p)
//   slot p   f32 d[world][cap] | i64 id[world][cap]          p = call number & 1
//
//   phase 1  each CTA stores its share of the local lists into EVERY rank's slot (plain stores to peer
//            memory: NVLink writes, 12 B x Q x k per peer), fences at system scope and adds 1 to
//            arrived[p][my rank] on every rank (red.release.sys)
//   phase 2  the CTA waits until arrived[p][r] of its OWN buffer reached the call's target for every r
//            (ld.acquire.sys spin, bounded by a clock limit that raises a status word instead of hanging)
//            and merges its share of the queries: k best by (value, shard, position) -- the order of
//            vdb_merge_topk, i.e. (value, id) when shard r owns smaller ids than shard r+1
//
// No CTA waits before it has pushed and the grid is at most one CTA per SM (all co-resident), so the wait
// cannot deadlock.  Slot reuse is safe with two parities: a rank can start call c+2 only after its call c+1
// finished, which needed every peer's pushes of c+1, which those peers issue after finishing call c.
// The arrival counters are never reset; the host tracks the running target (same on every rank because
// the query batch -- hence the grid -- is replicated).
#include "common.cuh"

namespace vdb {

constexpr int kXThreads = 256;
constexpr int kXWarps = kXThreads / 32;
constexpr int kXMaxWorld = 16;
constexpr int kXHeader = 256;                            // bytes: 2 * 16 counters, padded
constexpr long long kXSpinLimit = 6000000000ll;          // ~3 s of SM clocks

struct XchgArgs {
    const float* d;
    const int64_t* i;
    int Q, k, descending, rank, world, parity;
    uint32_t target;
    long long cap;                                       // elements per (rank) list plane
    const unsigned long long* peer;                      // [world] base addresses of the ranks' exchange buffers
    float* out_d;
    int64_t* out_i;
    int32_t* status;                                     // set to 1 when a wait ran into the clock limit
};

__host__ __device__ inline size_t xchg_slot_bytes(int world, long long cap) {
    return (((size_t)world * (size_t)cap * 12) + 255) & ~(size_t)255;
}

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_sys_add(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(kXThreads)
exchange_merge_kernel(const XchgArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* keys = reinterpret_cast<u64*>(smem_raw);        // [kXWarps][world * k]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (a.Q + gridDim.x - 1) / gridDim.x;
    const int q0 = min(a.Q, (int)blockIdx.x * per), q1 = min(a.Q, q0 + per);
    const long long e0 = (long long)q0 * a.k, e1 = (long long)q1 * a.k;
    const size_t slot = kXHeader + (size_t)a.parity * xchg_slot_bytes(a.world, a.cap);

    // ---- phase 1: my lists -> every rank's slot (own rank first, then the peers in ring order)
    for (int s = 0; s < a.world; ++s) {
        const int peer = (a.rank + s) % a.world;
        unsigned char* base = reinterpret_cast<unsigned char*>(a.peer[peer]) + slot;
        float* pd = reinterpret_cast<float*>(base) + (size_t)a.rank * a.cap;
        int64_t* pi = reinterpret_cast<int64_t*>(base + (size_t)a.world * a.cap * 4) + (size_t)a.rank * a.cap;
        for (long long e = e0 + tid; e < e1; e += kXThreads) {
            pd[e] = a.d[e];
            pi[e] = a.i[e];
        }
    }
    __threadfence_system();
    __syncthreads();
    if (tid < a.world)
        red_release_sys_add(reinterpret_cast<uint32_t*>(a.peer[tid]) + a.parity * kXMaxWorld + a.rank, 1u);

    // ---- phase 2: wait for every rank's CTAs, then merge my share of the queries
    int gave_up = 0;
    if (tid < a.world) {
        const uint32_t* f = reinterpret_cast<const uint32_t*>(a.peer[a.rank]) + a.parity * kXMaxWorld + tid;
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(f) - a.target) < 0) {
            if (clock64() - t0 > kXSpinLimit) { atomicExch(a.status, 1); gave_up = 1; break; }
            __nanosleep(100);
        }
    }
    if (__syncthreads_or(gave_up)) {
        // a peer never arrived: its slot holds a previous call's lists (or nothing).  Merging them would return a
        // plausible, WRONG answer -- return (inf, -1) for this CTA's queries; status tells the host (PeerExchange.check)
        for (long long e = e0 + tid; e < e1; e += kXThreads) {
            a.out_d[e] = a.descending ? -INFINITY : INFINITY;
            a.out_i[e] = -1;
        }
        return;
    }

    const int total = a.world * a.k;
    const unsigned char* mine = reinterpret_cast<const unsigned char*>(a.peer[a.rank]) + slot;
    const float* gd = reinterpret_cast<const float*>(mine);
    const int64_t* gi = reinterpret_cast<const int64_t*>(mine + (size_t)a.world * a.cap * 4);
    u64* wk = keys + (size_t)warp * total;
    for (int q = q0 + warp; q < q1; q += kXWarps) {
        // keys (value, shard * k + position); rows written by peers are read past L1 (ld.global.cg)
        for (int t = lane; t < total; t += 32) {
            const int l = t / a.k, p = t - l * a.k;
            const size_t src = (size_t)l * a.cap + (size_t)q * a.k + p;
            const int64_t id = __ldcg(gi + src);
            const float v = __ldcg(gd + src);
            wk[t] = id >= 0 ? pack_key(a.descending ? -v : v, (uint32_t)t) : VDB_KEY_MAX;
        }
        __syncwarp();
        for (int r = 0; r < a.k; ++r) {
            u64 best = VDB_KEY_MAX;
            for (int t = lane; t < total; t += 32) { const u64 v = wk[t]; best = v < best ? v : best; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const u64 other = __shfl_xor_sync(0xffffffffu, best, o);
                best = other < best ? other : best;
            }
            if (lane == 0) {
                const size_t o = (size_t)q * a.k + r;
                if (best == VDB_KEY_MAX) {
                    a.out_d[o] = a.descending ? -INFINITY : INFINITY;
                    a.out_i[o] = -1;
                } else {
                    const int t = (int)key_index(best);
                    const int l = t / a.k, p = t - l * a.k;
                    const size_t src = (size_t)l * a.cap + (size_t)q * a.k + p;
                    a.out_d[o] = __ldcg(gd + src);
                    a.out_i[o] = __ldcg(gi + src);
                    wk[t] = VDB_KEY_MAX;                       // taken
                }
            }
            __syncwarp();
        }
    }
}

static int xchg_grid(int Q) {
    const int want = (Q + kXWarps - 1) / kXWarps;
    const int sms = sm_count();
    return want < 1 ? 1 : (want < sms ? want : sms);
}

}  // namespace vdb

using namespace vdb;

extern "C" {

size_t vdb_exchange_bytes(int world, int64_t cap) {
    if (world <= 0 || world > kXMaxWorld || cap <= 0) return 0;
    return kXHeader + 2 * xchg_slot_bytes(world, cap);
}

int vdb_exchange_grid(int Q) { return Q > 0 ? xchg_grid(Q) : 0; }

int vdb_exchange_merge_topk(const float* d, const int64_t* i, int Q, int k, int descending, const uint64_t* peer_bufs,
                            int rank, int world, int64_t cap, int parity, uint32_t target, float* out_d, int64_t* out_i,
                            int32_t* status, void* stream) {
    VDB_REQUIRE(world > 0 && world <= kXMaxWorld && rank >= 0 && rank < world, VDB_ERR_INVALID,
                "vdb_exchange_merge_topk: bad rank %d / world %d (at most %d ranks)", rank, world, kXMaxWorld);
    VDB_REQUIRE(Q > 0 && k > 0 && (int64_t)Q * k <= cap, VDB_ERR_INVALID,
                "vdb_exchange_merge_topk: Q*k=%lld exceeds the exchange capacity %lld", (long long)Q * k, (long long)cap);
    VDB_REQUIRE(parity == 0 || parity == 1, VDB_ERR_INVALID, "parity must be 0 or 1");
    VDB_REQUIRE(d && i && peer_bufs && out_d && out_i && status, VDB_ERR_INVALID, "vdb_exchange_merge_topk: null pointer");
    const size_t smem = (size_t)kXWarps * world * k * 8;
    VDB_REQUIRE(smem <= (size_t)max_smem_optin(), VDB_ERR_UNSUPPORTED, "world*k=%d too large", world * k);
    XchgArgs a;
    a.d = d; a.i = i; a.Q = Q; a.k = k; a.descending = descending; a.rank = rank; a.world = world; a.parity = parity;
    a.target = target; a.cap = cap; a.peer = reinterpret_cast<const unsigned long long*>(peer_bufs);
    a.out_d = out_d; a.out_i = out_i; a.status = status;
    VDB_CUDA(cudaFuncSetAttribute(exchange_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    exchange_merge_kernel<<<xchg_grid(Q), kXThreads, smem, static_cast<cudaStream_t>(stream)>>>(a);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"
