// Shared device/host helpers for the vector_db_b200 search core (sm_100a only).
//
//  * error plumbing for the C-ABI (thread-local last-error string, VDB_* codes)
//  * order-preserving (score,index) -> u64 key packing: ascending key == ascending (score, idx),
//    the (dist, id) tuple order the reference sorts by (tests/indexing/test_hnsw.py:128-134,
//    src/vector_db/indexing/hnsw.py:366)
//  * BlockSelect: CTA-wide streaming top-R in shared memory (threshold filter + pending area +
//    bitonic compaction) used by the scan kernels and by the merge kernels; BlockSelectRadix: the same protocol
//    with a radix select instead of sorts and an optional barrier-free flag, for large shortlists (IVF-PQ)
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/vdb_b200.h"

namespace vdb {

// ------------------------------------------------------------------ errors
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define VDB_CUDA(expr)                                                          \
    do {                                                                        \
        cudaError_t _e = (expr);                                                \
        if (_e != cudaSuccess) return ::vdb::cuda_fail(_e, #expr, __FILE__, __LINE__); \
    } while (0)

#define VDB_REQUIRE(cond, code, ...)                                            \
    do {                                                                        \
        if (!(cond)) { ::vdb::set_error(__VA_ARGS__); return (code); }          \
    } while (0)

int sm_count();
int max_smem_optin();
void count_launch(int n);

// Optional per-kernel timing (bench.py): when enabled, a ProfScope brackets ONE kernel launch with a
// pair of CUDA events on the launching stream; vdb_profile_read() sums and clears a slot.
enum ProfSlot { kProfExactF32 = 0, kProfExactTC = 1, kProfAdcScan = 2, kProfPqEncode = 3, kProfHnswSearch = 4,
                kProfRerank = 5, kProfFinalize = 6, kProfHnswInsert = 7, kProfIvfScan = 8, kProfSlots = 9 };
struct ProfScope {
    int slot; cudaStream_t st; bool on;
    ProfScope(int slot_, cudaStream_t st_);
    ~ProfScope();
};

// ------------------------------------------------------------------ keys
typedef unsigned long long u64;
#define VDB_KEY_MAX 0xFFFFFFFFFFFFFFFFull

// float -> uint32 such that unsigned order == float order (-inf < ... < -0 < +0 < ... < +inf < NaN)
__host__ __device__ __forceinline__ uint32_t f32_to_ordered(float f) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4);
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ordered_to_f32(uint32_t u) {
    u = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
__host__ __device__ __forceinline__ u64 pack_key(float score, uint32_t idx) {
    return ((u64)f32_to_ordered(score) << 32) | (u64)idx;
}
__host__ __device__ __forceinline__ float key_score(u64 k) { return ordered_to_f32((uint32_t)(k >> 32)); }
__host__ __device__ __forceinline__ uint32_t key_index(u64 k) { return (uint32_t)(k & 0xFFFFFFFFull); }

__host__ __device__ __forceinline__ int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

#ifdef __CUDACC__
// ------------------------------------------------------------------ warp helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// 128-bit streaming load that does not allocate in L1 (read-once code / vector streams)
__device__ __forceinline__ uint4 ld_stream_u4(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// The same load for READ-ONLY data as a plain (non-volatile) asm: the compiler may hoist a group of them ahead of the
// arithmetic that consumes them (gather loops want every load of a row group in flight before the first use; the
// volatile form pins each load between the arithmetic statements around it).
__device__ __forceinline__ float4 ld_gather_f4(const float4* p) {
    float4 r;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
        : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// ------------------------------------------------------------------ CTA-wide bitonic sort (ascending), n = power of two
// `lists` independent arrays of n keys each, stored back to back, are sorted by the same network in one
// pass (one barrier per stage for all of them); every thread handles compare-exchange PAIRS, none idle.
__device__ __forceinline__ void block_bitonic_sort_lists(u64* buf, int n, int lists) {
    const int half = n >> 1;
    const int hs = 31 - __clz(half | 1);             // log2(half): n is a power of two (a division here was a third of the
    const int pairs = half * lists;                  // ADC sample pass: 45 stages x 3 pairs per thread, 8 sorts per unit)
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < pairs; t += blockDim.x) {
                const int l = t >> hs, tl = t & (half - 1);
                const int i = ((tl & ~(j - 1)) << 1) | (tl & (j - 1));
                u64* b = buf + (size_t)l * n;
                const u64 x = b[i], y = b[i | j];
                const bool up = ((i & k) == 0);
                if ((x > y) == up) { b[i] = y; b[i | j] = x; }
            }
            __syncthreads();
        }
    }
}
__device__ __forceinline__ void block_bitonic_sort(u64* buf, int n) { block_bitonic_sort_lists(buf, n, 1); }

// ------------------------------------------------------------------ BlockSelect
// Streaming "keep the R smallest keys" over a shared-memory array of C = 2^c entries.
//   [0, R)       best-so-far (sorted after every compaction)
//   [R, C)       pending area; threads append with an atomic counter
// Protocol: between two CTA-uniform calls to maybe_compact() the CTA may push at most `burst`
// keys; maybe_compact() sorts when fewer than `burst` pending slots remain, so push() never
// overflows.  Keys are unique (score,index) pairs, so `key < thr` is an exact filter.
struct BlockSelect {
    u64* buf;   // shared, C entries
    int* cnt;   // shared, pending count
    u64* thr;   // shared, current R-th best key (VDB_KEY_MAX until R keys were seen)
    int C, R, burst;

    __device__ __forceinline__ void init(u64* b, int* c, u64* t, int C_, int R_, int burst_) {
        buf = b; cnt = c; thr = t; C = C_; R = R_; burst = burst_;
        for (int i = threadIdx.x; i < C; i += blockDim.x) buf[i] = VDB_KEY_MAX;
        if (threadIdx.x == 0) { *cnt = 0; *thr = VDB_KEY_MAX; }
        // caller must __syncthreads()
    }
    __device__ __forceinline__ u64 threshold() const { return *thr; }
    __device__ __forceinline__ void push(u64 key) {
        if (key < *thr) {
            int p = atomicAdd(cnt, 1);
            buf[R + p] = key;
        }
    }
    // CTA-uniform, called by ALL threads after they issued their pushes of this burst.  The
    // barrier-with-OR makes the decision uniform: the thread that pushed last sees the full count.
    __device__ __forceinline__ void maybe_compact() {
        int pending = *(volatile int*)cnt;
        if (__syncthreads_or(pending + burst > C - R)) compact_after_barrier();
    }
    // requires: a barrier since the last push
    __device__ __forceinline__ void compact_after_barrier() {
        block_bitonic_sort(buf, C);
        if (threadIdx.x == 0) { *cnt = 0; *thr = buf[R - 1]; }
        __syncthreads();
    }
    __device__ __forceinline__ void finalize() {   // leaves buf[0..R) sorted ascending
        __syncthreads();
        compact_after_barrier();
    }
};

// ------------------------------------------------------------------ BlockSelectRadix
// The same streaming "keep the R smallest keys" protocol WITHOUT sorting, for large R next to few candidates
// (IVF-PQ shortlists: R = 1000 of ~16 k candidates per CTA, where the bitonic sorts of BlockSelect were 3/4 of
// the kernel's shared-memory traffic).  A compaction finds the R-th smallest key exactly with an 8-pass radix
// select over the 64-bit keys (one warp-aggregated shared histogram per byte) and then moves the keys <= it to
// the front with a stable in-place compaction.  buf[0, kept) holds the survivors in NO particular order.
// Requirements: keys unique and < VDB_KEY_MAX; blockDim.x a multiple of 32, >= 256 and <= 1024; `scratch` =
// kSelectRadixScratch ints of shared memory.
constexpr int kSelectRadixScratch = 256 + 64 + 2 + 1;

struct BlockSelectRadix {
    u64* buf;
    int* cnt;
    u64* thr;
    int* hist;    // 256 bins
    int* wsum;    // 2 x 32 warp totals
    int* sel;     // chosen bin, rank inside it
    int* flag;    // barrier-free protocol: raised by a pusher when the pending area is within one burst of full
    int C, R, burst;
    int kept;     // CTA-uniform: keys at the front of buf (0 until the first compaction, then R)

    __device__ __forceinline__ void init(u64* b, int* c, u64* t, int* scratch, int C_, int R_, int burst_) {
        buf = b; cnt = c; thr = t; hist = scratch; wsum = scratch + 256; sel = scratch + 320;
        flag = scratch + 322;
        C = C_; R = R_; burst = burst_; kept = 0;
        if (threadIdx.x == 0) { *cnt = 0; *thr = VDB_KEY_MAX; *flag = 0; }
        // caller must __syncthreads()
    }
    __device__ __forceinline__ void push(u64 key) {
        if (key < *thr) {
            const int p = atomicAdd(cnt, 1);
            buf[kept + p] = key;
        }
    }
    // Barrier-free variant (burst = blockDim.x, C >= R + 2 * burst): every thread polls flag_raised() before EACH
    // push_flagging() and, once it sees the flag, pushes nothing more until the CTA has met at a barrier and run
    // compact_after_barrier().  A thread whose slot lies at or past the trigger raises the flag itself, so it
    // pushes at most once after the trigger slot was taken: at most burst - 1 slots follow the trigger and the
    // buffer cannot overflow.  The threads need not be at the same batch when they meet.
    __device__ __forceinline__ bool flag_raised() const { return *(volatile int*)flag != 0; }
    __device__ __forceinline__ void push_flagging(u64 key) {
        if (key < *(volatile u64*)thr) {
            const int p = atomicAdd(cnt, 1);
            buf[kept + p] = key;
            if (kept + p + 1 + burst >= C) *(volatile int*)flag = 1;
        }
    }
    __device__ __forceinline__ void maybe_compact() {
        const int pending = *(volatile int*)cnt;
        if (__syncthreads_or(kept + pending + burst > C)) compact_after_barrier();
    }
    // the key with exactly `rank` keys <= it among buf[0, n); CTA-uniform result
    __device__ __forceinline__ u64 radix_select(int n, int rank) {
        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        u64 prefix = 0;
        for (int shift = 56; shift >= 0; shift -= 8) {
            for (int b = tid; b < 256; b += blockDim.x) hist[b] = 0;
            __syncthreads();
            const u64 himask = shift == 56 ? 0ull : (~0ull << (shift + 8));
            for (int base = 0; base < n; base += blockDim.x) {      // uniform trip count: every lane reaches the match
                const int i = base + tid;
                int digit = -1;
                if (i < n) {
                    const u64 key = buf[i];
                    if ((key & himask) == prefix) digit = (int)((key >> shift) & 255u);
                }
                const unsigned peers = __match_any_sync(0xFFFFFFFFu, digit);
                if (digit >= 0 && lane == __ffs(peers) - 1) atomicAdd(&hist[digit], __popc(peers));
            }
            __syncthreads();
            int v = 0, incl = 0;
            if (tid < 256) {
                v = hist[tid];
                incl = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                    if (lane >= o) incl += t;
                }
                if (lane == 31) wsum[warp] = incl;
            }
            __syncthreads();
            if (tid < 256) {
                int off = 0;
                for (int w = 0; w < warp; ++w) off += wsum[w];
                incl += off;
                if (incl - v < rank && rank <= incl) { sel[0] = tid; sel[1] = rank - (incl - v); }
            }
            __syncthreads();
            prefix |= (u64)(unsigned)sel[0] << shift;
            rank = sel[1];
        }
        return prefix;
    }
    // requires: a barrier since the last push
    __device__ __forceinline__ void compact_after_barrier() {
        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
        const int n = kept + *(volatile int*)cnt;
        if (n <= R) {                                  // nothing to drop (only finalize() gets here)
            __syncthreads();
            kept = n;
            if (tid == 0) { *cnt = 0; *flag = 0; }
            __syncthreads();
            return;
        }
        const u64 kth = radix_select(n, R);
        // stable compaction, one round of blockDim keys per barrier: a round writes only below its own end, and
        // its own keys are in registers by the barrier
        int out = 0, round = 0;
        for (int base = 0; base < n; base += blockDim.x, ++round) {
            const int i = base + tid;
            const u64 key = i < n ? buf[i] : VDB_KEY_MAX;
            const bool keep = key <= kth;
            const unsigned bal = __ballot_sync(0xFFFFFFFFu, keep);
            int* ws = wsum + (round & 1) * 32;
            if (lane == 0) ws[warp] = __popc(bal);
            __syncthreads();
            int off = out, total = 0;
            for (int w = 0; w < nw; ++w) { const int c = ws[w]; total += c; if (w < warp) off += c; }
            if (keep) buf[off + __popc(bal & ((1u << lane) - 1u))] = key;
            out += total;
        }
        kept = R;                                      // unique keys: exactly R survive
        if (tid == 0) { *cnt = 0; *thr = kth; *flag = 0; }
        __syncthreads();
    }
    // leaves the min(R, seen) best keys in buf[0, kept), unordered, and VDB_KEY_MAX in buf[kept, fill_to)
    __device__ __forceinline__ void finalize(int fill_to) {
        __syncthreads();
        compact_after_barrier();
        for (int i = kept + threadIdx.x; i < fill_to; i += blockDim.x) buf[i] = VDB_KEY_MAX;
        __syncthreads();
    }
};

__host__ __device__ __forceinline__ int select_capacity(int R, int burst) {
    int c = next_pow2(R + 2 * burst);
    return c < 256 ? 256 : c;
}

// (exact_f32.cu) select the k best of L keys per query and translate them to (value, id)
int finalize_keys(const u64* keys, int Q, int L, int k, const int64_t* ids, int64_t idx_offset, int negate,
                  float* out_d, int64_t* out_i, cudaStream_t st, const int32_t* qsel = nullptr,
                  const int32_t* qcount = nullptr, int sub_tq = 0, int sub_grid = 0, int64_t sub_N = 0);
// (exact_f32.cu) exact fp32 scan of a device-side list of queries (the tensor-core scan's uncertified ones)
size_t exact_f32_subset_workspace(int Q, int64_t N, int D, int k);
int exact_f32_subset(const float* q, int Q, const float* db, const int64_t* ids, int64_t N, int D, int k, int metric,
                     const int32_t* qsel, const int32_t* qcount, float* out_d, int64_t* out_i, void* ws, size_t ws_bytes,
                     cudaStream_t st);
// (exact_f32.cu) vdb_rerank_f32 with an optional per-query count of meaningful candidates (front-packed)
int rerank_f32_counted(const float* q, int Q, const float* db, const int64_t* ids, int64_t N, int D, const int32_t* cand,
                       const int32_t* cand_cnt, int R, int k, int metric, float* out_d, int64_t* out_i, void* stream);
#endif  // __CUDACC__

}  // namespace vdb
