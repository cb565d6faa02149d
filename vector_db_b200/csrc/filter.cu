// Metadata pre-filter on the GPU.  The reference answers a metadata filter with a Python
// scan that JSON-decodes every node (src/vector_db/services/storage_service.py:106-128) and hands the resulting
// id set to HNSW._search_layer (src/vector_db/indexing/hnsw.py:96,127).  Here a metadata key is a column of
// dictionary codes, int32 [rows] in HBM (code 0 = key missing or value None, which `metadata.get(k) != v`
// cannot tell apart either), and a filter {k: v, ...} is one pass per key that ANDs `column == code(v)` into the
// row bitmap every search kernel already consumes (exact scan, ADC scan, IVF list scan, HNSW traversal).
// HBM-bound integer work: 4 bytes read per row and key, 1 bit written (ballot per 32 rows, coalesced loads).
#include "common.cuh"

namespace vdb {

// bit r of bitmap: (first ? 1 : old bit) && column[r] == want && (alive == nullptr || alive bit r)
__global__ void filter_eq_kernel(const int32_t* __restrict__ column, int64_t n, int32_t want, const uint32_t* __restrict__ alive,
                                 uint32_t* __restrict__ bitmap, int first) {
    const int64_t words = (n + 31) >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    // a warp takes 32 consecutive words = 1024 rows per trip: lane l loads row 32 w + l of each word w (coalesced),
    // the ballot of word w is the word; lane w keeps it and the 32 lanes store 32 consecutive words
    for (int64_t base = warp0 * 32; base < words; base += nwarps * 32) {
        uint32_t mine = 0u;
#pragma unroll 4
        for (int w = 0; w < 32; ++w) {
            const int64_t r = ((base + w) << 5) + lane;
            const bool hit = (base + w < words) && r < n && __ldg(column + r) == want;
            const uint32_t b = __ballot_sync(0xFFFFFFFFu, hit);
            if (lane == w) mine = b;
        }
        const int64_t word = base + lane;
        if (word < words) {
            if (alive) mine &= alive[word];
            if (!first) mine &= bitmap[word];
            bitmap[word] = mine;
        }
    }
}

}  // namespace vdb

using namespace vdb;

extern "C" int vdb_filter_eq(const int32_t* column, int64_t n, int32_t want, const uint32_t* alive, uint32_t* bitmap, int first,
                             void* stream) {
    VDB_REQUIRE(n >= 0 && (n == 0 || (column && bitmap)), VDB_ERR_INVALID, "vdb_filter_eq: bad arguments (n=%lld)", (long long)n);
    if (n == 0) return VDB_OK;
    const int64_t words = (n + 31) >> 5;
    const int64_t trips = (words + 31) / 32;                    // warp trips
    int blocks = (int)((trips + 7) / 8);                        // 8 warps per block
    const int cap = sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    filter_eq_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(column, n, want, alive, bitmap, first);
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}
