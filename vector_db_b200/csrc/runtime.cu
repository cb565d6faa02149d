// Runtime plumbing of libvdb_b200.so: error strings, launch accounting, storage staging
// (reference layer-1 memmap AoS -> SoA in HBM) and the per-row preparation kernel
// (squared norms + bf16 mirror) that feeds the tensor-core scan.
#include <stdarg.h>

#include <string.h>

#include <atomic>
#include <mutex>
#include <thread>
#include <utility>
#include <vector>

#include "common.cuh"

namespace vdb {

static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    set_error("CUDA error %s (%s) at %s:%d in %s", cudaGetErrorName(e), cudaGetErrorString(e), file, line, what);
    return VDB_ERR_CUDA;
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

static std::atomic<int> g_prof_on{0};
static std::mutex g_prof_mu;
static std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_prof[kProfSlots];

ProfScope::ProfScope(int slot_, cudaStream_t st_) : slot(slot_), st(st_), on(g_prof_on.load() != 0) {
    if (!on) return;
    cudaEvent_t a, b;
    if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) { on = false; return; }
    cudaEventRecord(a, st);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_prof[slot].push_back({a, b});
}
ProfScope::~ProfScope() {
    if (!on) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    cudaEventRecord(g_prof[slot].back().second, st);
}

int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 148;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

int max_smem_optin() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 227 * 1024;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess || n <= 0)
            n = 227 * 1024;
    }
    return n;
}

// ---------------------------------------------------------------------------------------------
// sqnorm[r] = sum_t x[r][t]^2 (fp32, lane-strided + shuffle tree) ; mirror[r][t] = bf16_rn(x[r][t])
// One warp per row; 128-bit loads when D % 4 == 0.  HBM-bound: reads N*D*4, writes N*D*2 + N*4.
__global__ void prepare_rows_kernel(const float* __restrict__ x, int64_t N, int D,
                                    float* __restrict__ sqnorm, __nv_bfloat16* __restrict__ mirror) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5); r < N; r += warps) {
        const float* row = x + r * D;
        float acc = 0.f;
        if ((D & 3) == 0) {
            const float4* row4 = reinterpret_cast<const float4*>(row);
            for (int t = lane; t < (D >> 2); t += 32) {
                float4 v = ld_stream_f4(row4 + t);
                acc = fmaf(v.x, v.x, acc); acc = fmaf(v.y, v.y, acc);
                acc = fmaf(v.z, v.z, acc); acc = fmaf(v.w, v.w, acc);
                if (mirror) {
                    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
                    __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
                    uint2 pk;
                    pk.x = *reinterpret_cast<uint32_t*>(&lo);
                    pk.y = *reinterpret_cast<uint32_t*>(&hi);
                    *reinterpret_cast<uint2*>(mirror + r * D + 4 * t) = pk;
                }
            }
        } else {
            for (int t = lane; t < D; t += 32) {
                float v = row[t];
                acc = fmaf(v, v, acc);
                if (mirror) mirror[r * D + t] = __float2bfloat16_rn(v);
            }
        }
        acc = warp_sum(acc);
        if (sqnorm && lane == 0) sqnorm[r] = acc;
    }
}

// AoS records {int64 id; float32 emb[D]} (stride 8 + 4D, storage.py:135-144) -> SoA, one thread per 32-bit word.
// The copy engine moves whole chunks at the contiguous PCIe rate; splitting the columns there (cudaMemcpy2D with 8-byte
// and 4D-byte rows) ran at 11.8 GB/s against 55 GB/s for the same bytes as one copy -- so the split is this kernel,
// HBM-bound and ~1 ms per 1.5 GB.
__global__ void aos_split_kernel(const uint32_t* __restrict__ rec, int64_t n, int D, uint32_t* __restrict__ ids,
                                 uint32_t* __restrict__ vecs) {
    const int wpr = D + 2;                                   // 32-bit words per record
    const int64_t total = n * wpr;
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = w / wpr;
        const int c = (int)(w - r * wpr);
        const uint32_t v = rec[w];
        if (c < 2) { if (ids) ids[2 * r + c] = v; }
        else vecs[r * D + (c - 2)] = v;
    }
}

}  // namespace vdb

using namespace vdb;

extern "C" {

const char* vdb_last_error(void) { return vdb::g_err; }
int vdb_abi_version(void) { return 4; }
uint64_t vdb_launch_count(void) { return vdb::g_launches.load(); }

void vdb_profile_enable(int on) { vdb::g_prof_on.store(on); }

int vdb_profile_read(int slot, int* launches, float* total_ms) {
    VDB_REQUIRE(slot >= 0 && slot < kProfSlots && launches && total_ms, VDB_ERR_INVALID, "vdb_profile_read: bad slot");
    std::lock_guard<std::mutex> lk(vdb::g_prof_mu);
    float sum = 0.f;
    int n = 0;
    for (auto& pr : vdb::g_prof[slot]) {
        float ms = 0.f;
        if (cudaEventSynchronize(pr.second) == cudaSuccess && cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) {
            sum += ms;
            ++n;
        }
        cudaEventDestroy(pr.first);
        cudaEventDestroy(pr.second);
    }
    vdb::g_prof[slot].clear();
    *launches = n;
    *total_ms = sum;
    return VDB_OK;
}

int vdb_stage_aos_h2d(const void* host_records, int64_t n, int D, int64_t* ids, float* vecs, void* stream) {
    VDB_REQUIRE(n >= 0 && D > 0, VDB_ERR_INVALID, "vdb_stage_aos_h2d: bad n=%lld D=%d", (long long)n, D);
    if (n == 0) return VDB_OK;
    VDB_REQUIRE(host_records && vecs, VDB_ERR_INVALID, "vdb_stage_aos_h2d: null pointer");
    const size_t stride = 8 + 4 * (size_t)D;
    const char* base = static_cast<const char*>(host_records);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (ids) VDB_CUDA(cudaMemcpy2DAsync(ids, 8, base, stride, 8, (size_t)n, cudaMemcpyHostToDevice, s));
    VDB_CUDA(cudaMemcpy2DAsync(vecs, 4 * (size_t)D, base + 8, stride, 4 * (size_t)D, (size_t)n,
                               cudaMemcpyHostToDevice, s));
    return VDB_OK;
}

// Pageable source (the memmap itself): worker threads fill a ring of pinned buffers while the copy engine drains the
// previous ones.  A single-threaded fill runs at one core's page-cache memcpy rate (a few GB/s), an order of
// magnitude under the PCIe rate of the strided DMA that follows it -- the fill, not the link, bounded the upload.
int vdb_stage_aos_pageable_h2d(const void* records, int64_t n, int D, int64_t* ids, float* vecs, void* pinned_ring,
                               void* device_ring, size_t buffer_bytes, int depth, int threads, void* stream) {
    VDB_REQUIRE(n >= 0 && D > 0, VDB_ERR_INVALID, "vdb_stage_aos_pageable_h2d: bad n=%lld D=%d", (long long)n, D);
    if (n == 0) return VDB_OK;
    const size_t stride = 8 + 4 * (size_t)D;
    VDB_REQUIRE(records && vecs && pinned_ring && device_ring, VDB_ERR_INVALID, "vdb_stage_aos_pageable_h2d: null pointer");
    VDB_REQUIRE(depth >= 2 && depth <= 16 && buffer_bytes >= stride, VDB_ERR_INVALID,
                "vdb_stage_aos_pageable_h2d: ring of %d x %zu bytes cannot hold a record of %zu", depth, buffer_bytes, stride);
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;
    const int64_t per_buf = (int64_t)(buffer_bytes / stride);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    std::vector<cudaEvent_t> done((size_t)depth, nullptr);
    int rc = VDB_OK;
    const char* src = static_cast<const char*>(records);
    char* ring = static_cast<char*>(pinned_ring);
    int64_t chunk_i = 0;
    for (int64_t s = 0; s < n && rc == VDB_OK; s += per_buf, ++chunk_i) {
        const int64_t cnt = (n - s < per_buf) ? n - s : per_buf;
        const int b = (int)(chunk_i % depth);
        char* buf = ring + (size_t)b * buffer_bytes;
        if (done[b]) {                                   // the DMA that last read this buffer has finished
            if (cudaEventSynchronize(done[b]) != cudaSuccess) { rc = cuda_fail(cudaGetLastError(), "event wait", __FILE__, __LINE__); break; }
        } else if (cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming) != cudaSuccess) {
            rc = cuda_fail(cudaGetLastError(), "event create", __FILE__, __LINE__);
            break;
        }
        const size_t bytes = (size_t)cnt * stride;
        const char* from = src + (size_t)s * stride;
        int nt = threads;
        if ((size_t)nt * (1u << 20) > bytes) nt = (int)(bytes >> 20) + 1;      // at least 1 MB per thread
        if (nt <= 1) {
            memcpy(buf, from, bytes);
        } else {
            std::vector<std::thread> pool;
            pool.reserve((size_t)nt);
            const size_t per = ((bytes / (size_t)nt) + 4095) & ~(size_t)4095;  // page-aligned slices
            for (int t = 0; t < nt; ++t) {
                const size_t o = (size_t)t * per;
                if (o >= bytes) break;
                const size_t len = (o + per < bytes) ? per : bytes - o;
                pool.emplace_back([=] { memcpy(buf + o, from + o, len); });
            }
            for (auto& th : pool) th.join();
        }
        // one contiguous DMA per chunk into the device ring, then the column split on the device (stream order keeps
        // a device buffer's kernel ahead of the copy that reuses it `depth` chunks later)
        char* dbuf = static_cast<char*>(device_ring) + (size_t)b * buffer_bytes;
        cudaError_t e = cudaMemcpyAsync(dbuf, buf, bytes, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaEventRecord(done[b], st);
        if (e == cudaSuccess) {
            const int64_t words = cnt * (int64_t)(D + 2);
            int blocks = (int)((words + 1023) / 1024);
            if (blocks > sm_count() * 16) blocks = sm_count() * 16;
            aos_split_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const uint32_t*>(dbuf), cnt, D,
                                                     ids ? reinterpret_cast<uint32_t*>(ids + s) : nullptr,
                                                     reinterpret_cast<uint32_t*>(vecs + (size_t)s * D));
            count_launch(1);
            e = cudaGetLastError();
        }
        if (e != cudaSuccess) rc = cuda_fail(e, "chunk H2D copy + split", __FILE__, __LINE__);
    }
    // the ring belongs to the caller, who may free it on return: drain the copies that still read it
    for (auto ev : done) {
        if (!ev) continue;
        if (rc == VDB_OK && cudaEventSynchronize(ev) != cudaSuccess) rc = cuda_fail(cudaGetLastError(), "event wait", __FILE__, __LINE__);
        cudaEventDestroy(ev);
    }
    return rc;
}

int vdb_prepare_rows(const float* vecs, int64_t N, int D, float* sqnorm, void* bf16_mirror, void* stream) {
    VDB_REQUIRE(N >= 0 && D > 0, VDB_ERR_INVALID, "vdb_prepare_rows: bad N=%lld D=%d", (long long)N, D);
    if (N == 0) return VDB_OK;
    VDB_REQUIRE(vecs != nullptr, VDB_ERR_INVALID, "vdb_prepare_rows: null input");
    int blocks = sm_count() * 8;
    int64_t need = (N + 7) / 8;
    if (need < blocks) blocks = (int)need;
    prepare_rows_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        vecs, N, D, sqnorm, static_cast<__nv_bfloat16*>(bf16_mirror));
    count_launch(1);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

}  // extern "C"
