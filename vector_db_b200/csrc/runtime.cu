// Runtime plumbing of libvdb_b200.so: error strings, launch accounting, storage staging
// (reference layer-1 memmap AoS -> SoA in HBM) and the per-row preparation kernel
// (squared norms + bf16 mirror) that feeds the tensor-core scan.
#include <stdarg.h>

#include <string.h>

#include <atomic>
#include <mutex>
#include <string>
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
    VDB_REQUIRE(depth >= 2 && depth <= 128 && buffer_bytes >= stride, VDB_ERR_INVALID,
                "vdb_stage_aos_pageable_h2d: ring of %d x %zu bytes cannot hold a record of %zu", depth, buffer_bytes, stride);
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;
    if (threads > depth / 2) threads = depth / 2;        // every worker double-buffers
    const int64_t per_buf = (int64_t)(buffer_bytes / stride);
    const int64_t n_chunks = (n + per_buf - 1) / per_buf;
    if (threads > n_chunks) threads = (int)n_chunks;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int dev = 0;
    VDB_CUDA(cudaGetDevice(&dev));
    cudaEvent_t start;
    VDB_CUDA(cudaEventCreateWithFlags(&start, cudaEventDisableTiming));
    VDB_CUDA(cudaEventRecord(start, st));
    // Worker t owns chunks t, t + T, ... and buffers t, t + T, ... of both rings, and its own stream: fill (host memcpy out of
    // the page cache), one contiguous DMA, column split on the device -- no barrier between the workers, so the fills of
    // some overlap the DMAs of others all the time.  (One fill at a time by a thread pool spawned per chunk: 33 GB/s; a
    // single thread: 13.6 GB/s; the box's pinned H2D rate: 54 GB/s.)
    std::vector<cudaStream_t> streams((size_t)threads, nullptr);
    std::vector<cudaEvent_t> done((size_t)threads, nullptr);
    std::vector<int> rcs((size_t)threads, VDB_OK);
    std::vector<std::string> errs((size_t)threads);
    const int per_worker = depth / threads;                // buffers per worker (>= 2)
    const char* src = static_cast<const char*>(records);
    const int blocks_cap = sm_count() * 4;
    auto worker = [&](int t) {
        auto fail = [&](cudaError_t e, const char* what) {
            rcs[t] = VDB_ERR_CUDA;
            errs[t] = std::string(what) + ": " + cudaGetErrorString(e);
        };
        cudaError_t e = cudaSetDevice(dev);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&streams[t], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&done[t], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(streams[t], start, 0);
        if (e != cudaSuccess) { fail(e, "worker set-up"); return; }
        std::vector<cudaEvent_t> free_ev((size_t)per_worker, nullptr);
        int64_t k = 0;
        for (int64_t c = t; c < n_chunks; c += threads, ++k) {
            const int slot = (int)(k % per_worker);
            const int b = t + slot * threads;
            char* hbuf = static_cast<char*>(pinned_ring) + (size_t)b * buffer_bytes;
            char* dbuf = static_cast<char*>(device_ring) + (size_t)b * buffer_bytes;
            if (free_ev[slot]) e = cudaEventSynchronize(free_ev[slot]);         // the DMA that last read this buffer is done
            else e = cudaEventCreateWithFlags(&free_ev[slot], cudaEventDisableTiming);
            if (e != cudaSuccess) { fail(e, "buffer wait"); break; }
            const int64_t s0 = c * per_buf;
            const int64_t cnt = (n - s0 < per_buf) ? n - s0 : per_buf;
            const size_t bytes = (size_t)cnt * stride;
            memcpy(hbuf, src + (size_t)s0 * stride, bytes);
            e = cudaMemcpyAsync(dbuf, hbuf, bytes, cudaMemcpyHostToDevice, streams[t]);
            if (e == cudaSuccess) e = cudaEventRecord(free_ev[slot], streams[t]);
            if (e != cudaSuccess) { fail(e, "chunk H2D copy"); break; }
            const int64_t words = cnt * (int64_t)(D + 2);
            int blocks = (int)((words + 1023) / 1024);
            if (blocks > blocks_cap) blocks = blocks_cap;
            aos_split_kernel<<<blocks, 256, 0, streams[t]>>>(reinterpret_cast<const uint32_t*>(dbuf), cnt, D,
                                                             ids ? reinterpret_cast<uint32_t*>(ids + s0) : nullptr,
                                                             reinterpret_cast<uint32_t*>(vecs + (size_t)s0 * D));
            count_launch(1);
            e = cudaGetLastError();
            if (e != cudaSuccess) { fail(e, "column split launch"); break; }
        }
        // the rings belong to the caller, who may free them on return: drain this worker's copies and kernels
        e = cudaEventRecord(done[t], streams[t]);
        if (e == cudaSuccess) e = cudaStreamSynchronize(streams[t]);
        if (e != cudaSuccess && rcs[t] == VDB_OK) fail(e, "drain");
        for (auto ev : free_ev)
            if (ev) cudaEventDestroy(ev);
    };
    if (threads == 1) {
        worker(0);
    } else {
        std::vector<std::thread> pool;
        pool.reserve((size_t)threads);
        for (int t = 0; t < threads; ++t) pool.emplace_back(worker, t);
        for (auto& th : pool) th.join();
    }
    int rc = VDB_OK;
    for (int t = 0; t < threads; ++t) {
        if (done[t]) {
            if (rcs[t] == VDB_OK) cudaStreamWaitEvent(st, done[t], 0);           // later work on the caller's stream sees the rows
            cudaEventDestroy(done[t]);
        }
        if (streams[t]) cudaStreamDestroy(streams[t]);
        if (rcs[t] != VDB_OK && rc == VDB_OK) {
            rc = rcs[t];
            set_error("vdb_stage_aos_pageable_h2d: %s", errs[t].c_str());
        }
    }
    cudaEventDestroy(start);
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
