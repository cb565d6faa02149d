// Micro-benchmark (dev tool): the ceiling of the HNSW access pattern -- random gathers of whole fp32 rows (D = 768: 3 KB
// contiguous) out of a 1M-row matrix (3 GB, 24x the L2), one warp per ROWS_PER_WARP rows at a time, all 128-bit loads of
// those rows issued before any is consumed.  Prints GB/s for several occupancies / rows in flight / load flavours.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float4 ld_na(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

template <int RPW, bool NOALLOC>
__global__ void __launch_bounds__(128) gather(const float* __restrict__ db, const int* __restrict__ rows, int n_rows_total,
                                              int D, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    float acc = 0.f;
    for (int base = warp_global * RPW; base + RPW <= n_rows_total; base += warps * RPW) {
        float4 v[RPW][6];
#pragma unroll
        for (int j = 0; j < RPW; ++j) {
            const float4* p = reinterpret_cast<const float4*>(db + (size_t)rows[base + j] * D);
#pragma unroll
            for (int c = 0; c < 6; ++c) v[j][c] = NOALLOC ? ld_na(p + lane + 32 * c) : __ldg(p + lane + 32 * c);
        }
#pragma unroll
        for (int j = 0; j < RPW; ++j)
#pragma unroll
            for (int c = 0; c < 6; ++c) acc += v[j][c].x + v[j][c].y + v[j][c].z + v[j][c].w;
    }
    if (acc == 123.456f) out[0] = acc;
}

template <int RPW, bool NOALLOC>
void run(const char* name, int ctas_per_sm, const float* db, const int* rows, int n, float* out) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    gather<RPW, NOALLOC><<<148 * ctas_per_sm, 128>>>(db, rows, n, 768, out);
    cudaEventRecord(a);
    gather<RPW, NOALLOC><<<148 * ctas_per_sm, 128>>>(db, rows, n, 768, out);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    printf("%-34s ctas/SM %2d  %7.3f ms  %7.1f GB/s  %s\n", name, ctas_per_sm, ms, (double)n * 3072 / (ms * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    const int N = 1000000, D = 768, n = 8 << 20;          // 8M row gathers = 25.8 GB
    float *db, *out;
    int* rows;
    cudaMalloc(&db, (size_t)N * D * 4);
    cudaMemset(db, 0, (size_t)N * D * 4);
    cudaMalloc(&out, 16);
    int* h = (int*)malloc((size_t)n * 4);
    uint64_t s = 88172645463325252ull;
    for (int i = 0; i < n; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int)(s % N); }
    cudaMalloc(&rows, (size_t)n * 4);
    cudaMemcpy(rows, h, (size_t)n * 4, cudaMemcpyHostToDevice);
    for (int c : {4, 8, 12, 16}) {
        run<1, false>("1 row/warp in flight, ldg", c, db, rows, n, out);
        run<2, false>("2 rows/warp in flight, ldg", c, db, rows, n, out);
        run<2, true>("2 rows/warp in flight, no_allocate", c, db, rows, n, out);
        run<4, true>("4 rows/warp in flight, no_allocate", c, db, rows, n, out);
    }
    // sequential rows for comparison (streaming read of the same bytes)
    for (int i = 0; i < n; ++i) h[i] = i % N;
    cudaMemcpy(rows, h, (size_t)n * 4, cudaMemcpyHostToDevice);
    run<2, true>("SEQUENTIAL rows, 2/warp, no_alloc", 16, db, rows, n, out);
    return 0;
}
