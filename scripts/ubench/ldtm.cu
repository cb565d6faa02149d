// Dev probe (not product): TMEM read throughput of tcgen05.ld 32x32b.{x16,x32} with and without .pack::16b,
// one or two warps per SM sub-partition, back-to-back (one wait per load) or four loads per wait.
#include <cstdio>
#include <cstdint>
#include "../../vector_db_b200/csrc/sm100.cuh"
using namespace vdb::sm100;

#define LD32(PACK) \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32" PACK ".b32 " \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, " \
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];" \
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), \
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), \
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), \
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) \
        : "r"(taddr) : "memory")
#define LD16(PACK) \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16" PACK ".b32 " \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];" \
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), \
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) \
        : "r"(taddr) : "memory")

template <int MODE>   // 0: x32, 1: x32 pack, 2: x16, 3: x16 pack, 4: x32 pack, 4 loads per wait (different columns)
__global__ void k(long long* clk, uint32_t* sink, int iters) {
    __shared__ uint32_t tptr;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc(&tptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = tptr + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t r[32];
    uint32_t acc = 0;
    for (int i = 0; i < 32; ++i) r[i] = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const uint32_t taddr = base + (uint32_t)((it & 3) * 64);
        if (MODE == 0) { LD32(""); }
        if (MODE == 1) { LD32(".pack::16b"); }
        if (MODE == 2) { LD16(""); }
        if (MODE == 3) { LD16(".pack::16b"); }
        if (MODE == 4) {
            LD32(".pack::16b");
            acc ^= r[0];
            { const uint32_t taddr = base + 64u; LD32(".pack::16b"); } acc ^= r[1];
            { const uint32_t taddr = base + 128u; LD32(".pack::16b"); } acc ^= r[2];
            { const uint32_t taddr = base + 192u; LD32(".pack::16b"); }
        }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc ^= r[3] ^ r[17];
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) *clk = t1 - t0;
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tptr, 512);
}

template <int MODE>
void run(const char* name, int threads, double bytes_per_iter_per_warp, long long* dclk, uint32_t* sink) {
    const int iters = 4096;
    k<MODE><<<148, threads>>>(dclk, sink, iters);
    cudaDeviceSynchronize();
    k<MODE><<<148, threads>>>(dclk, sink, iters);
    cudaError_t e = cudaDeviceSynchronize();
    long long c = 0;
    cudaMemcpy(&c, dclk, 8, cudaMemcpyDeviceToHost);
    const int warps = threads / 32;
    printf("%-34s %2d warps: %7.1f clk/iter/warp, %6.1f TMEM-cell bytes/clk/SM  (%s)\n", name, warps, (double)c / iters,
           bytes_per_iter_per_warp * warps * iters / (double)c, cudaGetErrorString(e));
}

int main() {
    long long* dclk; uint32_t* sink;
    cudaMalloc(&dclk, 8); cudaMalloc(&sink, 148 * 1024 * 4);
    for (int threads : {128, 256}) {
        run<0>("x32 (32 cols, 4 KB cells)", threads, 4096, dclk, sink);
        run<1>("x32 pack16 (64 cols, 8 KB cells)", threads, 8192, dclk, sink);
        run<2>("x16 (16 cols, 2 KB cells)", threads, 2048, dclk, sink);
        run<3>("x16 pack16 (32 cols, 4 KB cells)", threads, 4096, dclk, sink);
        run<4>("4 x (x32 pack16) per wait", threads, 4 * 8192, dclk, sink);
    }
    return 0;
}
