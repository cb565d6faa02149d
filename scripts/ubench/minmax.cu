// Micro-benchmark (dev tool, not product): issue rates of the min/max flavours the PQ-encode fold could use.
// Each kernel runs ILP independent dependency chains per thread for ITERS iterations; 148 x 4 CTAs x 256 threads.
// Output: warp-instructions per clock per SM sub-partition for each flavour, and for pairs issued interleaved
// (a pair that sums to more than either alone runs on two pipes).
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>

constexpr int ILP = 8, ITERS = 4096;

enum { F_FMNMX, F_FMNMX3, F_HMNMX2, F_HMNMX2_3, F_VIMNMX3, F_FFMA, F_MIX_FMNMX3_FFMA, F_MIX_FMNMX3_HMNMX2, F_MIX_FMNMX3_VIMNMX3,
       F_MIX_HMNMX2_FFMA, F_IMNMX, F_F2FP, F_MIX_F2FP_VHMNMX, F_MIX_F2FP_FMNMX3, F_MIX_VHMNMX_FFMA, F_FADD, F_MIX_FADD_FMNMX3, F_F2FP_ONLY, F_MIX_F2FP_VHMNMX_IND, F_MIX_F2FP_FFMA_IND, F_PRMT, F_MIX_PRMT_FMNMX3, F_COUNT };

template <int WHICH>
__global__ void k(float* out, const float* in, long long* clocks) {
    float a[ILP], b[ILP];
    __half2 h[ILP], g[ILP];
    int x[ILP], y[ILP];
    for (int i = 0; i < ILP; ++i) {
        a[i] = in[threadIdx.x + i]; b[i] = in[threadIdx.x + 32 + i];
        h[i] = __floats2half2_rn(a[i], b[i]); g[i] = __floats2half2_rn(b[i], a[i]);
        x[i] = __float_as_int(a[i]); y[i] = __float_as_int(b[i]);
    }
    const float c = in[threadIdx.x + 64];
    const __half2 hc = __floats2half2_rn(c, c);
    const int ic = __float_as_int(c);
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (WHICH == F_FMNMX) a[i] = fminf(a[i], b[i] ), b[i] = fmaxf(b[i], a[i]);
            if (WHICH == F_FMNMX3) { a[i] = fminf(fminf(a[i], b[i]), c); b[i] = fmaxf(fmaxf(b[i], a[i]), c); }
            if (WHICH == F_HMNMX2) { h[i] = __hmin2(h[i], g[i]); g[i] = __hmax2(g[i], h[i]); }
            if (WHICH == F_HMNMX2_3) { h[i] = __hmin2(__hmin2(h[i], g[i]), hc); g[i] = __hmax2(__hmax2(g[i], h[i]), hc); }
            if (WHICH == F_VIMNMX3) { x[i] = min(min(x[i], y[i]), ic); y[i] = max(max(y[i], x[i]), ic); }
            if (WHICH == F_IMNMX) { x[i] = min(x[i], y[i]); y[i] = max(y[i], x[i]); }
            if (WHICH == F_FFMA) { a[i] = fmaf(a[i], b[i], c); b[i] = fmaf(b[i], a[i], c); }
            if (WHICH == F_MIX_FMNMX3_FFMA) { a[i] = fminf(fminf(a[i], b[i]), c); b[i] = fmaf(b[i], c, c); }
            if (WHICH == F_MIX_FMNMX3_HMNMX2) { a[i] = fminf(fminf(a[i], b[i]), c); h[i] = __hmin2(h[i], g[i]); g[i] = __hmax2(g[i], hc);  b[i] = fmaxf(fmaxf(b[i], a[i]), c); }
            if (WHICH == F_MIX_FMNMX3_VIMNMX3) { a[i] = fminf(fminf(a[i], b[i]), c); x[i] = min(min(x[i], y[i]), ic); }
            if (WHICH == F_F2FP) { h[i] = __floats2half2_rn(a[i], __low2float(h[i])); g[i] = __floats2half2_rn(b[i], __high2float(g[i])); }
            if (WHICH == F_MIX_F2FP_VHMNMX) { h[i] = __floats2half2_rn(a[i], b[i]); g[i] = __hmin2(__hmin2(g[i], h[i]), hc); a[i] += 1.0f; }
            if (WHICH == F_MIX_F2FP_FMNMX3) { h[i] = __floats2half2_rn(a[i], b[i]); a[i] = fminf(fminf(a[i], b[i]), c); b[i] = __low2float(h[i]); }
            if (WHICH == F_MIX_VHMNMX_FFMA) { g[i] = __hmin2(__hmin2(g[i], h[i]), hc); a[i] = fmaf(a[i], b[i], c); }
            if (WHICH == F_FADD) { a[i] = a[i] + b[i]; b[i] = b[i] + c; }
            if (WHICH == F_MIX_FADD_FMNMX3) { a[i] = a[i] + c; b[i] = fminf(fminf(b[i], a[i]), c); }
            if (WHICH == F_F2FP_ONLY) { h[i] = __floats2half2_rn(a[i], b[i]); a[i] = __uint_as_float(*reinterpret_cast<unsigned*>(&h[i])); g[i] = __floats2half2_rn(b[i], a[i]); b[i] = __uint_as_float(*reinterpret_cast<unsigned*>(&g[i])); }
            if (WHICH == F_MIX_F2FP_VHMNMX_IND) { h[i] = __floats2half2_rn(a[i], b[i]); a[i] = __uint_as_float(*reinterpret_cast<unsigned*>(&h[i])); g[i] = __hmin2(__hmin2(g[i], hc), hc); }
            if (WHICH == F_MIX_F2FP_FFMA_IND) { h[i] = __floats2half2_rn(a[i], c); a[i] = __uint_as_float(*reinterpret_cast<unsigned*>(&h[i])); b[i] = fmaf(b[i], c, c); }
            if (WHICH == F_PRMT) { x[i] = __byte_perm(x[i], y[i], 0x7632); y[i] = __byte_perm(y[i], x[i], 0x5410); }
            if (WHICH == F_MIX_PRMT_FMNMX3) { x[i] = __byte_perm(x[i], ic, 0x7632); a[i] = fminf(fminf(a[i], b[i]), c); }
            if (WHICH == F_MIX_HMNMX2_FFMA) { h[i] = __hmin2(h[i], g[i]); a[i] = fmaf(a[i], b[i], c); }
        }
    }
    long long t1 = clock64();
    float s = 0.f;
    for (int i = 0; i < ILP; ++i) s += a[i] + b[i] + __low2float(h[i]) + __high2float(g[i]) + (float)x[i] + (float)y[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clocks = t1 - t0;
}

template <int WHICH>
void run(const char* name, int ops_per_iter, float* out, float* in, long long* clk) {
    k<WHICH><<<148 * 4, 256>>>(out, in, clk);
    cudaDeviceSynchronize();
    k<WHICH><<<148 * 4, 256>>>(out, in, clk);
    cudaDeviceSynchronize();
    long long c;
    cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
    // per SM sub-partition: 4 CTAs x 8 warps / 4 SMSPs = 8 warps; warp-instructions = 8 * ITERS * ILP * ops
    double wi = 8.0 * ITERS * ILP * ops_per_iter;
    printf("%-28s %8lld clk  %.3f warp-instr/clk/SMSP (%d ops per chain step)\n", name, c, wi / (double)c, ops_per_iter);
}

int main() {
    float *out, *in;
    long long* clk;
    cudaMalloc(&out, 148 * 4 * 256 * 4);
    cudaMalloc(&in, 4096);
    cudaMemset(in, 0x3c, 4096);
    cudaMalloc(&clk, 8);
    run<F_FMNMX>("FMNMX (2-input f32)", 2, out, in, clk);
    run<F_FMNMX3>("FMNMX3 (3-input f32)", 2, out, in, clk);
    run<F_HMNMX2>("HMNMX2 (f16x2)", 2, out, in, clk);
    run<F_HMNMX2_3>("HMNMX2 3-input? (f16x2)", 2, out, in, clk);
    run<F_IMNMX>("IMNMX (2-input s32)", 2, out, in, clk);
    run<F_VIMNMX3>("VIMNMX3 (3-input s32)", 2, out, in, clk);
    run<F_FFMA>("FFMA", 2, out, in, clk);
    run<F_MIX_FMNMX3_FFMA>("FMNMX3 + FFMA", 2, out, in, clk);
    run<F_MIX_FMNMX3_HMNMX2>("2 FMNMX3 + 2 HMNMX2", 4, out, in, clk);
    run<F_MIX_FMNMX3_VIMNMX3>("FMNMX3 + VIMNMX3", 2, out, in, clk);
    run<F_MIX_HMNMX2_FFMA>("HMNMX2 + FFMA", 2, out, in, clk);
    run<F_F2FP>("F2FP pack (+cvt back)", 2, out, in, clk);
    run<F_MIX_F2FP_VHMNMX>("F2FP + VHMNMX + FADD", 3, out, in, clk);
    run<F_MIX_F2FP_FMNMX3>("F2FP + FMNMX3 + cvt", 3, out, in, clk);
    run<F_MIX_VHMNMX_FFMA>("VHMNMX + FFMA", 2, out, in, clk);
    run<F_F2FP_ONLY>("F2FP only (2 chains)", 2, out, in, clk);
    run<F_MIX_F2FP_VHMNMX_IND>("F2FP + VHMNMX independent", 2, out, in, clk);
    run<F_MIX_F2FP_FFMA_IND>("F2FP + FFMA independent", 2, out, in, clk);
    run<F_PRMT>("PRMT", 2, out, in, clk);
    run<F_MIX_PRMT_FMNMX3>("PRMT + FMNMX3", 2, out, in, clk);
    run<F_FADD>("FADD", 2, out, in, clk);
    run<F_MIX_FADD_FMNMX3>("FADD + FMNMX3", 2, out, in, clk);
    return 0;
}
