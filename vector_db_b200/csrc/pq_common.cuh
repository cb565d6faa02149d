// Arithmetic shared by the PQ encode kernels: the reference's distance in numpy's operation order.
#pragma once
#include "common.cuh"

namespace vdb {

// ------------------------------------------------------------------------------------------
// numpy pairwise_sum order over n rounded squares v[t] = rn(d[t]*d[t])   (n <= 128)
template <int DS>
__device__ __forceinline__ float np_sumsq(const float (&d)[DS]) {
    if (DS < 8) {
        float r = 0.f;
#pragma unroll
        for (int t = 0; t < DS; ++t) r = __fadd_rn(r, __fmul_rn(d[t], d[t]));
        return r;
    } else {
        float r[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) r[t] = __fmul_rn(d[t], d[t]);
        constexpr int main_end = DS - (DS % 8);
#pragma unroll
        for (int i = 8; i < main_end; i += 8) {
#pragma unroll
            for (int t = 0; t < 8; ++t) r[t] = __fadd_rn(r[t], __fmul_rn(d[i + t], d[i + t]));
        }
        float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                              __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
#pragma unroll
        for (int i = main_end; i < DS; ++i) res = __fadd_rn(res, __fmul_rn(d[i], d[i]));
        return res;
    }
}

}  // namespace vdb
