// K4 (backward half) — BatchNorm(batch statistics) + ReLU backward for the FreeSurfer MLP layers.
//
// Forward of a hidden layer (reference comps/fs/models.py:13-23: Linear(bias=False) -> BatchNorm1d
// (track_running_stats=False) -> ReLU) is ONE tcgen05 GEMM with the BatchNorm epilogue of
// csrc/gemm.cu: weights are the M operand, the batch is the N tile, so each TMEM lane holds one
// feature for the whole batch and the statistics need no reduction across threads.
// This file holds the matching backward through ReLU and BatchNorm.  Activations are kept
// feature-major ([features, batch], the natural GEMM output), so one thread owns one feature row.
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

__global__ void bn_relu_bwd_kernel(const float* __restrict__ daT, long long ldd, const float* __restrict__ xhatT,
                                   long long ldx, const float* __restrict__ gamma, const float* __restrict__ beta,
                                   const float* __restrict__ invstd, int N, int B, __nv_bfloat16* __restrict__ dzT,
                                   long long ldzt, __nv_bfloat16* __restrict__ dz, long long ldz,
                                   float* __restrict__ dgamma, float* __restrict__ dbeta, int accumulate) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    const float g = gamma[j], be = beta ? beta[j] : 0.f, istd = invstd[j];
    const float* da = daT + (long long)j * ldd;
    const float* xh = xhatT + (long long)j * ldx;
    float sg = 0.f, sb = 0.f;
    for (int b = 0; b < B; ++b) {
        const float x = xh[b];
        const float dy = (x * g + be > 0.f) ? da[b] : 0.f;
        sg += dy * x;
        sb += dy;
    }
    if (dgamma) dgamma[j] = accumulate ? dgamma[j] + sg : sg;
    if (dbeta) dbeta[j] = accumulate ? dbeta[j] + sb : sb;
    const float invB = 1.f / (float)B, gi = g * istd;
    for (int b = 0; b < B; ++b) {
        const float x = xh[b];
        const float dy = (x * g + be > 0.f) ? da[b] : 0.f;
        const __nv_bfloat16 v = __float2bfloat16(gi * (dy - sb * invB - x * sg * invB));
        if (dzT) dzT[(long long)j * ldzt + b] = v;
        if (dz) dz[(long long)b * ldz + j] = v;
    }
}

}  // namespace

DNB_EXPORT int dnb_bn_relu_backward(const float* daT, long long ldd, const float* xhatT, long long ldx,
                                    const float* gamma, const float* beta, const float* invstd, int N, int B,
                                    void* dzT, long long ldzt, void* dz, long long ldz, float* dgamma, float* dbeta,
                                    int accumulate, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(daT && xhatT && gamma && invstd && N > 0 && B > 0, "dnb_bn_relu_backward: bad arguments");
    DNB_REQUIRE(dzT || dz, "dnb_bn_relu_backward: no output requested");
    bn_relu_bwd_kernel<<<dnb_ceil_div(N, 128), 128, 0, st>>>(daT, ldd, xhatT, ldx, gamma, beta, invstd, N, B,
                                                             reinterpret_cast<__nv_bfloat16*>(dzT), ldzt,
                                                             reinterpret_cast<__nv_bfloat16*>(dz), ldz, dgamma, dbeta,
                                                             accumulate);
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}
