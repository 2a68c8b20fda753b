// K3a / K3c — the three passes of a device-mode powerSGD step (engine semantics §2.3 E4;
// engine name comps/__init__.py:16, results nnlogs.ipynb:48-50), each fused down to the HBM traffic the
// algorithm needs.  With rank r <= 8 every product has r flops per matrix element loaded
// (~1 flop/byte): the bound is HBM bandwidth, not the tensor cores, so these are streaming kernels.
//
//   pass P : M = grad + E (error feedback), written back into E;  P = M·Q       reads grad, E; writes E     (12 B/elem)
//            -> all-reduce(P) over the sites, orthogonalise (csrc/dsgd.cu reduce-only mode, csrc/lowrank.cu)
//   pass Q : Q' = Mᵀ·P̂                                                           reads E                      ( 4 B/elem)
//            -> all-reduce(Q')
//   finish : M̂ = P̂·Q̄ᵀ (rank-r, formed on the fly) -> grad;  E = M − M̂             reads E; writes E, grad     (12 B/elem)
//
// No bf16 copies, no transposes, no temporaries; the products are exact fp32.  A fourth small kernel moves the
// dense (1-D) gradients to and from the exchange buffer in one launch.
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

constexpr int kMaxR = DNB_PSGD_MAX_RANK;

// ---- pass P: one warp per row ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) psgd_p_kernel(const float* __restrict__ grad, float* __restrict__ E,
                                                     const float* __restrict__ Q, float* __restrict__ P, int n, int m,
                                                     int r) {
    dnb::pdl_wait();
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const bool vec = (m & 3) == 0 && ((reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(E)) & 15) == 0;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
        float acc[kMaxR];
#pragma unroll
        for (int k = 0; k < kMaxR; ++k) acc[k] = 0.f;
        const float* g = grad + (long long)i * m;
        float* e = E + (long long)i * m;
        if (vec) {
            for (int j = lane * 4; j < m; j += 128) {
                const float4 a = __ldcs(reinterpret_cast<const float4*>(g + j));
                const float4 b = *reinterpret_cast<const float4*>(e + j);
                const float mv[4] = {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w};
                *reinterpret_cast<float4*>(e + j) = make_float4(mv[0], mv[1], mv[2], mv[3]);
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const float* q = Q + (long long)(j + t) * r;
#pragma unroll
                    for (int k = 0; k < kMaxR; ++k)
                        if (k < r) acc[k] = fmaf(mv[t], __ldg(q + k), acc[k]);
                }
            }
        } else {
            for (int j = lane; j < m; j += 32) {
                const float mv = g[j] + e[j];
                e[j] = mv;
                const float* q = Q + (long long)j * r;
#pragma unroll
                for (int k = 0; k < kMaxR; ++k)
                    if (k < r) acc[k] = fmaf(mv, __ldg(q + k), acc[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < kMaxR; ++k)
            if (k < r) {
                const float s = dnb::warp_sum(acc[k]);
                if (lane == 0) P[(long long)i * r + k] = s;
            }
    }
}

// ---- pass Q: one thread per column, a block of rows per CTA, fp32 atomics into the (pre-zeroed) exchange buffer
constexpr int kQRows = 128;
__global__ void __launch_bounds__(256) psgd_q_kernel(const float* __restrict__ E, const float* __restrict__ Phat,
                                                     float* __restrict__ Qout, int n, int m, int r) {
    dnb::pdl_wait();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i0 = blockIdx.y * kQRows;
    const int i1 = min(n, i0 + kQRows);
    if (j >= m) return;
    float acc[kMaxR];
#pragma unroll
    for (int k = 0; k < kMaxR; ++k) acc[k] = 0.f;
#pragma unroll 4
    for (int i = i0; i < i1; ++i) {
        const float mv = E[(long long)i * m + j];
        const float* p = Phat + (long long)i * r;
#pragma unroll
        for (int k = 0; k < kMaxR; ++k)
            if (k < r) acc[k] = fmaf(mv, __ldg(p + k), acc[k]);
    }
#pragma unroll
    for (int k = 0; k < kMaxR; ++k)
        if (k < r) atomicAdd(Qout + (long long)j * r + k, acc[k]);
}

// ---- finish: M̂ = P̂·Q̄ᵀ -> grad, E = M − M̂ ------------------------------------------------------------------
__global__ void __launch_bounds__(256) psgd_finish_kernel(float* __restrict__ E, float* __restrict__ grad,
                                                          const float* __restrict__ Phat, const float* __restrict__ Qbar,
                                                          int n, int m, int r) {
    dnb::pdl_wait();
    const bool vec = (m & 3) == 0 && ((reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(E)) & 15) == 0;
    const long long total = (long long)n * m;
    const long long stride = (long long)gridDim.x * blockDim.x;
    if (vec) {
        const int m4 = m >> 2;
        for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total >> 2; idx += stride) {
            const int i = (int)(idx / m4), j = (int)(idx % m4) * 4;
            float p[kMaxR];
#pragma unroll
            for (int k = 0; k < kMaxR; ++k) p[k] = k < r ? __ldg(Phat + (long long)i * r + k) : 0.f;
            float mh[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const float* q = Qbar + (long long)(j + t) * r;
                float s = 0.f;
#pragma unroll
                for (int k = 0; k < kMaxR; ++k)
                    if (k < r) s = fmaf(p[k], __ldg(q + k), s);
                mh[t] = s;
            }
            float4* ep = reinterpret_cast<float4*>(E + (long long)i * m + j);
            const float4 mv = *ep;
            *ep = make_float4(mv.x - mh[0], mv.y - mh[1], mv.z - mh[2], mv.w - mh[3]);
            *reinterpret_cast<float4*>(grad + (long long)i * m + j) = make_float4(mh[0], mh[1], mh[2], mh[3]);
        }
    } else {
        for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
            const int i = (int)(idx / m), j = (int)(idx % m);
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < kMaxR; ++k)
                if (k < r) s = fmaf(__ldg(Phat + (long long)i * r + k), __ldg(Qbar + (long long)j * r + k), s);
            E[idx] -= s;
            grad[idx] = s;
        }
    }
}

// ---- segment copies (dense gradients <-> exchange buffer), one launch ------------------------------------------
struct SegArgs {
    const float* src[DNB_PSGD_MAX_SEGS];
    float* dst[DNB_PSGD_MAX_SEGS];
    int len[DNB_PSGD_MAX_SEGS];
    int count;
};
__global__ void __launch_bounds__(256) seg_copy_kernel(const SegArgs a) {
    dnb::pdl_wait();
    const int k = blockIdx.y;
    const float* src = nullptr;
    float* dst = nullptr;
    int len = 0;
#pragma unroll
    for (int i = 0; i < DNB_PSGD_MAX_SEGS; ++i)
        if (i == k) { src = a.src[i]; dst = a.dst[i]; len = a.len[i]; }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < len; i += gridDim.x * blockDim.x) dst[i] = src[i];
}

}  // namespace

DNB_EXPORT int dnb_psgd_pass(const DnbPsgdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_psgd_pass: null descriptor");
    DNB_REQUIRE(d->n > 0 && d->m > 0 && d->r >= 1 && d->r <= kMaxR, "dnb_psgd_pass: bad shape %d x %d, rank %d (max %d)", d->n,
                d->m, d->r, kMaxR);
    DNB_REQUIRE(d->E != nullptr, "dnb_psgd_pass: E is null");
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (d->pass == 0) {
        DNB_REQUIRE(d->grad && d->Q && d->P, "pass P needs grad, Q and P");
        int blocks = (d->n + 7) / 8;                 // 8 warps (rows) per CTA
        if (blocks > sms * 8) blocks = sms * 8;
        DNB_CUDA_CHECK(dnb_launch(psgd_p_kernel, dim3(blocks), dim3(256), 0, st, 0, (const float*)d->grad, d->E,
                                  (const float*)d->Q, d->P, d->n, d->m, d->r));
    } else if (d->pass == 1) {
        DNB_REQUIRE(d->P && d->Q, "pass Q needs P (orthonormalised) and the Q output");
        dim3 grid((d->m + 255) / 256, (d->n + kQRows - 1) / kQRows);
        DNB_CUDA_CHECK(dnb_launch(psgd_q_kernel, grid, dim3(256), 0, st, 0, (const float*)d->E, (const float*)d->P, d->Q,
                                  d->n, d->m, d->r));
    } else if (d->pass == 2) {
        DNB_REQUIRE(d->grad && d->Q && d->P, "finish needs grad, P and Q");
        long long work = ((long long)d->n * d->m + 3) / 4;
        int blocks = (int)((work + 255) / 256);
        if (blocks > sms * 16) blocks = sms * 16;
        if (blocks < 1) blocks = 1;
        DNB_CUDA_CHECK(dnb_launch(psgd_finish_kernel, dim3(blocks), dim3(256), 0, st, 0, d->E, d->grad, (const float*)d->P,
                                  (const float*)d->Q, d->n, d->m, d->r));
    } else {
        DNB_REQUIRE(false, "pass must be 0 (P), 1 (Q) or 2 (finish)");
    }
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

DNB_EXPORT int dnb_seg_copy(const float* const* src, float* const* dst, const int* len, int count, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(src && dst && len && count >= 0, "dnb_seg_copy: bad arguments");
    for (int off = 0; off < count; off += DNB_PSGD_MAX_SEGS) {
        SegArgs a;
        memset(&a, 0, sizeof(a));
        const int c = count - off < DNB_PSGD_MAX_SEGS ? count - off : DNB_PSGD_MAX_SEGS;
        int mx = 1;
        for (int i = 0; i < c; ++i) {
            DNB_REQUIRE(src[off + i] && dst[off + i] && len[off + i] >= 0, "dnb_seg_copy: bad segment %d", off + i);
            a.src[i] = src[off + i]; a.dst[i] = dst[off + i]; a.len[i] = len[off + i];
            if (len[off + i] > mx) mx = len[off + i];
        }
        a.count = c;
        int bx = (mx + 255) / 256;
        if (bx > 128) bx = 128;
        DNB_CUDA_CHECK(dnb_launch(seg_copy_kernel, dim3(bx, c), dim3(256), 0, st, 0, a));
        DNB_LAUNCH_CHECK();
        dnb_count_launch(1);
    }
    return 0;
}

DNB_EXPORT int dnb_memset_async(void* p, int value, size_t bytes, void* stream_) {
    DNB_REQUIRE(p != nullptr || bytes == 0, "dnb_memset_async: null pointer");
    DNB_CUDA_CHECK(cudaMemsetAsync(p, value, bytes, reinterpret_cast<cudaStream_t>(stream_)));
    return 0;
}
