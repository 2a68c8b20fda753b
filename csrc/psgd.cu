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

// Q [m, r] row-major -> shared memory, k-major: Qs[k * ldq + j].  A thread that owns four consecutive columns then
// reads its four Q values of one k as ONE conflict-free 16-byte load (lanes 16 bytes apart); reading Q [m, r] in place
// instead costs one 32-bit load per value with the lanes 16*r bytes apart: 16 L1 wavefronts per instruction, which
// capped the first version of these kernels at 1.3 TB/s.
__device__ __forceinline__ void stage_q(const float* __restrict__ Q, float* Qs, int m, int r, int ldq) {
    for (int idx = threadIdx.x; idx < m * r; idx += blockDim.x) Qs[(idx % r) * ldq + idx / r] = __ldg(Q + idx);
    __syncthreads();
}

// ---- pass P: persistent CTAs, one warp per row, Q staged in shared memory --------------------------------------
template <int R>
__global__ void __launch_bounds__(256) psgd_p_smem_kernel(const float* __restrict__ grad, float* __restrict__ E,
                                                          const float* __restrict__ Q, float* __restrict__ P, int n, int m,
                                                          int ldq) {
    extern __shared__ float Qs[];
    dnb::pdl_wait();
    stage_q(Q, Qs, m, R, ldq);
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
        float acc[R];
#pragma unroll
        for (int k = 0; k < R; ++k) acc[k] = 0.f;
        const float* g = grad + (long long)i * m;
        float* e = E + (long long)i * m;
#pragma unroll 4
        for (int j = lane * 4; j < m; j += 128) {
            const float4 a = __ldcs(reinterpret_cast<const float4*>(g + j));
            const float4 b = *reinterpret_cast<const float4*>(e + j);
            const float4 mv = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
            *reinterpret_cast<float4*>(e + j) = mv;
#pragma unroll
            for (int k = 0; k < R; ++k) {
                const float4 q = *reinterpret_cast<const float4*>(Qs + k * ldq + j);
                acc[k] = fmaf(mv.x, q.x, fmaf(mv.y, q.y, fmaf(mv.z, q.z, fmaf(mv.w, q.w, acc[k]))));
            }
        }
#pragma unroll
        for (int k = 0; k < R; ++k) {
            const float s = dnb::warp_sum(acc[k]);
            if (lane == 0) P[(long long)i * R + k] = s;
        }
    }
}

// generic fallback (any m, any r <= 8, Q read in place): one warp per row
__global__ void __launch_bounds__(256) psgd_p_kernel(const float* __restrict__ grad, float* __restrict__ E,
                                                     const float* __restrict__ Q, float* __restrict__ P, int n, int m,
                                                     int r) {
    dnb::pdl_wait();
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
        float acc[kMaxR];
#pragma unroll
        for (int k = 0; k < kMaxR; ++k) acc[k] = 0.f;
        const float* g = grad + (long long)i * m;
        float* e = E + (long long)i * m;
        for (int j = lane; j < m; j += 32) {
            const float mv = g[j] + e[j];
            e[j] = mv;
            const float* q = Q + (long long)j * r;
#pragma unroll
            for (int k = 0; k < kMaxR; ++k)
                if (k < r) acc[k] = fmaf(mv, __ldg(q + k), acc[k]);
        }
#pragma unroll
        for (int k = 0; k < kMaxR; ++k)
            if (k < r) {
                const float s = dnb::warp_sum(acc[k]);
                if (lane == 0) P[(long long)i * r + k] = s;
            }
    }
}

// ---- pass Q: a thread owns 4 consecutive columns (16-byte loads), a CTA a block of rows; fp32 atomics into the
// (pre-zeroed) exchange buffer.  P̂ rows are warp-uniform loads.
constexpr int kQRows = 256;
template <int R>
__global__ void __launch_bounds__(256) psgd_q_vec_kernel(const float* __restrict__ E, const float* __restrict__ Phat,
                                                         float* __restrict__ Qout, int n, int m) {
    // CTA = 64 column groups (4 columns each) x 4 row phases over a block of kQRows rows; the four phases are summed in
    // shared memory, then one set of atomics per column group (the row blocks of different CTAs meet in Qout)
    __shared__ float red[3][64][4 * R + 1];
    dnb::pdl_wait();
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    const int j = (blockIdx.x * 64 + tx) * 4;
    const int i0 = blockIdx.y * kQRows;
    const int i1 = min(n, i0 + kQRows);
    float acc[4][R];
#pragma unroll
    for (int t = 0; t < 4; ++t)
#pragma unroll
        for (int k = 0; k < R; ++k) acc[t][k] = 0.f;
    if (j < m) {
#pragma unroll 4
        for (int i = i0 + ty; i < i1; i += 4) {
            const float4 mv = *reinterpret_cast<const float4*>(E + (long long)i * m + j);
            const float* p = Phat + (long long)i * R;
#pragma unroll
            for (int k = 0; k < R; ++k) {
                const float pk = __ldg(p + k);
                acc[0][k] = fmaf(mv.x, pk, acc[0][k]);
                acc[1][k] = fmaf(mv.y, pk, acc[1][k]);
                acc[2][k] = fmaf(mv.z, pk, acc[2][k]);
                acc[3][k] = fmaf(mv.w, pk, acc[3][k]);
            }
        }
    }
    if (ty > 0) {
#pragma unroll
        for (int t = 0; t < 4; ++t)
#pragma unroll
            for (int k = 0; k < R; ++k) red[ty - 1][tx][t * R + k] = acc[t][k];
    }
    __syncthreads();
    if (ty == 0 && j < m) {
#pragma unroll
        for (int t = 0; t < 4; ++t)
#pragma unroll
            for (int k = 0; k < R; ++k)
                atomicAdd(Qout + (long long)(j + t) * R + k,
                          acc[t][k] + red[0][tx][t * R + k] + red[1][tx][t * R + k] + red[2][tx][t * R + k]);
    }
}

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

// ---- finish: M̂ = P̂·Q̄ᵀ -> grad, E = M − M̂; persistent CTAs, one warp per row, Q̄ staged in shared memory ----------
template <int R>
__global__ void __launch_bounds__(256) psgd_finish_smem_kernel(float* __restrict__ E, float* __restrict__ grad,
                                                               const float* __restrict__ Phat, const float* __restrict__ Qbar,
                                                               int n, int m, int ldq) {
    extern __shared__ float Qs[];
    dnb::pdl_wait();
    stage_q(Qbar, Qs, m, R, ldq);
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
        float p[R];
#pragma unroll
        for (int k = 0; k < R; ++k) p[k] = __ldg(Phat + (long long)i * R + k);
        float* e = E + (long long)i * m;
        float* g = grad + (long long)i * m;
#pragma unroll 4
        for (int j = lane * 4; j < m; j += 128) {
            float4 mh = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int k = 0; k < R; ++k) {
                const float4 q = *reinterpret_cast<const float4*>(Qs + k * ldq + j);
                mh.x = fmaf(p[k], q.x, mh.x); mh.y = fmaf(p[k], q.y, mh.y);
                mh.z = fmaf(p[k], q.z, mh.z); mh.w = fmaf(p[k], q.w, mh.w);
            }
            const float4 mv = *reinterpret_cast<const float4*>(e + j);
            *reinterpret_cast<float4*>(e + j) = make_float4(mv.x - mh.x, mv.y - mh.y, mv.z - mh.z, mv.w - mh.w);
            __stcs(reinterpret_cast<float4*>(g + j), mh);
        }
    }
}

__global__ void __launch_bounds__(256) psgd_finish_kernel(float* __restrict__ E, float* __restrict__ grad,
                                                          const float* __restrict__ Phat, const float* __restrict__ Qbar,
                                                          int n, int m, int r) {
    dnb::pdl_wait();
    const long long total = (long long)n * m;
    const long long stride = (long long)gridDim.x * blockDim.x;
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

template <typename K, typename... Args>
cudaError_t launch_smem(K kernel, int blocks, size_t smem, cudaStream_t st, Args... args) {
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    return dnb_launch(kernel, dim3(blocks), dim3(256), smem, st, 0, args...);
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
    // fast paths: rows 16-byte aligned, Q fits shared memory, rank 1 / 2 / 4 / 8
    const bool vec = (d->m % 4) == 0 && ((reinterpret_cast<uintptr_t>(d->E) | reinterpret_cast<uintptr_t>(d->grad)) & 15) == 0;
    const int ldq = (d->m + 3) / 4 * 4 + 4;
    const size_t qbytes = (size_t)d->r * ldq * sizeof(float);
    const bool fast = vec && qbytes <= 96 * 1024 && (d->r == 1 || d->r == 2 || d->r == 4 || d->r == 8);
#define DNB_PSGD_R(KERNEL, ...)                                                                                       \
    do {                                                                                                              \
        switch (d->r) {                                                                                               \
            case 1: DNB_CUDA_CHECK(launch_smem(KERNEL<1>, blocks, qbytes, st, __VA_ARGS__)); break;                   \
            case 2: DNB_CUDA_CHECK(launch_smem(KERNEL<2>, blocks, qbytes, st, __VA_ARGS__)); break;                   \
            case 4: DNB_CUDA_CHECK(launch_smem(KERNEL<4>, blocks, qbytes, st, __VA_ARGS__)); break;                   \
            default: DNB_CUDA_CHECK(launch_smem(KERNEL<8>, blocks, qbytes, st, __VA_ARGS__)); break;                  \
        }                                                                                                             \
    } while (0)
    if (d->pass == 0) {
        DNB_REQUIRE(d->grad && d->Q && d->P, "pass P needs grad, Q and P");
        if (fast) {
            int blocks = (d->n + 7) / 8;
            if (blocks > sms * 3) blocks = sms * 3;      // persistent: every CTA stages Q once and walks many rows
            DNB_PSGD_R(psgd_p_smem_kernel, (const float*)d->grad, d->E, (const float*)d->Q, d->P, d->n, d->m, ldq);
        } else {
            int blocks = (d->n + 7) / 8;
            if (blocks > sms * 8) blocks = sms * 8;
            DNB_CUDA_CHECK(dnb_launch(psgd_p_kernel, dim3(blocks), dim3(256), 0, st, 0, (const float*)d->grad, d->E,
                                      (const float*)d->Q, d->P, d->n, d->m, d->r));
        }
    } else if (d->pass == 1) {
        DNB_REQUIRE(d->P && d->Q, "pass Q needs P (orthonormalised) and the Q output");
        if (vec && (d->r == 1 || d->r == 2 || d->r == 4)) {
            dim3 grid((d->m / 4 + 63) / 64, (d->n + kQRows - 1) / kQRows);
            switch (d->r) {
                case 1: DNB_CUDA_CHECK(dnb_launch(psgd_q_vec_kernel<1>, grid, dim3(256), 0, st, 0, (const float*)d->E, (const float*)d->P, d->Q, d->n, d->m)); break;
                case 2: DNB_CUDA_CHECK(dnb_launch(psgd_q_vec_kernel<2>, grid, dim3(256), 0, st, 0, (const float*)d->E, (const float*)d->P, d->Q, d->n, d->m)); break;
                default: DNB_CUDA_CHECK(dnb_launch(psgd_q_vec_kernel<4>, grid, dim3(256), 0, st, 0, (const float*)d->E, (const float*)d->P, d->Q, d->n, d->m)); break;
            }
        } else {
            dim3 grid((d->m + 255) / 256, (d->n + kQRows - 1) / kQRows);
            DNB_CUDA_CHECK(dnb_launch(psgd_q_kernel, grid, dim3(256), 0, st, 0, (const float*)d->E, (const float*)d->P, d->Q,
                                      d->n, d->m, d->r));
        }
    } else if (d->pass == 2) {
        DNB_REQUIRE(d->grad && d->Q && d->P, "finish needs grad, P and Q");
        if (fast) {
            int blocks = (d->n + 7) / 8;
            if (blocks > sms * 3) blocks = sms * 3;
            DNB_PSGD_R(psgd_finish_smem_kernel, d->E, d->grad, (const float*)d->P, (const float*)d->Q, d->n, d->m, ldq);
        } else {
            long long work = (long long)d->n * d->m;
            int blocks = (int)((work + 255) / 256);
            if (blocks > sms * 16) blocks = sms * 16;
            if (blocks < 1) blocks = 1;
            DNB_CUDA_CHECK(dnb_launch(psgd_finish_kernel, dim3(blocks), dim3(256), 0, st, 0, d->E, d->grad, (const float*)d->P,
                                      (const float*)d->Q, d->n, d->m, d->r));
        }
    } else {
        DNB_REQUIRE(false, "pass must be 0 (P), 1 (Q) or 2 (finish)");
    }
#undef DNB_PSGD_R
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
