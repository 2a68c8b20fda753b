// K8 — classifier head of the ICA-LSTM, forward and backward.
//
// Reference: comps/icalstm/models.py:95-103 —
//   Dropout(.25) -> Linear(N0,N1) -> BatchNorm1d(N1) -> ReLU -> Linear(N1,N2) -> ReLU -> Linear(N2,N3)
// with N0 = hidden (384), N1 = 256, N2 = 64, N3 = classes, batch 16-64.  Everything here is
// latency-bound small-matrix work (M = batch), far below a tcgen05 tile, and it sits on the
// serial chain LSTM -> head -> loss -> head' -> LSTM'.  So the layers are spread over many CTAs
// with warp-level dot products (lanes split K with 16-byte bf16 weight loads, shuffle reduce) and
// the BatchNorm statistics — per feature over the batch — are computed by the warp that owns the
// feature, with no extra pass:
//   forward  : [16 CTAs] Linear1 + BN + ReLU      ->  [1 CTA] Linear2 + ReLU + Linear3
//   backward : [1 CTA] layers 3, 2 and BN         ->  [38 CTAs] dW1 (feature tiles) and dx (column tiles)
// Weights are read as bf16 (the parameter shadow); accumulation, saved activations and all
// gradients are fp32.  Parameter gradients are accumulated (+=) when `accumulate` is set.
#include "common.cuh"
#include "../include/dinunet_b200.h"

#include <mutex>

void dnb_count_launch(int n);

namespace {

constexpr int BT = 16;          // batch rows handled per register tile
constexpr int FPW = 4;          // features per warp
constexpr int kMaxB = 64;

struct HeadFwdArgs {
    const float* x; const float* drop;
    const __nv_bfloat16 *W1, *W2, *W3;
    const float *b1, *b2, *b3, *bn_w, *bn_b;
    float *run_mean, *run_var;
    float momentum, eps; int training;
    float *a0, *xhat, *invstd, *a1, *a2, *logits;
    int B, N0, N1, N2, N3;
};

struct HeadBwdArgs {
    const float* dlogits;
    const float *a0, *drop, *xhat, *invstd, *a1, *a2;
    const __nv_bfloat16 *W1, *W2, *W3;
    const float* bn_w;
    float *dW1, *db1, *dbn_w, *dbn_b, *dW2, *db2, *dW3, *db3, *dx;
    float* dz1;                              // [B,N1] scratch between the two backward launches
    int training, accumulate;
    int B, N0, N1, N2, N3;
};

__device__ __forceinline__ void put(float* p, float v, int acc) { *p = acc ? (*p + v) : v; }

__device__ __forceinline__ void bf16x8_to_f32(const uint4& raw, float (&w)[8]) {
    const uint32_t u[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        w[2 * i] = __uint_as_float(u[i] << 16);
        w[2 * i + 1] = __uint_as_float(u[i] & 0xFFFF0000u);
    }
}

// One warp: dot products of FPW weight rows (bf16, row pitch ldw) with BT activation rows held in
// shared memory (fp32, row pitch K).  Lanes split K in 8-element chunks; result summed over the
// warp, every lane ends up with all FPW x BT sums.
__device__ __forceinline__ void warp_dots(const __nv_bfloat16* __restrict__ W, int ldw, int nfeat, const float* xs,
                                          int K, int b0, int B, float (&acc)[FPW][BT]) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int f = 0; f < FPW; ++f)
#pragma unroll
        for (int b = 0; b < BT; ++b) acc[f][b] = 0.f;
    for (int c = lane; c < K / 8; c += 32) {
        float w[FPW][8];
#pragma unroll
        for (int f = 0; f < FPW; ++f) {
            if (f < nfeat) bf16x8_to_f32(*reinterpret_cast<const uint4*>(W + (size_t)f * ldw + c * 8), w[f]);
            else {
#pragma unroll
                for (int i = 0; i < 8; ++i) w[f][i] = 0.f;
            }
        }
#pragma unroll
        for (int b = 0; b < BT; ++b) {
            if (b0 + b < B) {
                const float4 x0 = *reinterpret_cast<const float4*>(xs + (size_t)(b0 + b) * K + c * 8);
                const float4 x1 = *reinterpret_cast<const float4*>(xs + (size_t)(b0 + b) * K + c * 8 + 4);
#pragma unroll
                for (int f = 0; f < FPW; ++f)
                    acc[f][b] += w[f][0] * x0.x + w[f][1] * x0.y + w[f][2] * x0.z + w[f][3] * x0.w + w[f][4] * x1.x +
                                 w[f][5] * x1.y + w[f][6] * x1.z + w[f][7] * x1.w;
            }
        }
    }
#pragma unroll
    for (int f = 0; f < FPW; ++f)
#pragma unroll
        for (int b = 0; b < BT; ++b) acc[f][b] = dnb::warp_sum(acc[f][b]);
}

// ------------------------------------------------------------------ forward, launch 1: Linear1 + BN + ReLU
constexpr int kL1Warps = 4;
__global__ void __launch_bounds__(kL1Warps * 32) head_l1_kernel(const HeadFwdArgs a) {
    extern __shared__ float sm[];
    float* xs = sm;                                   // [B][N0] dropped input
    float* zs = xs + a.B * a.N0;                      // [warps][FPW][kMaxB] pre-BN values
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, B = a.B;
    {
        // vectorised, unrolled staging: independent 16-byte loads in flight instead of a dependent chain
        const int n4 = B * a.N0 / 4;
        const float4* x4 = reinterpret_cast<const float4*>(a.x);
        const float4* d4 = reinterpret_cast<const float4*>(a.drop);
#pragma unroll 4
        for (int i = tid; i < n4; i += kL1Warps * 32) {
            float4 v = x4[i];
            if (d4) {
                const float4 d = d4[i];
                v.x *= d.x; v.y *= d.y; v.z *= d.z; v.w *= d.w;
            }
            reinterpret_cast<float4*>(xs)[i] = v;
            if (a.a0 && blockIdx.x == 0) reinterpret_cast<float4*>(a.a0)[i] = v;
        }
    }
    __syncthreads();
    const int j0 = (blockIdx.x * kL1Warps + warp) * FPW;
    if (j0 >= a.N1) return;
    const int nfeat = min(FPW, a.N1 - j0);
    float* z = zs + warp * FPW * kMaxB;
    for (int b0 = 0; b0 < B; b0 += BT) {
        float acc[FPW][BT];
        warp_dots(a.W1 + (size_t)j0 * a.N0, a.N0, nfeat, xs, a.N0, b0, B, acc);
#pragma unroll
        for (int f = 0; f < FPW; ++f)
#pragma unroll
            for (int b = 0; b < BT; ++b)
                if (lane == b && b0 + b < B) z[f * kMaxB + b0 + b] = acc[f][b];
    }
    __syncwarp();
    for (int f = 0; f < nfeat; ++f) {
        const int j = j0 + f;
        const float bias = a.b1[j];
        float v0 = lane < B ? z[f * kMaxB + lane] + bias : 0.f;
        float v1 = lane + 32 < B ? z[f * kMaxB + lane + 32] + bias : 0.f;
        const float mean = dnb::warp_sum(v0 + v1) / (float)B;
        const float d0 = lane < B ? v0 - mean : 0.f, d1 = lane + 32 < B ? v1 - mean : 0.f;
        const float var = dnb::warp_sum(d0 * d0 + d1 * d1) / (float)B;
        float use_mean = mean, use_var = var;
        if (a.training) {
            if (a.run_mean && lane == 0) {
                a.run_mean[j] = (1.f - a.momentum) * a.run_mean[j] + a.momentum * mean;
                const float unbiased = B > 1 ? var * (float)B / (float)(B - 1) : var;
                a.run_var[j] = (1.f - a.momentum) * a.run_var[j] + a.momentum * unbiased;
            }
        } else {
            use_mean = a.run_mean[j];
            use_var = a.run_var[j];
        }
        const float istd = rsqrtf(use_var + a.eps);
        if (lane == 0) a.invstd[j] = istd;
        const float g = a.bn_w[j], be = a.bn_b[j];
        if (lane < B) {
            const float xh = (v0 - use_mean) * istd;
            a.xhat[lane * a.N1 + j] = xh;
            a.a1[lane * a.N1 + j] = fmaxf(xh * g + be, 0.f);
        }
        if (lane + 32 < B) {
            const float xh = (v1 - use_mean) * istd;
            a.xhat[(lane + 32) * a.N1 + j] = xh;
            a.a1[(lane + 32) * a.N1 + j] = fmaxf(xh * g + be, 0.f);
        }
    }
}

// ------------------------------------------------------------------ forward, launch 2: Linear2 + ReLU + Linear3
constexpr int kL23Threads = 256;
__global__ void __launch_bounds__(kL23Threads) head_l23_kernel(const HeadFwdArgs a) {
    extern __shared__ float sm[];
    float* s1 = sm;                       // [B][N1]
    float* s2 = s1 + a.B * a.N1;          // [B][N2]
    float* s3 = s2 + a.B * a.N2;          // [N3][N2] last-layer weights
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, B = a.B;
    {
        const int n4 = B * a.N1 / 4;
        const float4* src = reinterpret_cast<const float4*>(a.a1);
#pragma unroll 4
        for (int i = tid; i < n4; i += kL23Threads) reinterpret_cast<float4*>(s1)[i] = src[i];
#pragma unroll 2
        for (int i = tid; i < a.N3 * a.N2; i += kL23Threads) s3[i] = __bfloat162float(a.W3[i]);
    }
    __syncthreads();
    for (int j0 = warp * FPW; j0 < a.N2; j0 += (kL23Threads / 32) * FPW) {
        const int nfeat = min(FPW, a.N2 - j0);
        for (int b0 = 0; b0 < B; b0 += BT) {
            float acc[FPW][BT];
            warp_dots(a.W2 + (size_t)j0 * a.N1, a.N1, nfeat, s1, a.N1, b0, B, acc);
#pragma unroll
            for (int f = 0; f < FPW; ++f)
#pragma unroll
                for (int b = 0; b < BT; ++b)
                    if (lane == b && f < nfeat && b0 + b < B) {
                        const float v = fmaxf(acc[f][b] + a.b2[j0 + f], 0.f);
                        s2[(b0 + b) * a.N2 + j0 + f] = v;
                        a.a2[(b0 + b) * a.N2 + j0 + f] = v;
                    }
        }
    }
    __syncthreads();
    for (int idx = tid; idx < B * a.N3; idx += kL23Threads) {
        const int b = idx / a.N3, c = idx % a.N3;
        const float* w = s3 + c * a.N2;
        float acc = a.b3[c];
#pragma unroll 8
        for (int k = 0; k < a.N2; ++k) acc += w[k] * s2[b * a.N2 + k];
        a.logits[idx] = acc;
    }
}

// ------------------------------------------------------------------ backward, launch 1: layers 3, 2 and BN
constexpr int kB23Threads = 256;
template <int BTT>
__global__ void __launch_bounds__(kB23Threads, 1) head_bwd_l23_kernel(const HeadBwdArgs a) {
    extern __shared__ float sm[];
    const int B = a.B, N1 = a.N1, N2 = a.N2, N3 = a.N3, tid = threadIdx.x, acc_mode = a.accumulate;
    float* sdl = sm;                       // [B][N3]
    float* sa2 = sdl + B * N3;             // [B][N2]
    float* sdz2 = sa2 + B * N2;            // [B][N2]
    float* sw3 = sdz2 + B * N2;            // [N3][N2]
    for (int i = tid; i < B * N3; i += kB23Threads) sdl[i] = a.dlogits[i];
#pragma unroll 4
    for (int i = tid; i < B * N2; i += kB23Threads) sa2[i] = a.a2[i];
    for (int i = tid; i < N3 * N2; i += kB23Threads) sw3[i] = __bfloat162float(a.W3[i]);
    __syncthreads();
    for (int idx = tid; idx < N3 * N2; idx += kB23Threads) {
        const int c = idx / N2, k = idx % N2;
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += sdl[b * N3 + c] * sa2[b * N2 + k];
        put(a.dW3 + idx, s, acc_mode);
    }
    for (int c = tid; c < N3; c += kB23Threads) {
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += sdl[b * N3 + c];
        put(a.db3 + c, s, acc_mode);
    }
    for (int idx = tid; idx < B * N2; idx += kB23Threads) {
        const int b = idx / N2, k = idx % N2;
        float s = 0.f;
        for (int c = 0; c < N3; ++c) s += sdl[b * N3 + c] * sw3[c * N2 + k];
        sdz2[idx] = sa2[idx] > 0.f ? s : 0.f;
    }
    __syncthreads();
    for (int k = tid; k < N2; k += kB23Threads) {
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += sdz2[b * N2 + k];
        put(a.db2 + k, s, acc_mode);
    }
    // thread = feature j of N1: all batch rows in registers -> BN backward needs no reduction
    for (int j = tid; j < N1; j += kB23Threads) {
        float a1r[BTT], xh[BTT], da1[BTT];
#pragma unroll
        for (int b = 0; b < BTT; ++b) {
            a1r[b] = b < B ? a.a1[b * N1 + j] : 0.f;
            xh[b] = b < B ? a.xhat[b * N1 + j] : 0.f;
            da1[b] = 0.f;
        }
        for (int k0 = 0; k0 < N2; k0 += 8) {
            float w8[8], old8[8];
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {              // 8 independent loads per trip
                const int k = k0 + kk;
                w8[kk] = k < N2 ? __bfloat162float(a.W2[(size_t)k * N1 + j]) : 0.f;
                old8[kk] = (acc_mode && k < N2) ? a.dW2[(size_t)k * N1 + j] : 0.f;
            }
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                const int k = k0 + kk;
                if (k < N2) {
                    float s = 0.f;
#pragma unroll
                    for (int b = 0; b < BTT; ++b)
                        if (b < B) {
                            const float dz = sdz2[b * N2 + k];
                            s += dz * a1r[b];
                            da1[b] += dz * w8[kk];
                        }
                    a.dW2[(size_t)k * N1 + j] = old8[kk] + s;
                }
            }
        }
        float dgamma = 0.f, dbeta = 0.f;
#pragma unroll
        for (int b = 0; b < BTT; ++b)
            if (b < B) {
                const float dy = a1r[b] > 0.f ? da1[b] : 0.f;
                da1[b] = dy;
                dgamma += dy * xh[b];
                dbeta += dy;
            }
        put(a.dbn_w + j, dgamma, acc_mode);
        put(a.dbn_b + j, dbeta, acc_mode);
        const float gi = a.bn_w[j] * a.invstd[j];
        const float invB = 1.f / (float)B;
        float db1 = 0.f;
#pragma unroll
        for (int b = 0; b < BTT; ++b)
            if (b < B) {
                const float dz = a.training ? gi * (da1[b] - dbeta * invB - xh[b] * dgamma * invB) : gi * da1[b];
                a.dz1[b * N1 + j] = dz;
                db1 += dz;
            }
        put(a.db1 + j, db1, acc_mode);
    }
}

// ------------------------------------------------------------------ backward, launch 2: dW1 and dx
constexpr int kB1Threads = 256;
constexpr int kJTile = 8;        // features per dW1 block
constexpr int kKTile = 64;       // input columns per dx block
__global__ void __launch_bounds__(kB1Threads) head_bwd_l1_kernel(const HeadBwdArgs a, int n_jblocks) {
    extern __shared__ float sm[];
    const int B = a.B, N0 = a.N0, N1 = a.N1, tid = threadIdx.x;
    if ((int)blockIdx.x < n_jblocks) {
        // dW1[j][k] (+)= sum_b dz1[b][j] * a0[b][k] for 8 features, all k (coalesced rows)
        float* a0s = sm;                        // [B][N0]
        float* dzt = a0s + B * N0;              // [kJTile][B]
        const int j0 = blockIdx.x * kJTile;
        {
            const int n4 = B * N0 / 4;
            const float4* src = reinterpret_cast<const float4*>(a.a0);
#pragma unroll 4
            for (int i = tid; i < n4; i += kB1Threads) reinterpret_cast<float4*>(a0s)[i] = src[i];
        }
        for (int i = tid; i < kJTile * B; i += kB1Threads) {
            const int jj = i / B, b = i % B;
            dzt[i] = (j0 + jj < N1) ? a.dz1[b * N1 + j0 + jj] : 0.f;
        }
        __syncthreads();
        for (int k = tid; k < N0; k += kB1Threads) {
            float s[kJTile];
#pragma unroll
            for (int jj = 0; jj < kJTile; ++jj) s[jj] = 0.f;
            for (int b = 0; b < B; ++b) {
                const float xv = a0s[b * N0 + k];
#pragma unroll
                for (int jj = 0; jj < kJTile; ++jj) s[jj] += dzt[jj * B + b] * xv;
            }
            float old[kJTile];
#pragma unroll
            for (int jj = 0; jj < kJTile; ++jj)
                old[jj] = (a.accumulate && j0 + jj < N1) ? a.dW1[(size_t)(j0 + jj) * N0 + k] : 0.f;
#pragma unroll
            for (int jj = 0; jj < kJTile; ++jj)
                if (j0 + jj < N1) a.dW1[(size_t)(j0 + jj) * N0 + k] = old[jj] + s[jj];
        }
    } else {
        // dx[b][k] = drop * sum_j dz1[b][j] * W1[j][k] for 64 columns; 4 thread groups split j
        float* dzs = sm;                        // [N1][B]  (feature-major: one float4 run per feature)
        float* red = dzs + N1 * B;              // [4][kKTile][B]
        const int k0 = (blockIdx.x - n_jblocks) * kKTile;
#pragma unroll 4
        for (int i = tid; i < N1 * B; i += kB1Threads) {
            const int b = i / N1, j = i % N1;   // coalesced read of dz1[b][j]
            dzs[j * B + b] = a.dz1[i];
        }
        __syncthreads();
        const int kk = tid % kKTile, jp = tid / kKTile;
        const int k = k0 + kk;
        const int jper = (N1 + 3) / 4;
        for (int b0 = 0; b0 < B; b0 += BT) {
            float acc[BT];
#pragma unroll
            for (int b = 0; b < BT; ++b) acc[b] = 0.f;
            if (k < N0) {
                const int jend = min(N1, (jp + 1) * jper);
                for (int j = jp * jper; j < jend; j += 8) {
                    float w8[8];
#pragma unroll
                    for (int jj = 0; jj < 8; ++jj)           // 8 weight loads in flight per trip
                        w8[jj] = (j + jj < jend) ? __bfloat162float(a.W1[(size_t)(j + jj) * N0 + k]) : 0.f;
#pragma unroll
                    for (int jj = 0; jj < 8; ++jj)
                        if (j + jj < jend) {
#pragma unroll
                            for (int b = 0; b < BT; ++b)
                                if (b0 + b < B) acc[b] += dzs[(j + jj) * B + b0 + b] * w8[jj];
                        }
                }
            }
#pragma unroll
            for (int b = 0; b < BT; ++b)
                if (b0 + b < B) red[(jp * kKTile + kk) * B + b0 + b] = acc[b];
        }
        __syncthreads();
        for (int i = tid; i < kKTile * B; i += kB1Threads) {
            const int b = i / kKTile, c = i % kKTile;
            if (k0 + c < N0) {
                float s = 0.f;
#pragma unroll
                for (int p = 0; p < 4; ++p) s += red[(p * kKTile + c) * B + b];
                const size_t o = (size_t)b * N0 + k0 + c;
                a.dx[o] = s * (a.drop ? a.drop[o] : 1.f);
            }
        }
    }
}

}  // namespace

DNB_EXPORT int dnb_head_forward(const DnbHeadFwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_head_forward: null descriptor");
    DNB_REQUIRE(d->B >= 1 && d->B <= kMaxB, "head: batch must be in [1,%d] (got %d)", kMaxB, d->B);
    DNB_REQUIRE(d->N0 % 8 == 0 && d->N1 % 8 == 0, "head: N0 and N1 must be multiples of 8");
    DNB_REQUIRE(!d->training || d->B > 1, "head: BatchNorm in training mode needs more than one sample");
    HeadFwdArgs a;
    a.x = d->x; a.drop = d->drop;
    a.W1 = (const __nv_bfloat16*)d->W1; a.W2 = (const __nv_bfloat16*)d->W2; a.W3 = (const __nv_bfloat16*)d->W3;
    a.b1 = d->b1; a.b2 = d->b2; a.b3 = d->b3; a.bn_w = d->bn_w; a.bn_b = d->bn_b;
    a.run_mean = d->run_mean; a.run_var = d->run_var; a.momentum = d->momentum; a.eps = d->eps; a.training = d->training;
    a.a0 = d->a0; a.xhat = d->xhat; a.invstd = d->invstd; a.a1 = d->a1; a.a2 = d->a2; a.logits = d->logits;
    a.B = d->B; a.N0 = d->N0; a.N1 = d->N1; a.N2 = d->N2; a.N3 = d->N3;
    DNB_REQUIRE(a.x && a.W1 && a.W2 && a.W3 && a.xhat && a.invstd && a.a1 && a.a2 && a.logits, "head fwd: null buffer");
    DNB_REQUIRE(a.training || (a.run_mean && a.run_var), "head fwd: eval mode needs running statistics");
    {
        const size_t smem = ((size_t)d->B * d->N0 + kL1Warps * FPW * kMaxB) * sizeof(float);
        DNB_CUDA_CHECK(cudaFuncSetAttribute(head_l1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const int blocks = dnb_ceil_div(d->N1, kL1Warps * FPW);
        head_l1_kernel<<<blocks, kL1Warps * 32, smem, st>>>(a);
        DNB_LAUNCH_CHECK();
    }
    {
        const size_t smem = ((size_t)d->B * (d->N1 + d->N2) + (size_t)d->N3 * d->N2) * sizeof(float);
        DNB_CUDA_CHECK(cudaFuncSetAttribute(head_l23_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        head_l23_kernel<<<1, kL23Threads, smem, st>>>(a);
        DNB_LAUNCH_CHECK();
    }
    dnb_count_launch(2);
    return 0;
}

DNB_EXPORT int dnb_head_backward(const DnbHeadBwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_head_backward: null descriptor");
    DNB_REQUIRE(d->B >= 1 && d->B <= kMaxB, "head: batch must be in [1,%d] (got %d)", kMaxB, d->B);
    HeadBwdArgs a;
    a.dlogits = d->dlogits; a.a0 = d->a0; a.drop = d->drop; a.xhat = d->xhat; a.invstd = d->invstd; a.a1 = d->a1; a.a2 = d->a2;
    a.W1 = (const __nv_bfloat16*)d->W1; a.W2 = (const __nv_bfloat16*)d->W2; a.W3 = (const __nv_bfloat16*)d->W3;
    a.bn_w = d->bn_w;
    a.dW1 = d->dW1; a.db1 = d->db1; a.dbn_w = d->dbn_w; a.dbn_b = d->dbn_b; a.dW2 = d->dW2; a.db2 = d->db2;
    a.dW3 = d->dW3; a.db3 = d->db3; a.dx = d->dx; a.dz1 = d->dz1;
    a.training = d->training; a.accumulate = d->accumulate;
    a.B = d->B; a.N0 = d->N0; a.N1 = d->N1; a.N2 = d->N2; a.N3 = d->N3;
    DNB_REQUIRE(a.dlogits && a.a0 && a.xhat && a.invstd && a.a1 && a.a2 && a.dW1 && a.db1 && a.dbn_w && a.dbn_b &&
                    a.dW2 && a.db2 && a.dW3 && a.db3 && a.dx && a.dz1,
                "head bwd: null buffer");
    {
        const size_t smem = ((size_t)d->B * (d->N3 + 2 * d->N2) + (size_t)d->N3 * d->N2) * sizeof(float);
        if (d->B <= 16) head_bwd_l23_kernel<16><<<1, kB23Threads, smem, st>>>(a);
        else if (d->B <= 32) head_bwd_l23_kernel<32><<<1, kB23Threads, smem, st>>>(a);
        else head_bwd_l23_kernel<64><<<1, kB23Threads, smem, st>>>(a);
        DNB_LAUNCH_CHECK();
    }
    {
        const int n_j = dnb_ceil_div(d->N1, kJTile), n_k = dnb_ceil_div(d->N0, kKTile);
        const size_t smem_j = ((size_t)d->B * d->N0 + kJTile * d->B) * sizeof(float);
        const size_t smem_k = ((size_t)d->N1 * d->B + 4 * kKTile * d->B) * sizeof(float);
        const size_t smem = smem_j > smem_k ? smem_j : smem_k;
        DNB_CUDA_CHECK(cudaFuncSetAttribute(head_bwd_l1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        head_bwd_l1_kernel<<<n_j + n_k, kB1Threads, smem, st>>>(a, n_j);
        DNB_LAUNCH_CHECK();
    }
    dnb_count_launch(2);
    return 0;
}

// ------------------------------------------------------------------ allocation-free backward
// dnb_head_backward needs a [B, N1] fp32 scratch (`dz1`) to hand the BatchNorm-input gradient from
// its first launch to its second, and leaves providing it to the caller.  Some callers have no way
// to provide one: a host process whose device memory is carved up front by another runtime or pool
// and that is not allowed to call cudaMalloc behind its back, or a service with a fixed memory
// budget in which an extra allocation per call (which may synchronise the device, and may fail long
// after start-up) is unacceptable.  For them the scratch lives in the library image itself.
//
// The array below is file-scope __device__ memory: every GPU that loads this library holds its
// 128 KiB until the process exits, used or not.  In host code the NAME of the array is only a
// host-side handle, not a device address — handing it to a kernel would fault (or, where the GPU
// can reach host memory, silently touch the wrong bytes) — so the device address is looked up with
// cudaGetSymbolAddress, once per device, while that device is current, and cached.
namespace {
constexpr int kStaticMaxB = 64;       // batch rows the built-in scratch can hold
constexpr int kStaticMaxN1 = 512;     // hidden width (BatchNorm features) it can hold
constexpr int kStaticMaxDevices = 64;
static __device__ __align__(256) float g_head_dz1_scratch[kStaticMaxB * kStaticMaxN1];   // 131072 bytes

float* g_scratch_dev_ptr[kStaticMaxDevices] = {nullptr};
std::mutex g_scratch_mutex;

// Device address of the built-in scratch on the CURRENT device (looked up on first use).
int static_scratch_for_current_device(float** out) {
    int dev = -1;
    DNB_CUDA_CHECK(cudaGetDevice(&dev));
    DNB_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "device ordinal %d outside [0,%d)", dev, kStaticMaxDevices);
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (g_scratch_dev_ptr[dev] == nullptr) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_head_dz1_scratch);
        if (e != cudaSuccess || p == nullptr) {
            dnb_set_error("cudaGetSymbolAddress(built-in head scratch) failed on device %d: %s", dev,
                          cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : -1;
        }
        g_scratch_dev_ptr[dev] = static_cast<float*>(p);
    }
    *out = g_scratch_dev_ptr[dev];
    return 0;
}
}  // namespace

// Same computation and same outputs as dnb_head_backward, but the [B, N1] scratch is the
// library's built-in array; `desc->dz1` is ignored.  Performs no cudaMalloc / cudaFree and calls no
// allocating routine.  Limits: B <= 64 and N1 <= 512 (larger shapes are refused with an error —
// use dnb_head_backward with your own scratch instead).
//
//  * The first call on each device resolves the array's device address (cudaGetSymbolAddress).
//    Make that first call during start-up, before any stream capture is begun on that device.
//  * All calls on a device share the one array and it keeps whatever the previous call left in
//    it.  Nothing in it has to start at zero: the first launch writes every element the second
//    launch reads (rows < B, columns < N1), so stale contents are never consumed.
//  * Because the array is shared, two calls whose kernels can overlap on the same GPU (different
//    streams, or different host threads) race on it.  Issue them on one stream, or serialise them.
DNB_EXPORT int dnb_head_backward_noalloc(const DnbHeadBwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_head_backward_noalloc: null descriptor");
    DNB_REQUIRE(d->B >= 1 && d->B <= kStaticMaxB,
                "dnb_head_backward_noalloc: batch %d exceeds the built-in scratch (max %d rows); "
                "use dnb_head_backward with caller-provided dz1", d->B, kStaticMaxB);
    DNB_REQUIRE(d->N1 >= 1 && d->N1 <= kStaticMaxN1,
                "dnb_head_backward_noalloc: N1 = %d exceeds the built-in scratch (max %d features); "
                "use dnb_head_backward with caller-provided dz1", d->N1, kStaticMaxN1);
    float* scratch = nullptr;
    if (int rc = static_scratch_for_current_device(&scratch)) return rc;

    HeadBwdArgs a;
    a.dlogits = d->dlogits; a.a0 = d->a0; a.drop = d->drop; a.xhat = d->xhat; a.invstd = d->invstd; a.a1 = d->a1; a.a2 = d->a2;
    a.W1 = (const __nv_bfloat16*)d->W1; a.W2 = (const __nv_bfloat16*)d->W2; a.W3 = (const __nv_bfloat16*)d->W3;
    a.bn_w = d->bn_w;
    a.dW1 = d->dW1; a.db1 = d->db1; a.dbn_w = d->dbn_w; a.dbn_b = d->dbn_b; a.dW2 = d->dW2; a.db2 = d->db2;
    a.dW3 = d->dW3; a.db3 = d->db3; a.dx = d->dx;
    a.dz1 = scratch;                                   // device address, never the host-side symbol
    a.training = d->training; a.accumulate = d->accumulate;
    a.B = d->B; a.N0 = d->N0; a.N1 = d->N1; a.N2 = d->N2; a.N3 = d->N3;
    DNB_REQUIRE(a.dlogits && a.a0 && a.xhat && a.invstd && a.a1 && a.a2 && a.dW1 && a.db1 && a.dbn_w && a.dbn_b &&
                    a.dW2 && a.db2 && a.dW3 && a.db3 && a.dx,
                "dnb_head_backward_noalloc: null buffer");
    {
        const size_t smem = ((size_t)d->B * (d->N3 + 2 * d->N2) + (size_t)d->N3 * d->N2) * sizeof(float);
        if (d->B <= 16) head_bwd_l23_kernel<16><<<1, kB23Threads, smem, st>>>(a);
        else if (d->B <= 32) head_bwd_l23_kernel<32><<<1, kB23Threads, smem, st>>>(a);
        else head_bwd_l23_kernel<64><<<1, kB23Threads, smem, st>>>(a);
        DNB_LAUNCH_CHECK();
    }
    {
        const int n_j = dnb_ceil_div(d->N1, kJTile), n_k = dnb_ceil_div(d->N0, kKTile);
        const size_t smem_j = ((size_t)d->B * d->N0 + kJTile * d->B) * sizeof(float);
        const size_t smem_k = ((size_t)d->N1 * d->B + 4 * kKTile * d->B) * sizeof(float);
        const size_t smem = smem_j > smem_k ? smem_j : smem_k;
        DNB_CUDA_CHECK(cudaFuncSetAttribute(head_bwd_l1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        head_bwd_l1_kernel<<<n_j + n_k, kB1Threads, smem, st>>>(a, n_j);
        DNB_LAUNCH_CHECK();
    }
    dnb_count_launch(2);
    return 0;
}
