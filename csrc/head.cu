// K8 — classifier head of the ICA-LSTM, forward and backward, one launch each.
//
// Reference: comps/icalstm/models.py:95-103 —
//   Dropout(.25) -> Linear(N0,N1) -> BatchNorm1d(N1) -> ReLU -> Linear(N1,N2) -> ReLU -> Linear(N2,N3)
// with N0 = hidden (384), N1 = 256, N2 = 64, N3 = classes.  The batch is tiny (16-64 rows), so
// the whole head is latency-bound: a single CTA keeps all activations in shared memory / registers
// and walks the three layers without touching HBM in between.  BatchNorm statistics are per
// feature over the batch, and each thread owns one feature for all batch rows, so mean/var and
// the BN backward need no cross-thread reduction.  Weights are read as bf16 (the parameter
// shadow), accumulation and all outputs are fp32.
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

constexpr int kHeadThreads = 256;

struct HeadFwdArgs {
    const float* x; const float* drop;       // [B,N0]; drop: per-element scale (0 or 1/(1-p)) or NULL
    const __nv_bfloat16 *W1, *W2, *W3;        // [N1,N0], [N2,N1], [N3,N2]
    const float *b1, *b2, *b3, *bn_w, *bn_b;
    float *run_mean, *run_var;
    float momentum, eps; int training;
    float *a0, *xhat, *invstd, *a1, *a2, *logits;   // saved / outputs
    int B, N0, N1, N2, N3;
};

struct HeadBwdArgs {
    const float* dlogits;                    // [B,N3]
    const float *a0, *drop, *xhat, *invstd, *a1, *a2;
    const __nv_bfloat16 *W1, *W2, *W3;
    const float* bn_w;
    float *dW1, *db1, *dbn_w, *dbn_b, *dW2, *db2, *dW3, *db3, *dx;
    int training, accumulate;
    int B, N0, N1, N2, N3;
};

__device__ __forceinline__ void put(float* p, float v, int acc) { *p = acc ? (*p + v) : v; }

template <int BT>
__global__ void __launch_bounds__(kHeadThreads, 1) head_fwd_kernel(const HeadFwdArgs a) {
    extern __shared__ float sm[];
    float* s0 = sm;                         // [B][N0]
    float* s1 = s0 + a.B * a.N0;            // [B][N1]
    float* s2 = s1 + a.B * a.N1;            // [B][N2]
    const int tid = threadIdx.x, B = a.B;

    for (int i = tid; i < B * a.N0; i += kHeadThreads) {
        const float v = a.x[i] * (a.drop ? a.drop[i] : 1.f);
        s0[i] = v;
        if (a.a0) a.a0[i] = v;
    }
    __syncthreads();

    // ---- Linear 1 + BatchNorm + ReLU: one feature per thread
    for (int j = tid; j < a.N1; j += kHeadThreads) {
        float acc[BT];
#pragma unroll
        for (int b = 0; b < BT; ++b) acc[b] = 0.f;
        const __nv_bfloat16* w = a.W1 + (size_t)j * a.N0;
        for (int k = 0; k < a.N0; k += 8) {
            const uint4 raw = *reinterpret_cast<const uint4*>(w + k);
            const uint32_t ww[4] = {raw.x, raw.y, raw.z, raw.w};
            float wf[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                wf[2 * i] = __uint_as_float(ww[i] << 16);
                wf[2 * i + 1] = __uint_as_float(ww[i] & 0xFFFF0000u);
            }
#pragma unroll
            for (int b = 0; b < BT; ++b) {
                if (b < B) {
                    const float4 x0 = *reinterpret_cast<const float4*>(s0 + b * a.N0 + k);
                    const float4 x1 = *reinterpret_cast<const float4*>(s0 + b * a.N0 + k + 4);
                    acc[b] += wf[0] * x0.x + wf[1] * x0.y + wf[2] * x0.z + wf[3] * x0.w + wf[4] * x1.x +
                              wf[5] * x1.y + wf[6] * x1.z + wf[7] * x1.w;
                }
            }
        }
        const float bias = a.b1[j];
        float mean = 0.f, var = 0.f;
#pragma unroll
        for (int b = 0; b < BT; ++b)
            if (b < B) { acc[b] += bias; mean += acc[b]; }
        mean /= (float)B;
#pragma unroll
        for (int b = 0; b < BT; ++b)
            if (b < B) var += (acc[b] - mean) * (acc[b] - mean);
        var /= (float)B;
        float use_mean = mean, use_var = var;
        if (a.training) {
            if (a.run_mean) {
                a.run_mean[j] = (1.f - a.momentum) * a.run_mean[j] + a.momentum * mean;
                const float unbiased = B > 1 ? var * (float)B / (float)(B - 1) : var;
                a.run_var[j] = (1.f - a.momentum) * a.run_var[j] + a.momentum * unbiased;
            }
        } else {
            use_mean = a.run_mean[j];
            use_var = a.run_var[j];
        }
        const float istd = rsqrtf(use_var + a.eps);
        a.invstd[j] = istd;
        const float g = a.bn_w[j], be = a.bn_b[j];
#pragma unroll
        for (int b = 0; b < BT; ++b)
            if (b < B) {
                const float xh = (acc[b] - use_mean) * istd;
                const float y = fmaxf(xh * g + be, 0.f);
                a.xhat[b * a.N1 + j] = xh;
                a.a1[b * a.N1 + j] = y;
                s1[b * a.N1 + j] = y;
            }
    }
    __syncthreads();

    // ---- Linear 2 + ReLU: (batch row, feature) pairs spread over the threads
    for (int idx = tid; idx < B * a.N2; idx += kHeadThreads) {
        const int b = idx / a.N2, j = idx % a.N2;
        const __nv_bfloat16* w = a.W2 + (size_t)j * a.N1;
        float acc = a.b2[j];
        for (int k = 0; k < a.N1; k += 2) {
            const uint32_t raw = *reinterpret_cast<const uint32_t*>(w + k);
            acc += __uint_as_float(raw << 16) * s1[b * a.N1 + k] + __uint_as_float(raw & 0xFFFF0000u) * s1[b * a.N1 + k + 1];
        }
        acc = fmaxf(acc, 0.f);
        s2[idx] = acc;
        a.a2[idx] = acc;
    }
    __syncthreads();

    // ---- Linear 3
    for (int idx = tid; idx < B * a.N3; idx += kHeadThreads) {
        const int b = idx / a.N3, c = idx % a.N3;
        const __nv_bfloat16* w = a.W3 + (size_t)c * a.N2;
        float acc = a.b3[c];
        for (int k = 0; k < a.N2; ++k) acc += __bfloat162float(w[k]) * s2[b * a.N2 + k];
        a.logits[idx] = acc;
    }
}

template <int BT>
__global__ void __launch_bounds__(kHeadThreads, 1) head_bwd_kernel(const HeadBwdArgs a) {
    extern __shared__ float sm[];
    const int B = a.B, N0 = a.N0, N1 = a.N1, N2 = a.N2, N3 = a.N3, tid = threadIdx.x, acc_mode = a.accumulate;
    float* sdl = sm;                       // [B][N3]
    float* sa2 = sdl + B * N3;             // [B][N2]
    float* sdz2 = sa2 + B * N2;            // [B][N2]
    float* sdz1 = sdz2 + B * N2;           // [B][N1]
    float* sa0 = sdz1 + B * N1;            // [B][N0]

    for (int i = tid; i < B * N3; i += kHeadThreads) sdl[i] = a.dlogits[i];
    for (int i = tid; i < B * N2; i += kHeadThreads) sa2[i] = a.a2[i];
    for (int i = tid; i < B * N0; i += kHeadThreads) sa0[i] = a.a0[i];
    __syncthreads();

    // ---- layer 3
    for (int idx = tid; idx < N3 * N2; idx += kHeadThreads) {
        const int c = idx / N2, k = idx % N2;
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += sdl[b * N3 + c] * sa2[b * N2 + k];
        put(a.dW3 + idx, s, acc_mode);
    }
    for (int c = tid; c < N3; c += kHeadThreads) {
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += sdl[b * N3 + c];
        put(a.db3 + c, s, acc_mode);
    }
    for (int idx = tid; idx < B * N2; idx += kHeadThreads) {
        const int b = idx / N2, k = idx % N2;
        float s = 0.f;
        for (int c = 0; c < N3; ++c) s += sdl[b * N3 + c] * __bfloat162float(a.W3[c * N2 + k]);
        sdz2[idx] = sa2[idx] > 0.f ? s : 0.f;
    }
    __syncthreads();

    // ---- layer 2 + BatchNorm backward: thread = feature j of N1
    for (int k = tid; k < N2; k += kHeadThreads) {
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += sdz2[b * N2 + k];
        put(a.db2 + k, s, acc_mode);
    }
    for (int j = tid; j < N1; j += kHeadThreads) {
        float a1r[BT], xh[BT], da1[BT];
#pragma unroll
        for (int b = 0; b < BT; ++b) {
            a1r[b] = b < B ? a.a1[b * N1 + j] : 0.f;
            xh[b] = b < B ? a.xhat[b * N1 + j] : 0.f;
            da1[b] = 0.f;
        }
        for (int k = 0; k < N2; ++k) {
            const float w = __bfloat162float(a.W2[(size_t)k * N1 + j]);
            float s = 0.f;
#pragma unroll
            for (int b = 0; b < BT; ++b)
                if (b < B) {
                    const float dz = sdz2[b * N2 + k];
                    s += dz * a1r[b];
                    da1[b] += dz * w;
                }
            put(a.dW2 + (size_t)k * N1 + j, s, acc_mode);
        }
        float dgamma = 0.f, dbeta = 0.f;
#pragma unroll
        for (int b = 0; b < BT; ++b)
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
        for (int b = 0; b < BT; ++b)
            if (b < B) {
                const float dz = a.training ? gi * (da1[b] - dbeta * invB - xh[b] * dgamma * invB) : gi * da1[b];
                sdz1[b * N1 + j] = dz;
                db1 += dz;
            }
        put(a.db1 + j, db1, acc_mode);
    }
    __syncthreads();

    // ---- layer 1: thread = input feature k of N0 (coalesced weight reads / gradient writes)
    for (int k = tid; k < N0; k += kHeadThreads) {
        float a0r[BT], dx[BT];
#pragma unroll
        for (int b = 0; b < BT; ++b) {
            a0r[b] = b < B ? sa0[b * N0 + k] : 0.f;
            dx[b] = 0.f;
        }
        for (int j = 0; j < N1; ++j) {
            const float w = __bfloat162float(a.W1[(size_t)j * N0 + k]);
            float s = 0.f;
#pragma unroll
            for (int b = 0; b < BT; ++b)
                if (b < B) {
                    const float dz = sdz1[b * N1 + j];
                    s += dz * a0r[b];
                    dx[b] += dz * w;
                }
            put(a.dW1 + (size_t)j * N0 + k, s, acc_mode);
        }
#pragma unroll
        for (int b = 0; b < BT; ++b)
            if (b < B) a.dx[b * N0 + k] = dx[b] * (a.drop ? a.drop[b * N0 + k] : 1.f);
    }
}

template <typename K, typename A>
int launch1(K kernel, const A& args, size_t smem, cudaStream_t st) {
    DNB_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kernel<<<1, kHeadThreads, smem, st>>>(args);
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

}  // namespace

DNB_EXPORT int dnb_head_forward(const DnbHeadFwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_head_forward: null descriptor");
    DNB_REQUIRE(d->B >= 1 && d->B <= 64, "head: batch must be in [1,64] (got %d)", d->B);
    DNB_REQUIRE(d->N0 % 8 == 0 && d->N1 % 2 == 0, "head: N0 must be a multiple of 8 and N1 even");
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
    const size_t smem = (size_t)d->B * (d->N0 + d->N1 + d->N2) * sizeof(float);
    if (d->B <= 16) return launch1(head_fwd_kernel<16>, a, smem, st);
    if (d->B <= 32) return launch1(head_fwd_kernel<32>, a, smem, st);
    return launch1(head_fwd_kernel<64>, a, smem, st);
}

DNB_EXPORT int dnb_head_backward(const DnbHeadBwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_head_backward: null descriptor");
    DNB_REQUIRE(d->B >= 1 && d->B <= 64, "head: batch must be in [1,64] (got %d)", d->B);
    HeadBwdArgs a;
    a.dlogits = d->dlogits; a.a0 = d->a0; a.drop = d->drop; a.xhat = d->xhat; a.invstd = d->invstd; a.a1 = d->a1; a.a2 = d->a2;
    a.W1 = (const __nv_bfloat16*)d->W1; a.W2 = (const __nv_bfloat16*)d->W2; a.W3 = (const __nv_bfloat16*)d->W3;
    a.bn_w = d->bn_w;
    a.dW1 = d->dW1; a.db1 = d->db1; a.dbn_w = d->dbn_w; a.dbn_b = d->dbn_b; a.dW2 = d->dW2; a.db2 = d->db2;
    a.dW3 = d->dW3; a.db3 = d->db3; a.dx = d->dx;
    a.training = d->training; a.accumulate = d->accumulate;
    a.B = d->B; a.N0 = d->N0; a.N1 = d->N1; a.N2 = d->N2; a.N3 = d->N3;
    DNB_REQUIRE(a.dlogits && a.a0 && a.xhat && a.invstd && a.a1 && a.a2 && a.dW1 && a.db1 && a.dbn_w && a.dbn_b &&
                    a.dW2 && a.db2 && a.dW3 && a.db3 && a.dx,
                "head bwd: null buffer");
    const size_t smem = (size_t)d->B * (d->N3 + 2 * d->N2 + d->N1 + d->N0) * sizeof(float);
    if (d->B <= 16) return launch1(head_bwd_kernel<16>, a, smem, st);
    if (d->B <= 32) return launch1(head_bwd_kernel<32>, a, smem, st);
    return launch1(head_bwd_kernel<64>, a, smem, st);
}
