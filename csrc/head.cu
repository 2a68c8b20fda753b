// K8 — classifier head of the ICA-LSTM, forward and backward.
//
// Reference: comps/icalstm/models.py:95-103 —
//   Dropout(.25) -> Linear(N0,N1) -> BatchNorm1d(N1) -> ReLU -> Linear(N1,N2) -> ReLU -> Linear(N2,N3)
// with N0 = hidden (384), N1 = 256, N2 = 64, N3 = classes, batch 16-64.  This is latency-bound
// small-matrix work (M = batch) on the serial chain LSTM -> head -> loss -> head' -> LSTM', and every
// kernel runs once per step, so what matters is (a) many warps in flight and (b) few instructions per
// warp (a long unrolled instruction stream executed once by 8 warps costs ~6 cycles per instruction:
// the first versions spent 20-30 us per launch that way).  Hence: ONE output element per thread with a
// short loop over K, 16 consecutive lanes = 16 batch rows of one feature, so BatchNorm's per-feature
// batch statistics are 16-lane shuffle reductions.
//   forward  : [N1/8 CTAs] Linear1 + BN + ReLU        -> [N2/16 CTAs] Linear2 + ReLU + partial Linear3
//   backward : [N1/16 CTAs] layers 3, 2 and BN         -> [N1/8 + N0/16 CTAs] dW1 and dx
// Weights are read as bf16 (the parameter shadow); accumulation, saved activations and all gradients
// are fp32.  Parameter gradients are accumulated (+=, fire-and-forget RED) when `accumulate` is set.
#include "common.cuh"
#include "../include/dinunet_b200.h"

#include <mutex>

void dnb_count_launch(int n);

namespace {

constexpr int kMaxB = 64;        // batch rows supported (4 chunks of 16 lanes)
constexpr int kChunks = kMaxB / 16;

struct HeadFwdArgs {
    const float* x; const float* drop;
    const __nv_bfloat16 *W1, *W2, *W3;
    const float *b1, *b2, *b3, *bn_w, *bn_b;
    float *run_mean, *run_var;
    float momentum, eps; int training;
    float *a0, *xhat, *invstd, *a1, *a2, *logits;
    int B, N0, N1, N2, N3;
    float* drop_out; float drop_p; unsigned long long drop_seed; unsigned long long* drop_state;
    long long* nbt;
};

struct HeadBwdArgs {
    const float* dlogits;
    const float *a0, *drop, *xhat, *invstd, *a1, *a2;
    const __nv_bfloat16 *W1, *W2, *W3;
    const float* bn_w;
    float *dW1, *db1, *dbn_w, *dbn_b, *dW2, *db2, *dW3, *db3, *dx;
    float* dz1;                              // [B,N1] scratch between the two backward launches
    float* dz2;                              // optional [B,N2]: gradient at the second layer's pre-activation (rankDAD pairs)
    int training, accumulate;
    int B, N0, N1, N2, N3;
};

// accumulate with a fire-and-forget RED (no load -> no L2 round trip on the critical path)
__device__ __forceinline__ void put(float* p, float v, int acc) {
    if (acc) atomicAdd(p, v);
    else *p = v;
}

// sum over the 16 lanes that share a feature (xor offsets < 16 stay inside a half-warp)
__device__ __forceinline__ float sum16(float v) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Counter-based dropout: 64 random bits per group of 4 elements from splitmix64(seed, step, group index);
// element e keeps its value (scaled by 1/(1-p)) when its 16-bit lane is >= p * 65536.
__device__ __forceinline__ uint64_t drop_bits(uint64_t seed, uint64_t step, uint32_t group) {
    uint64_t z = seed + step * 0xD1B54A32D192ED03ull + (uint64_t)group * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__device__ __forceinline__ float dot8(const uint4& w, const float4& x0, const float4& x1) {
    return __uint_as_float(w.x << 16) * x0.x + __uint_as_float(w.x & 0xFFFF0000u) * x0.y +
           __uint_as_float(w.y << 16) * x0.z + __uint_as_float(w.y & 0xFFFF0000u) * x0.w +
           __uint_as_float(w.z << 16) * x1.x + __uint_as_float(w.z & 0xFFFF0000u) * x1.y +
           __uint_as_float(w.w << 16) * x1.z + __uint_as_float(w.w & 0xFFFF0000u) * x1.w;
}

// ------------------------------------------------------------------ forward, launch 1: Linear1 + BN + ReLU
// CTA = 8 features x 16 lanes.  Thread (f, l) owns batch rows l, l+16, ... of feature f.
constexpr int kF1Feat = 8;
constexpr int kF1Threads = kF1Feat * 16;
__global__ void __launch_bounds__(kF1Threads) head_f1_kernel(const HeadFwdArgs a) {
    dnb::pdl_wait();       // inputs come from the preceding kernel
    dnb::pdl_launch();     // short kernel: let the successor get scheduled right away
    extern __shared__ float sm[];
    const int B = a.B, N0 = a.N0, ldx = N0 + 4;       // +4 floats: rows of different lanes land in different banks
    float* xs = sm;                                    // [B][ldx] dropped input
    const int tid = threadIdx.x, f = tid >> 4, l = tid & 15;
    {
        const int per_row = N0 / 4;
        const float4* x4 = reinterpret_cast<const float4*>(a.x);
        const float4* d4 = reinterpret_cast<const float4*>(a.drop);
        const bool gen = d4 == nullptr && a.drop_state != nullptr && a.drop_p > 0.f;
        const uint64_t step = gen ? *a.drop_state : 0ull;       // advanced by launch 2 (after every CTA has read it)
        const uint32_t thr = gen ? (uint32_t)(a.drop_p * 65536.f) : 0u;
        const float keep = gen ? 1.f / (1.f - a.drop_p) : 1.f;
#pragma unroll 12      // B = 16, N0 = 384: 12 trips, all loads in flight at once
        for (int i = tid; i < B * per_row; i += kF1Threads) {
            float4 v = x4[i];
            if (d4) {
                const float4 d = d4[i];
                v.x *= d.x; v.y *= d.y; v.z *= d.z; v.w *= d.w;
            } else if (gen) {
                const uint64_t r = drop_bits(a.drop_seed, step, (uint32_t)i);
                float4 d;
                d.x = ((uint32_t)(r & 0xFFFF) >= thr) ? keep : 0.f;
                d.y = ((uint32_t)((r >> 16) & 0xFFFF) >= thr) ? keep : 0.f;
                d.z = ((uint32_t)((r >> 32) & 0xFFFF) >= thr) ? keep : 0.f;
                d.w = ((uint32_t)(r >> 48) >= thr) ? keep : 0.f;
                v.x *= d.x; v.y *= d.y; v.z *= d.z; v.w *= d.w;
                if (blockIdx.x == 0) reinterpret_cast<float4*>(a.drop_out)[i] = d;
            }
            const int r = i / per_row, c = i % per_row;
            *reinterpret_cast<float4*>(xs + r * ldx + c * 4) = v;
            if (a.a0 && blockIdx.x == 0) reinterpret_cast<float4*>(a.a0)[i] = v;
        }
        if (blockIdx.x == 0)                           // Linear3 accumulates into logits in launch 2: start from the bias
            for (int i = tid; i < B * a.N3; i += kF1Threads) a.logits[i] = a.b3[i % a.N3];
    }
    __syncthreads();
    const int j = blockIdx.x * kF1Feat + f;
    if (j >= a.N1) return;
    const uint4* w = reinterpret_cast<const uint4*>(a.W1 + (size_t)j * N0);
    float z[kChunks];
#pragma unroll
    for (int c = 0; c < kChunks; ++c) z[c] = 0.f;
#pragma unroll
    for (int c = 0; c < kChunks; ++c) {
        if (c * 16 < B) {
            const int b = c * 16 + l;
            const float* xr = xs + (b < B ? b : 0) * ldx;
            float acc = 0.f;
#pragma unroll 12      // weight rows come from L2: keep a dozen 16-byte loads in flight per thread
            for (int k = 0; k < N0 / 8; ++k)
                acc += dot8(w[k], *reinterpret_cast<const float4*>(xr + k * 8),
                            *reinterpret_cast<const float4*>(xr + k * 8 + 4));
            z[c] = b < B ? acc + a.b1[j] : 0.f;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChunks; ++c) s += z[c];
    const float mean = sum16(s) / (float)B;
    float q = 0.f;
#pragma unroll
    for (int c = 0; c < kChunks; ++c) q += (c * 16 + l < B) ? (z[c] - mean) * (z[c] - mean) : 0.f;
    const float var = sum16(q) / (float)B;
    float use_mean = mean, use_var = var;
    if (a.training) {
        if (a.run_mean && l == 0) {
            a.run_mean[j] = (1.f - a.momentum) * a.run_mean[j] + a.momentum * mean;
            a.run_var[j] = (1.f - a.momentum) * a.run_var[j] + a.momentum * (B > 1 ? var * (float)B / (float)(B - 1) : var);
        }
    } else {
        use_mean = a.run_mean[j];
        use_var = a.run_var[j];
    }
    const float istd = rsqrtf(use_var + a.eps);
    if (l == 0) a.invstd[j] = istd;
    const float g = a.bn_w[j], be = a.bn_b[j];
#pragma unroll
    for (int c = 0; c < kChunks; ++c) {
        const int b = c * 16 + l;
        if (b < B) {
            const float xh = (z[c] - use_mean) * istd;
            a.xhat[b * a.N1 + j] = xh;
            a.a1[b * a.N1 + j] = fmaxf(xh * g + be, 0.f);
        }
    }
}

// ------------------------------------------------------------------ forward, launch 2: Linear2 + ReLU + Linear3
// CTA = 16 features of layer 2 x 16 lanes; Linear3 is accumulated into `logits` (preset to b3 by
// launch 1) with one atomic per (row, class, CTA).
constexpr int kF2Feat = 16;
constexpr int kF2Threads = kF2Feat * 16;
__global__ void __launch_bounds__(kF2Threads) head_f2_kernel(const HeadFwdArgs a) {
    dnb::pdl_wait();       // inputs come from the preceding kernel
    dnb::pdl_launch();     // short kernel: let the successor get scheduled right away
    extern __shared__ float sm[];
    const int B = a.B, N1 = a.N1, N2 = a.N2, N3 = a.N3, lda = N1 + 4;
    float* s1 = sm;                        // [B][lda] layer-1 activations
    float* s2 = s1 + B * lda;              // [B][kF2Feat] this CTA's layer-2 activations
    const int tid = threadIdx.x, f = tid >> 4, l = tid & 15;
    {
        const int per_row = N1 / 4;
        const float4* src = reinterpret_cast<const float4*>(a.a1);
#pragma unroll 4
        for (int i = tid; i < B * per_row; i += kF2Threads) {
            const int r = i / per_row, c = i % per_row;
            *reinterpret_cast<float4*>(s1 + r * lda + c * 4) = src[i];
        }
    }
    __syncthreads();
    const int j0 = blockIdx.x * kF2Feat, j = j0 + f;
    if (j < N2) {
        const uint4* w = reinterpret_cast<const uint4*>(a.W2 + (size_t)j * N1);
        const float bias = a.b2[j];
#pragma unroll 1
        for (int c = 0; c * 16 < B; ++c) {
            const int b = c * 16 + l;
            const float* xr = s1 + (b < B ? b : 0) * lda;
            float acc = 0.f;
#pragma unroll 8
            for (int k = 0; k < N1 / 8; ++k)
                acc += dot8(w[k], *reinterpret_cast<const float4*>(xr + k * 8),
                            *reinterpret_cast<const float4*>(xr + k * 8 + 4));
            if (b < B) {
                const float v = fmaxf(acc + bias, 0.f);
                s2[b * kF2Feat + f] = v;
                a.a2[b * N2 + j] = v;
            }
        }
    }
    __syncthreads();
    if (blockIdx.x == 0 && tid == 0) {     // bookkeeping that must follow launch 1 as a whole
        if (a.drop == nullptr && a.drop_state != nullptr && a.drop_p > 0.f) *a.drop_state += 1ull;
        if (a.nbt != nullptr && a.training) *a.nbt += 1ll;
    }
    const int nf = min(kF2Feat, N2 - j0);
    for (int idx = tid; idx < B * N3; idx += kF2Threads) {
        const int b = idx / N3, c = idx % N3;
        float acc = 0.f;
        for (int k = 0; k < nf; ++k) acc += __bfloat162float(a.W3[(size_t)c * N2 + j0 + k]) * s2[b * kF2Feat + k];
        atomicAdd(a.logits + idx, acc);
    }
}

// ------------------------------------------------------------------ backward, launch 1: layers 3, 2 and BatchNorm
// CTA = 16 features of layer 1 x 16 lanes (batch).  Every CTA rebuilds the tiny dz2 = dL/d(pre-ReLU2)
// [B, N2] in shared memory; CTA 0 also produces the layer-3 / bias-2 gradients.  Zeroes dx for launch 2.
constexpr int kB1Feat = 16;
constexpr int kB1Threads = kB1Feat * 16;
__global__ void __launch_bounds__(kB1Threads) head_b1_kernel(const HeadBwdArgs a) {
    dnb::pdl_wait();       // inputs come from the preceding kernel
    dnb::pdl_launch();     // short kernel: let the successor get scheduled right away
    extern __shared__ float sm[];
    const int B = a.B, N0 = a.N0, N1 = a.N1, N2 = a.N2, N3 = a.N3, acc_mode = a.accumulate;
    const int ldz = N2 + 1;
    float* sdl = sm;                       // [B][N3]
    float* sdz2 = sdl + B * N3;            // [B][ldz]
    float* sa1 = sdz2 + B * ldz;           // [B][kB1Feat] this CTA's layer-1 activations
    const int tid = threadIdx.x, f = tid >> 4, l = tid & 15;
    for (int i = tid; i < B * N3; i += kB1Threads) sdl[i] = a.dlogits[i];
    for (int i = blockIdx.x * kB1Threads + tid; i < B * N0; i += gridDim.x * kB1Threads) a.dx[i] = 0.f;
    __syncthreads();
    for (int idx = tid; idx < B * N2; idx += kB1Threads) {
        const int b = idx / N2, k = idx % N2;
        float s = 0.f;
        for (int c = 0; c < N3; ++c) s += sdl[b * N3 + c] * __bfloat162float(a.W3[c * N2 + k]);
        sdz2[b * ldz + k] = a.a2[idx] > 0.f ? s : 0.f;
    }
    __syncthreads();
    if (blockIdx.x == gridDim.x - 1) {     // one extra block: layer-3 and bias-2 gradients, off the other blocks' path
        for (int idx = tid; idx < N3 * N2; idx += kB1Threads) {
            const int c = idx / N2, k = idx % N2;
            float s = 0.f;
            for (int b = 0; b < B; ++b) s += sdl[b * N3 + c] * a.a2[b * N2 + k];
            put(a.dW3 + idx, s, acc_mode);
        }
        for (int c = tid; c < N3; c += kB1Threads) {
            float s = 0.f;
            for (int b = 0; b < B; ++b) s += sdl[b * N3 + c];
            put(a.db3 + c, s, acc_mode);
        }
        for (int k = tid; k < N2; k += kB1Threads) {
            float s = 0.f;
            for (int b = 0; b < B; ++b) s += sdz2[b * ldz + k];
            put(a.db2 + k, s, acc_mode);
        }
        if (a.dz2 != nullptr)
            for (int idx = tid; idx < B * N2; idx += kB1Threads) a.dz2[idx] = sdz2[(idx / N2) * ldz + idx % N2];
        return;
    }
    const int j0 = blockIdx.x * kB1Feat, j = j0 + f;     // N1 % 16 == 0: every thread has a valid feature
    float dy[kChunks], xh[kChunks];
    float sg = 0.f, sb = 0.f;
#pragma unroll
    for (int c = 0; c < kChunks; ++c) {
        dy[c] = 0.f;
        xh[c] = 0.f;
        if (c * 16 < B) {
            const int b = c * 16 + l;
            const bool ok = b < B;
            const float* dz = sdz2 + (ok ? b : 0) * ldz;
            float acc = 0.f;
#pragma unroll 16
            for (int k = 0; k < N2; ++k) acc += dz[k] * __bfloat162float(a.W2[(size_t)k * N1 + j]);
            const float a1v = ok ? a.a1[b * N1 + j] : 0.f;
            if (ok) sa1[b * kB1Feat + f] = a1v;
            xh[c] = ok ? a.xhat[b * N1 + j] : 0.f;
            dy[c] = (ok && a1v > 0.f) ? acc : 0.f;
            sg += dy[c] * xh[c];
            sb += dy[c];
        }
    }
    const float dgamma = sum16(sg), dbeta = sum16(sb);
    const float gi = a.bn_w[j] * a.invstd[j], invB = 1.f / (float)B;
    float s1 = 0.f;
#pragma unroll
    for (int c = 0; c < kChunks; ++c) {
        const int b = c * 16 + l;
        if (b < B) {
            const float dz = a.training ? gi * (dy[c] - dbeta * invB - xh[c] * dgamma * invB) : gi * dy[c];
            a.dz1[b * N1 + j] = dz;
            s1 += dz;
        }
    }
    s1 = sum16(s1);
    if (l == 0) {
        put(a.dbn_w + j, dgamma, acc_mode);
        put(a.dbn_b + j, dbeta, acc_mode);
        put(a.db1 + j, s1, acc_mode);
    }
    __syncthreads();
    // dW2[k][j0..j0+15] (+)= sum_b dz2[b][k] * a1[b][j]
    for (int idx = tid; idx < N2 * kB1Feat; idx += kB1Threads) {
        const int k = idx / kB1Feat, ff = idx % kB1Feat;
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += sdz2[b * ldz + k] * sa1[b * kB1Feat + ff];
        put(a.dW2 + (size_t)k * N1 + j0 + ff, s, acc_mode);
    }
}

// ------------------------------------------------------------------ backward, launch 2: dW1 and dx
// blocks [0, n_j)      : dW1 rows for 8 features, thread = input column (coalesced)
// blocks [n_j, +n_k*8) : dx for 128 input columns x one eighth of the features, thread = (8 columns, batch lane);
//                        partial sums are added to dx (zeroed by launch 1)
constexpr int kB2Threads = 256;
constexpr int kJTile = 8;
constexpr int kKTile = 128;
constexpr int kJSplit = 8;
__global__ void __launch_bounds__(kB2Threads) head_b2_kernel(const HeadBwdArgs a, int n_jblocks) {
    dnb::pdl_wait();       // inputs come from the preceding kernel
    dnb::pdl_launch();     // short kernel: let the successor get scheduled right away
    extern __shared__ float sm[];
    const int B = a.B, N0 = a.N0, N1 = a.N1, tid = threadIdx.x;
    if ((int)blockIdx.x < n_jblocks) {
        float* a0s = sm;                        // [B][N0]
        float* dzt = a0s + B * N0;              // [kJTile][B]
        const int j0 = blockIdx.x * kJTile;
        {
            const int n4 = B * N0 / 4;
            const float4* src = reinterpret_cast<const float4*>(a.a0);
#pragma unroll 4
            for (int i = tid; i < n4; i += kB2Threads) reinterpret_cast<float4*>(a0s)[i] = src[i];
        }
        for (int i = tid; i < kJTile * B; i += kB2Threads) {
            const int jj = i / B, b = i % B;
            dzt[i] = (j0 + jj < N1) ? a.dz1[b * N1 + j0 + jj] : 0.f;
        }
        __syncthreads();
        for (int k = tid; k < N0; k += kB2Threads) {
            float s[kJTile];
#pragma unroll
            for (int jj = 0; jj < kJTile; ++jj) s[jj] = 0.f;
#pragma unroll 2
            for (int b = 0; b < B; ++b) {
                const float xv = a0s[b * N0 + k];
#pragma unroll
                for (int jj = 0; jj < kJTile; ++jj) s[jj] += dzt[jj * B + b] * xv;
            }
#pragma unroll
            for (int jj = 0; jj < kJTile; ++jj)
                if (j0 + jj < N1) put(a.dW1 + (size_t)(j0 + jj) * N0 + k, s[jj], a.accumulate);
        }
    } else {
        const int blk = blockIdx.x - n_jblocks;
        const int k0 = (blk / kJSplit) * kKTile, jpart = blk % kJSplit;
        const int jper = (N1 + kJSplit - 1) / kJSplit;
        const int jb = jpart * jper, je = min(N1, jb + jper);
        float* dzs = sm;                        // [B][jper + 1]
        const int ldj = jper + 1;
        for (int i = tid; i < B * (je - jb); i += kB2Threads) {
            const int b = i / (je - jb), jj = i % (je - jb);
            dzs[b * ldj + jj] = a.dz1[b * N1 + jb + jj];
        }
        __syncthreads();
        const int kg = tid >> 4, l = tid & 15;
        const int k = k0 + kg * 8;
        if (k < N0) {
#pragma unroll 1
            for (int c = 0; c * 16 < B; ++c) {
                const int b = c * 16 + l;
                const float* dz = dzs + (b < B ? b : 0) * ldj;
                float acc[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll 8
                for (int j = jb; j < je; ++j) {
                    const uint4 w = *reinterpret_cast<const uint4*>(a.W1 + (size_t)j * N0 + k);
                    const float d = dz[j - jb];
                    acc[0] += d * __uint_as_float(w.x << 16); acc[1] += d * __uint_as_float(w.x & 0xFFFF0000u);
                    acc[2] += d * __uint_as_float(w.y << 16); acc[3] += d * __uint_as_float(w.y & 0xFFFF0000u);
                    acc[4] += d * __uint_as_float(w.z << 16); acc[5] += d * __uint_as_float(w.z & 0xFFFF0000u);
                    acc[6] += d * __uint_as_float(w.w << 16); acc[7] += d * __uint_as_float(w.w & 0xFFFF0000u);
                }
                if (b < B) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const size_t o = (size_t)b * N0 + k + i;
                        atomicAdd(a.dx + o, acc[i] * (a.drop ? a.drop[o] : 1.f));
                    }
                }
            }
        }
    }
}

int launch_head_backward(const HeadBwdArgs& a, cudaStream_t st) {
    {
        const size_t smem = ((size_t)a.B * (a.N3 + a.N2 + 1 + kB1Feat)) * sizeof(float);
        DNB_CUDA_CHECK(dnb_launch(head_b1_kernel, dim3(a.N1 / kB1Feat + 1), dim3(kB1Threads), smem, st, 0, a));
        DNB_LAUNCH_CHECK();
    }
    {
        const int n_j = dnb_ceil_div(a.N1, kJTile), n_k = dnb_ceil_div(a.N0, kKTile);
        const size_t smem_j = ((size_t)a.B * a.N0 + kJTile * a.B) * sizeof(float);
        const size_t smem_k = (size_t)a.B * ((a.N1 + kJSplit - 1) / kJSplit + 1) * sizeof(float);
        const size_t smem = smem_j > smem_k ? smem_j : smem_k;
        DNB_CUDA_CHECK(cudaFuncSetAttribute(head_b2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        DNB_CUDA_CHECK(dnb_launch(head_b2_kernel, dim3(n_j + n_k * kJSplit), dim3(kB2Threads), smem, st, 0, a, n_j));
        DNB_LAUNCH_CHECK();
    }
    dnb_count_launch(2);
    return 0;
}

}  // namespace

DNB_EXPORT int dnb_head_forward(const DnbHeadFwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_head_forward: null descriptor");
    DNB_REQUIRE(d->B >= 1 && d->B <= kMaxB, "head: batch must be in [1,%d] (got %d)", kMaxB, d->B);
    DNB_REQUIRE(d->N0 % 8 == 0 && d->N1 % 16 == 0, "head: N0 must be a multiple of 8 and N1 of 16");
    DNB_REQUIRE(!d->training || d->B > 1, "head: BatchNorm in training mode needs more than one sample");
    HeadFwdArgs a;
    a.x = d->x; a.drop = d->drop;
    a.W1 = (const __nv_bfloat16*)d->W1; a.W2 = (const __nv_bfloat16*)d->W2; a.W3 = (const __nv_bfloat16*)d->W3;
    a.b1 = d->b1; a.b2 = d->b2; a.b3 = d->b3; a.bn_w = d->bn_w; a.bn_b = d->bn_b;
    a.run_mean = d->run_mean; a.run_var = d->run_var; a.momentum = d->momentum; a.eps = d->eps; a.training = d->training;
    a.a0 = d->a0; a.xhat = d->xhat; a.invstd = d->invstd; a.a1 = d->a1; a.a2 = d->a2; a.logits = d->logits;
    a.B = d->B; a.N0 = d->N0; a.N1 = d->N1; a.N2 = d->N2; a.N3 = d->N3;
    a.drop_out = d->drop_out; a.drop_p = d->drop_p; a.drop_seed = d->drop_seed; a.drop_state = d->drop_state;
    a.nbt = d->num_batches_tracked;
    if (a.drop == nullptr && a.drop_state != nullptr && a.drop_p > 0.f) {
        DNB_REQUIRE(a.drop_p < 1.f && a.drop_out != nullptr && a.a0 != nullptr,
                    "head fwd: in-kernel dropout needs 0 < p < 1 and the drop_out / a0 buffers");
    }
    DNB_REQUIRE(a.x && a.W1 && a.W2 && a.W3 && a.xhat && a.invstd && a.a1 && a.a2 && a.logits, "head fwd: null buffer");
    DNB_REQUIRE(a.training || (a.run_mean && a.run_var), "head fwd: eval mode needs running statistics");
    {
        const size_t smem = (size_t)d->B * (d->N0 + 4) * sizeof(float);
        DNB_CUDA_CHECK(cudaFuncSetAttribute(head_f1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        DNB_CUDA_CHECK(dnb_launch(head_f1_kernel, dim3(d->N1 / kF1Feat), dim3(kF1Threads), smem, st, 0, a));
        DNB_LAUNCH_CHECK();
    }
    {
        const size_t smem = (size_t)d->B * (d->N1 + 4 + kF2Feat) * sizeof(float);
        DNB_CUDA_CHECK(cudaFuncSetAttribute(head_f2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        DNB_CUDA_CHECK(dnb_launch(head_f2_kernel, dim3(dnb_ceil_div(d->N2, kF2Feat)), dim3(kF2Threads), smem, st, 0, a));
        DNB_LAUNCH_CHECK();
    }
    dnb_count_launch(2);
    return 0;
}

DNB_EXPORT int dnb_head_backward(const DnbHeadBwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_head_backward: null descriptor");
    DNB_REQUIRE(d->B >= 1 && d->B <= kMaxB, "head: batch must be in [1,%d] (got %d)", kMaxB, d->B);
    DNB_REQUIRE(d->N0 % 8 == 0 && d->N1 % 16 == 0, "head: N0 must be a multiple of 8 and N1 of 16");
    HeadBwdArgs a;
    a.dlogits = d->dlogits; a.a0 = d->a0; a.drop = d->drop; a.xhat = d->xhat; a.invstd = d->invstd; a.a1 = d->a1; a.a2 = d->a2;
    a.W1 = (const __nv_bfloat16*)d->W1; a.W2 = (const __nv_bfloat16*)d->W2; a.W3 = (const __nv_bfloat16*)d->W3;
    a.bn_w = d->bn_w;
    a.dW1 = d->dW1; a.db1 = d->db1; a.dbn_w = d->dbn_w; a.dbn_b = d->dbn_b; a.dW2 = d->dW2; a.db2 = d->db2;
    a.dW3 = d->dW3; a.db3 = d->db3; a.dx = d->dx; a.dz1 = d->dz1; a.dz2 = d->dz2;
    a.training = d->training; a.accumulate = d->accumulate;
    a.B = d->B; a.N0 = d->N0; a.N1 = d->N1; a.N2 = d->N2; a.N3 = d->N3;
    DNB_REQUIRE(a.dlogits && a.a0 && a.xhat && a.invstd && a.a1 && a.a2 && a.dW1 && a.db1 && a.dbn_w && a.dbn_b &&
                    a.dW2 && a.db2 && a.dW3 && a.db3 && a.dx && a.dz1,
                "head bwd: null buffer");
    return launch_head_backward(a, st);
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
//  * The address is the one of the CURRENT device: make the device that owns `stream` current before the
//    call (a stream of another device gets that other device's launch with this device's pointer).
//  * The cached address does not survive cudaDeviceReset(): a process that resets a device must not call
//    this entry point on it again.
DNB_EXPORT int dnb_head_backward_noalloc(const DnbHeadBwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_head_backward_noalloc: null descriptor");
    DNB_REQUIRE(d->B >= 1 && d->B <= kStaticMaxB,
                "dnb_head_backward_noalloc: batch %d exceeds the built-in scratch (max %d rows); "
                "use dnb_head_backward with caller-provided dz1", d->B, kStaticMaxB);
    DNB_REQUIRE(d->N1 >= 1 && d->N1 <= kStaticMaxN1,
                "dnb_head_backward_noalloc: N1 = %d exceeds the built-in scratch (max %d features); "
                "use dnb_head_backward with caller-provided dz1", d->N1, kStaticMaxN1);
    DNB_REQUIRE(d->N0 % 8 == 0 && d->N1 % 16 == 0, "dnb_head_backward_noalloc: N0 must be a multiple of 8 and N1 of 16");
    float* scratch = nullptr;
    if (int rc = static_scratch_for_current_device(&scratch)) return rc;

    HeadBwdArgs a;
    a.dlogits = d->dlogits; a.a0 = d->a0; a.drop = d->drop; a.xhat = d->xhat; a.invstd = d->invstd; a.a1 = d->a1; a.a2 = d->a2;
    a.W1 = (const __nv_bfloat16*)d->W1; a.W2 = (const __nv_bfloat16*)d->W2; a.W3 = (const __nv_bfloat16*)d->W3;
    a.bn_w = d->bn_w;
    a.dW1 = d->dW1; a.db1 = d->db1; a.dbn_w = d->dbn_w; a.dbn_b = d->dbn_b; a.dW2 = d->dW2; a.db2 = d->db2;
    a.dW3 = d->dW3; a.db3 = d->db3; a.dx = d->dx;
    a.dz1 = scratch;                                   // device address, never the host-side symbol
    a.dz2 = d->dz2;
    a.training = d->training; a.accumulate = d->accumulate;
    a.B = d->B; a.N0 = d->N0; a.N1 = d->N1; a.N2 = d->N2; a.N3 = d->N3;
    DNB_REQUIRE(a.dlogits && a.a0 && a.xhat && a.invstd && a.a1 && a.a2 && a.dW1 && a.db1 && a.dbn_w && a.dbn_b &&
                    a.dW2 && a.db2 && a.dW3 && a.db3 && a.dx,
                "dnb_head_backward_noalloc: null buffer");
    return launch_head_backward(a, st);
}
