// K4, default size — the whole FreeSurfer MLP in ONE launch forward and ONE launch backward
// (reference comps/fs/models.py:13-31: per hidden size Linear(bias=False) -> BatchNorm1d(batch statistics, also in
// eval) -> ReLU, then fc_out with bias; default 66 -> 256 -> 128 -> 64 -> 32 -> 2, batch 16, local.py:33).
//
// The default net has 61 k parameters and ~1 MFLOP per forward: at this size the step is pure launch / dependency
// latency (the per-layer path was ~25 launches).  One CTA therefore walks the whole network with the activations
// resident in shared memory (fp32, batch-major) and the bf16 weights streamed from L2 (120 KB).  A 128-row tensor-core tile has nothing to
// offer here (16 batch columns, <= 256 features: the MMA would be >90 % padding and the chain of layers is serial anyway),
// so this kernel uses the FMA pipes; the wide variant of the same net (66 -> 4096 x 6) runs on the tcgen05 GEMM (gemm.cu).
// Limits: batch <= 64, every width <= 256, <= 8 hidden layers, <= 16 classes.
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

constexpr int kThreads = 512;
constexpr int kMaxW = DNB_MLP_SMALL_MAX_WIDTH;     // 256
constexpr int kMaxB = DNB_MLP_SMALL_MAX_BATCH;     // 64
constexpr int kMaxL = DNB_MLP_SMALL_MAX_LAYERS;    // 8
constexpr int kMaxC = DNB_MLP_SMALL_MAX_CLASSES;   // 16

struct Args {
    const float* x;
    const __nv_bfloat16* W[kMaxL];
    const float* gamma[kMaxL];
    const float* beta[kMaxL];
    const __nv_bfloat16* Wfc;
    const float* bfc;
    float* act[kMaxL];        // [B, width] post-ReLU
    float* xhat[kMaxL];       // [B, width]
    float* invstd[kMaxL];     // [width]
    float* logits;            // [B, C]
    // backward
    const float* dlogits;     // [B, C]
    float* dW[kMaxL];
    float* dgamma[kMaxL];
    float* dbeta[kMaxL];
    float* dWfc;
    float* dbfc;
    float* dx;                // optional [B, in0]
    int width[kMaxL];
    int B, in0, L, C, accumulate;
    float eps;
};

__device__ __forceinline__ float bf(const __nv_bfloat16* p) { return __bfloat162float(*p); }

__device__ __forceinline__ void put(float* p, float v, int acc) {
    if (acc) *p += v;
    else *p = v;
}

// ---- asynchronous staging of global data into shared memory (cp.async, 16 bytes per request) -------------------------
// The first version of these kernels read weights and saved activations straight from global memory inside the inner
// loops: every loop iteration then waits for an L2 round trip (~0.3 us), and the step took 0.24 ms for ~3 MFLOP.
// Everything an inner loop touches is therefore staged into shared memory first — the next layer's weights while the
// current layer computes (two buffers, when they fit).
__device__ __forceinline__ void stage_async(void* smem_dst, const void* gsrc, int bytes, bool commit = true) {
    const bool ok = ((reinterpret_cast<uintptr_t>(gsrc) | reinterpret_cast<uintptr_t>(smem_dst)) & 15) == 0;
    const int n16 = ok ? bytes >> 4 : 0;
    const uint32_t d0 = dnb::smem_u32(smem_dst);
    const char* g = reinterpret_cast<const char*>(gsrc);
    for (int i = threadIdx.x; i < n16; i += kThreads)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + i * 16), "l"(g + (long long)i * 16) : "memory");
    // tail (and the whole block when misaligned): plain 2-byte copies, visible after the next __syncthreads
    const unsigned short* gs = reinterpret_cast<const unsigned short*>(gsrc);
    unsigned short* ds = reinterpret_cast<unsigned short*>(smem_dst);
    for (int i = n16 * 8 + threadIdx.x; i < bytes >> 1; i += kThreads) ds[i] = gs[i];
    if (commit) asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void stage_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Shared-memory planes are batch-major: plane[b * width + f].  A thread that owns feature f then walks the batch with
// conflict-free accesses (consecutive threads = consecutive f), and a warp that owns one output feature reads the
// activations of one batch row with its lanes along k (consecutive words).
//
// nxt[b][f] = sum_k cur[b][k] * W[f][k]: one warp per output feature, lanes along k, 16 batch rows per pass, shuffle
// reduction.  W is the staged shared-memory copy (bf16, [out][in]).
__device__ __forceinline__ void dense_fwd(const float* cur, float* nxt, const __nv_bfloat16* W, int in, int out, int B) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int f = warp; f < out; f += kThreads / 32) {
        const __nv_bfloat16* w = W + f * in;
        for (int b0 = 0; b0 < B; b0 += 16) {
            float acc[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = 0.f;
            for (int k = lane; k < in; k += 32) {
                const float wk = bf(w + k);
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (b0 + j < B) acc[j] = fmaf(wk, cur[(b0 + j) * in + k], acc[j]);
            }
            float mine = 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float s = dnb::warp_sum(acc[j]);
                if (lane == j) mine = s;
            }
            if (lane < 16 && b0 + lane < B) nxt[(b0 + lane) * out + f] = mine;
        }
    }
}

// smem: [plane cur][plane nxt][W buffer 0][W buffer 1 (if dbl)];  plane = B * kMaxW floats, W buffer = kMaxW*kMaxW bf16
__global__ void __launch_bounds__(kThreads, 1) mlp_small_fwd_kernel(const Args a, int dbl, int wcap) {
    extern __shared__ __align__(16) float sm[];
    const int B = a.B, plane = B * kMaxW;
    float* cur = sm;
    float* nxt = sm + plane;
    __nv_bfloat16* const w0 = reinterpret_cast<__nv_bfloat16*>(sm + 2 * plane);
    __nv_bfloat16* const w1 = dbl ? w0 + wcap : w0;
    dnb::pdl_wait();
    stage_async(w0, a.W[0], a.width[0] * a.in0 * 2);
    for (int i = threadIdx.x; i < a.in0 * B; i += kThreads) cur[i] = a.x[i];
    int in = a.in0;
#pragma unroll 1
    for (int l = 0; l < a.L; ++l) {
        const int out = a.width[l];
        const bool more = l + 1 < a.L;
        if (dbl && more) {                       // prefetch the next layer's weights into the other buffer
            stage_async(((l + 1) & 1) ? w1 : w0, a.W[l + 1], a.width[l + 1] * out * 2);
            stage_wait<1>();
        } else {
            stage_wait<0>();
        }
        __syncthreads();
        dense_fwd(cur, nxt, (dbl && (l & 1)) ? w1 : w0, in, out, B);
        __syncthreads();
        if (!dbl && more) stage_async(w0, a.W[l + 1], a.width[l + 1] * out * 2);   // overlaps the BatchNorm phase
        // BatchNorm over the batch (biased variance) + ReLU, one thread per feature
        for (int f = threadIdx.x; f < out; f += kThreads) {
            float s = 0.f;
            for (int b = 0; b < B; ++b) s += nxt[b * out + f];
            const float mean = s / (float)B;
            float v = 0.f;
            for (int b = 0; b < B; ++b) {
                const float d = nxt[b * out + f] - mean;
                v = fmaf(d, d, v);
            }
            const float istd = rsqrtf(v / (float)B + a.eps);
            const float g = a.gamma[l][f], be = a.beta[l][f];
            a.invstd[l][f] = istd;
            for (int b = 0; b < B; ++b) {
                const float xh = (nxt[b * out + f] - mean) * istd;
                const float y = fmaxf(fmaf(xh, g, be), 0.f);
                nxt[b * out + f] = y;
                a.xhat[l][(long long)b * out + f] = xh;
                a.act[l][(long long)b * out + f] = y;
            }
        }
        __syncthreads();
        float* t = cur; cur = nxt; nxt = t;
        in = out;
    }
    // fc_out (with bias): one warp per (class, batch row), lanes along k
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int item = warp; item < a.C * B; item += kThreads / 32) {
        const int c = item % a.C, b = item / a.C;
        const __nv_bfloat16* w = a.Wfc + (long long)c * in;
        float s = 0.f;
        for (int k = lane; k < in; k += 32) s = fmaf(bf(w + k), cur[b * in + k], s);
        s = dnb::warp_sum(s);
        if (lane == 0) a.logits[(long long)b * a.C + c] = s + a.bfc[c];
    }
}

// smem: [da][dz][dl 4 KB][W buffer 0][W 1 (if dbl)][a_prev 0 (if stage_ap)][a_prev 1 (if stage_ap && dbl)]
__global__ void __launch_bounds__(kThreads, 1) mlp_small_bwd_kernel(const Args a, int dbl, int stage_ap, int wcap) {
    extern __shared__ __align__(16) float sm[];
    const int B = a.B, C = a.C, acc = a.accumulate & 1, acc_fc = a.accumulate & 2, plane = B * kMaxW;
    float* da = sm;                       // [batch][feature] gradient w.r.t. the current layer's output
    float* dz = sm + plane;               // [batch][feature] gradient w.r.t. its pre-BatchNorm values
    float* dl = sm + 2 * plane;           // [batch][class]
    __nv_bfloat16* const w0 = reinterpret_cast<__nv_bfloat16*>(dl + kMaxB * kMaxC);
    __nv_bfloat16* const w1 = dbl ? w0 + wcap : w0;
    float* const ap0 = reinterpret_cast<float*>(w0 + (dbl ? 2 : 1) * wcap);
    float* const ap1 = dbl ? ap0 + plane : ap0;
    dnb::pdl_wait();
    const int last = a.width[a.L - 1];
    auto stage_layer = [&](int l, int slot) {
        const int in = l > 0 ? a.width[l - 1] : a.in0;
        stage_async(slot ? w1 : w0, a.W[l], a.width[l] * in * 2, false);
        if (stage_ap) stage_async(slot ? ap1 : ap0, l > 0 ? a.act[l - 1] : a.x, B * in * 4, false);
        asm volatile("cp.async.commit_group;" ::: "memory");       // one group per layer
    };
    stage_layer(a.L - 1, dbl ? ((a.L - 1) & 1) : 0);
    for (int i = threadIdx.x; i < B * C; i += kThreads) dl[i] = a.dlogits[i];
    __syncthreads();
    const float* aL = a.act[a.L - 1];
    // fc_out: weight / bias gradients, gradient w.r.t. the last hidden activations
    for (int item = threadIdx.x; item < C * last; item += kThreads) {
        const int k = item % last, c = item / last;
        float s = 0.f;
#pragma unroll 8
        for (int b = 0; b < B; ++b) s = fmaf(dl[b * C + c], __ldg(aL + (long long)b * last + k), s);
        put(a.dWfc + item, s, acc_fc);
    }
    for (int c = threadIdx.x; c < C; c += kThreads) {
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += dl[b * C + c];
        put(a.dbfc + c, s, acc_fc);
    }
    for (int item = threadIdx.x; item < last * B; item += kThreads) {
        const int k = item % last, b = item / last;
        float s = 0.f;
        for (int c = 0; c < C; ++c) s = fmaf(dl[b * C + c], bf(a.Wfc + (long long)c * last + k), s);
        da[b * last + k] = s;
    }
    __syncthreads();
#pragma unroll 1
    for (int l = a.L - 1; l >= 0; --l) {
        const int out = a.width[l], in = l > 0 ? a.width[l - 1] : a.in0;
        const int slot = dbl ? (l & 1) : 0;
        if (dbl && l > 0) stage_layer(l - 1, (l - 1) & 1);       // prefetch the next (lower) layer's operands
        // ReLU + BatchNorm backward, one thread per feature (its saved values are read with several loads in flight)
        for (int f = threadIdx.x; f < out; f += kThreads) {
            const float g = a.gamma[l][f], istd = a.invstd[l][f];
            float s1 = 0.f, s2 = 0.f;
#pragma unroll 8
            for (int b = 0; b < B; ++b) {
                const float y = __ldg(a.act[l] + (long long)b * out + f);
                const float xh = __ldg(a.xhat[l] + (long long)b * out + f);
                const float dy = y > 0.f ? da[b * out + f] : 0.f;
                s1 += dy;
                s2 = fmaf(dy, xh, s2);
                dz[b * out + f] = dy;
            }
            put(a.dgamma[l] + f, s2, acc);
            put(a.dbeta[l] + f, s1, acc);
            const float k0 = g * istd / (float)B;
#pragma unroll 8
            for (int b = 0; b < B; ++b) {
                const float xh = __ldg(a.xhat[l] + (long long)b * out + f);
                dz[b * out + f] = k0 * ((float)B * dz[b * out + f] - s1 - xh * s2);
            }
        }
        if (dbl && l > 0) stage_wait<1>();
        else stage_wait<0>();
        __syncthreads();
        // weight gradient dW[f][k] = sum_b dz[b][f] * a_prev[b][k]   (thread = one (f, k); k fastest: coalesced stores)
        const float* ap = stage_ap ? (slot ? ap1 : ap0) : (l > 0 ? a.act[l - 1] : a.x);
        for (int item = threadIdx.x; item < out * in; item += kThreads) {
            const int k = item % in, f = item / in;
            float s = 0.f;
#pragma unroll 8
            for (int b = 0; b < B; ++b) s = fmaf(dz[b * out + f], ap[b * in + k], s);
            put(a.dW[l] + item, s, acc);
        }
        // gradient w.r.t. the previous activations da_prev[b][k] = sum_f dz[b][f] * W[f][k]
        // (thread = column k x 8 batch rows; staged weights: consecutive threads read consecutive bf16, dz reads broadcast)
        if (l > 0 || a.dx != nullptr) {
            const int nbg = (B + 7) >> 3;
            const __nv_bfloat16* Ws = slot ? w1 : w0;
            for (int item = threadIdx.x; item < in * nbg; item += kThreads) {
                const int k = item % in, b0 = (item / in) * 8;
                float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
                for (int f = 0; f < out; ++f) {
                    const float wk = bf(Ws + f * in + k);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (b0 + j < B) s[j] = fmaf(wk, dz[(b0 + j) * out + f], s[j]);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (b0 + j < B) da[(b0 + j) * in + k] = s[j];
            }
        }
        __syncthreads();
        if (!dbl && l > 0) stage_layer(l - 1, 0);
    }
    if (a.dx != nullptr)
        for (int i = threadIdx.x; i < B * a.in0; i += kThreads) a.dx[i] = da[i];
}

int fill(const DnbMlpSmallDesc* d, Args& a, bool backward) {
    DNB_REQUIRE(d != nullptr, "dnb_mlp_small: null descriptor");
    DNB_REQUIRE(d->B >= 2 && d->B <= kMaxB, "mlp_small: batch must be in [2,%d] (got %d)", kMaxB, d->B);
    DNB_REQUIRE(d->L >= 1 && d->L <= kMaxL, "mlp_small: 1..%d hidden layers (got %d)", kMaxL, d->L);
    DNB_REQUIRE(d->in0 >= 1 && d->in0 <= kMaxW, "mlp_small: input width must be <= %d (got %d)", kMaxW, d->in0);
    DNB_REQUIRE(d->C >= 1 && d->C <= kMaxC, "mlp_small: classes must be <= %d (got %d)", kMaxC, d->C);
    memset(&a, 0, sizeof(a));
    for (int l = 0; l < d->L; ++l) {
        DNB_REQUIRE(d->width[l] >= 1 && d->width[l] <= kMaxW, "mlp_small: width[%d] = %d outside [1,%d]", l, d->width[l], kMaxW);
        DNB_REQUIRE(d->W[l] && d->gamma[l] && d->beta[l] && d->act[l] && d->xhat[l] && d->invstd[l], "mlp_small: null buffer in layer %d", l);
        a.W[l] = reinterpret_cast<const __nv_bfloat16*>(d->W[l]);
        a.gamma[l] = d->gamma[l]; a.beta[l] = d->beta[l];
        a.act[l] = d->act[l]; a.xhat[l] = d->xhat[l]; a.invstd[l] = d->invstd[l];
        a.width[l] = d->width[l];
        if (backward) {
            DNB_REQUIRE(d->dW[l] && d->dgamma[l] && d->dbeta[l], "mlp_small backward: null gradient buffer in layer %d", l);
            a.dW[l] = d->dW[l]; a.dgamma[l] = d->dgamma[l]; a.dbeta[l] = d->dbeta[l];
        }
    }
    DNB_REQUIRE(d->x && d->Wfc && d->bfc, "mlp_small: null input / fc_out buffer");
    a.x = d->x; a.Wfc = reinterpret_cast<const __nv_bfloat16*>(d->Wfc); a.bfc = d->bfc;
    a.logits = d->logits; a.dlogits = d->dlogits; a.dWfc = d->dWfc; a.dbfc = d->dbfc; a.dx = d->dx;
    if (backward) DNB_REQUIRE(d->dlogits && d->dWfc && d->dbfc, "mlp_small backward: null dlogits / fc_out gradient");
    else DNB_REQUIRE(d->logits, "mlp_small forward: logits is null");
    a.B = d->B; a.in0 = d->in0; a.L = d->L; a.C = d->C; a.accumulate = d->accumulate; a.eps = d->eps;
    return 0;
}

constexpr int kSmemLimit = 227 * 1024;

// bf16 elements one weight buffer must hold: the largest layer of THIS network (the default net: 256 x 128 = 64 KB)
int weight_capacity(const Args& a) {
    int cap = 0, in = a.in0;
    for (int l = 0; l < a.L; ++l) {
        const int n = a.width[l] * in;
        if (n > cap) cap = n;
        in = a.width[l];
    }
    return (cap + 7) & ~7;
}

int launch_fwd(const Args& a, cudaStream_t st) {
    const int plane = a.B * kMaxW * 4, wcap = weight_capacity(a), kWBytes = wcap * 2;
    int dbl = 1;
    int need = 2 * plane + 2 * kWBytes;
    if (need > kSmemLimit) { dbl = 0; need = 2 * plane + kWBytes; }
    DNB_REQUIRE(need <= kSmemLimit, "mlp_small forward: batch %d with these widths needs %d bytes of shared memory", a.B, need);
    static int configured = 0;
    if (configured < need) {
        DNB_CUDA_CHECK(cudaFuncSetAttribute(mlp_small_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
        configured = kSmemLimit;
    }
    DNB_CUDA_CHECK(dnb_launch(mlp_small_fwd_kernel, dim3(1), dim3(kThreads), (size_t)need, st, 0, a, dbl, wcap));
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

int launch_bwd(const Args& a, cudaStream_t st) {
    const int plane = a.B * kMaxW * 4, fixed = 2 * plane + kMaxB * kMaxC * 4, wcap = weight_capacity(a), kWBytes = wcap * 2;
    int dbl = 1, stage_ap = 1;
    int need = fixed + 2 * kWBytes + 2 * plane;
    if (need > kSmemLimit) { dbl = 0; need = fixed + kWBytes + plane; }
    if (need > kSmemLimit) { stage_ap = 0; need = fixed + kWBytes; }
    DNB_REQUIRE(need <= kSmemLimit, "mlp_small backward: batch %d with these widths needs %d bytes of shared memory", a.B, need);
    static int configured = 0;
    if (configured < need) {
        DNB_CUDA_CHECK(cudaFuncSetAttribute(mlp_small_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
        configured = kSmemLimit;
    }
    DNB_CUDA_CHECK(dnb_launch(mlp_small_bwd_kernel, dim3(1), dim3(kThreads), (size_t)need, st, 0, a, dbl, stage_ap, wcap));
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

}  // namespace

DNB_EXPORT int dnb_mlp_small_forward(const DnbMlpSmallDesc* d, void* stream_) {
    Args a;
    if (int rc = fill(d, a, false)) return rc;
    return launch_fwd(a, reinterpret_cast<cudaStream_t>(stream_));
}

DNB_EXPORT int dnb_mlp_small_backward(const DnbMlpSmallDesc* d, void* stream_) {
    Args a;
    if (int rc = fill(d, a, true)) return rc;
    return launch_bwd(a, reinterpret_cast<cudaStream_t>(stream_));
}
