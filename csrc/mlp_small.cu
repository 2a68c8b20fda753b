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
constexpr int kBufFloats = kMaxW * kMaxB;          // one [feature][batch] plane
constexpr int kSmemBytes = 2 * kBufFloats * 4 + kMaxB * kMaxC * 4;

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

// Shared-memory planes are batch-major: plane[b * width + f].  A thread that owns feature f then walks the batch with
// conflict-free accesses (consecutive threads = consecutive f), and a warp that owns one output feature reads the
// activations of one batch row with its lanes along k (consecutive words).
//
// nxt[b][f] = sum_k cur[b][k] * W[f][k]: one warp per output feature, lanes along k (the bf16 weight row is read as
// consecutive 2-byte loads = coalesced whatever `in` is, e.g. 66), 16 batch rows per pass, shuffle reduction.
__device__ __forceinline__ void dense_fwd(const float* cur, float* nxt, const __nv_bfloat16* W, int in, int out, int B) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int f = warp; f < out; f += kThreads / 32) {
        const __nv_bfloat16* w = W + (long long)f * in;
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

__global__ void __launch_bounds__(kThreads, 1) mlp_small_fwd_kernel(const Args a) {
    extern __shared__ __align__(16) float sm[];
    float* cur = sm;
    float* nxt = sm + kBufFloats;
    dnb::pdl_wait();
    const int B = a.B;
    for (int i = threadIdx.x; i < a.in0 * B; i += kThreads) cur[i] = a.x[i];
    __syncthreads();
    int in = a.in0;
#pragma unroll 1
    for (int l = 0; l < a.L; ++l) {
        const int out = a.width[l];
        dense_fwd(cur, nxt, a.W[l], in, out, B);
        __syncthreads();
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

__global__ void __launch_bounds__(kThreads, 1) mlp_small_bwd_kernel(const Args a) {
    extern __shared__ __align__(16) float sm[];
    float* da = sm;                       // [batch][feature] gradient w.r.t. the current layer's output
    float* dz = sm + kBufFloats;          // [batch][feature] gradient w.r.t. its pre-BatchNorm values
    float* dl = sm + 2 * kBufFloats;      // [batch][class]
    dnb::pdl_wait();
    const int B = a.B, C = a.C, acc = a.accumulate & 1, acc_fc = a.accumulate & 2;
    const int last = a.L > 0 ? a.width[a.L - 1] : a.in0;
    for (int i = threadIdx.x; i < B * C; i += kThreads) dl[i] = a.dlogits[i];
    __syncthreads();
    const float* aL = a.L > 0 ? a.act[a.L - 1] : a.x;
    // fc_out: weight / bias gradients, gradient w.r.t. the last hidden activations
    for (int item = threadIdx.x; item < C * last; item += kThreads) {
        const int k = item % last, c = item / last;
        float s = 0.f;
        for (int b = 0; b < B; ++b) s = fmaf(dl[b * C + c], aL[(long long)b * last + k], s);
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
        // ReLU + BatchNorm backward, one thread per feature
        for (int f = threadIdx.x; f < out; f += kThreads) {
            const float g = a.gamma[l][f], istd = a.invstd[l][f];
            float s1 = 0.f, s2 = 0.f;
            for (int b = 0; b < B; ++b) {
                const float y = a.act[l][(long long)b * out + f];
                const float dy = y > 0.f ? da[b * out + f] : 0.f;
                const float xh = a.xhat[l][(long long)b * out + f];
                s1 += dy;
                s2 = fmaf(dy, xh, s2);
                dz[b * out + f] = dy;
            }
            put(a.dgamma[l] + f, s2, acc);
            put(a.dbeta[l] + f, s1, acc);
            const float k0 = g * istd / (float)B;
            for (int b = 0; b < B; ++b) {
                const float xh = a.xhat[l][(long long)b * out + f];
                dz[b * out + f] = k0 * ((float)B * dz[b * out + f] - s1 - xh * s2);
            }
        }
        __syncthreads();
        // weight gradient dW[f][k] = sum_b dz[b][f] * a_prev[b][k]   (thread = one (f, k); k fastest: coalesced)
        const float* ap = l > 0 ? a.act[l - 1] : a.x;
        for (int item = threadIdx.x; item < out * in; item += kThreads) {
            const int k = item % in, f = item / in;
            float s = 0.f;
#pragma unroll 4
            for (int b = 0; b < B; ++b) s = fmaf(dz[b * out + f], ap[(long long)b * in + k], s);
            put(a.dW[l] + item, s, acc);
        }
        // gradient w.r.t. the previous activations da_prev[b][k] = sum_f dz[b][f] * W[f][k]
        // (thread = column k x 8 batch rows; the weight reads are coalesced over k, dz reads are broadcasts)
        if (l > 0 || a.dx != nullptr) {
            const int nbg = (B + 7) >> 3;
            for (int item = threadIdx.x; item < in * nbg; item += kThreads) {
                const int k = item % in, b0 = (item / in) * 8;
                float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                const __nv_bfloat16* w = a.W[l] + k;
                for (int f = 0; f < out; ++f) {
                    const float wk = bf(w + (long long)f * in);
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

template <typename K>
int launch(K kernel, const Args& a, cudaStream_t st) {
    static bool configured[2] = {false, false};
    const int which = (void*)kernel == (void*)mlp_small_fwd_kernel ? 0 : 1;
    if (!configured[which]) {
        DNB_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        configured[which] = true;
    }
    DNB_CUDA_CHECK(dnb_launch(kernel, dim3(1), dim3(kThreads), kSmemBytes, st, 0, a));
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

}  // namespace

DNB_EXPORT int dnb_mlp_small_forward(const DnbMlpSmallDesc* d, void* stream_) {
    Args a;
    if (int rc = fill(d, a, false)) return rc;
    return launch(mlp_small_fwd_kernel, a, reinterpret_cast<cudaStream_t>(stream_));
}

DNB_EXPORT int dnb_mlp_small_backward(const DnbMlpSmallDesc* d, void* stream_) {
    Args a;
    if (int rc = fill(d, a, true)) return rc;
    return launch(mlp_small_bwd_kernel, a, reinterpret_cast<cudaStream_t>(stream_));
}
