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

// ---- shared-memory layouts -----------------------------------------------------------------------------------------
// Activation / gradient planes: fp32, batch-major, row stride lda(width) = round4(width) + 4 floats, round4(B) rows
// (pad rows and columns are zero).  Weight buffers: bf16 [round4(out)][ldw(in)] with ldw(in) = even(in) + 2: the
// 4-byte pad per row puts the rows that the lanes of a warp read at the same time into different banks.
// The first version of these kernels read weights and saved activations from global memory inside the inner loops and
// reduced with shuffles (one warp per output feature): 546 k warp instructions at 0.15 IPC, 121 us for the forward.
// Now everything an inner loop touches is staged into shared memory with cp.async (the next layer's weights while the
// current layer computes, when two buffers fit) and every thread owns a 4 x 4 register tile of the output.
__host__ __device__ __forceinline__ int r4(int n) { return (n + 3) & ~3; }
__host__ __device__ __forceinline__ int lda_of(int width) { return r4(width) + 4; }
__host__ __device__ __forceinline__ int ldw_of(int in) { return ((in + 1) & ~1) + 2; }

// W [out][in] bf16 (global) -> padded shared-memory copy; rows >= out and the odd-width pad column are zeroed
__device__ __forceinline__ void stage_w(__nv_bfloat16* dst, const __nv_bfloat16* W, int out, int in) {
    const int ldw = ldw_of(in), rows = r4(out);
    const bool fast = (in & 1) == 0 && (reinterpret_cast<uintptr_t>(W) & 3) == 0;
    if (fast) {
        const int pairs = in >> 1;
        const uint32_t d0 = dnb::smem_u32(dst);
        for (int i = threadIdx.x; i < out * pairs; i += kThreads) {
            const int f = i / pairs, p = i - f * pairs;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + (f * ldw + 2 * p) * 2),
                         "l"(W + (long long)f * in + 2 * p)
                         : "memory");
        }
    } else {
        for (int i = threadIdx.x; i < out * in; i += kThreads) {
            const int f = i / in, k = i - f * in;
            dst[f * ldw + k] = W[i];
        }
        if (in & 1)
            for (int f = threadIdx.x; f < out; f += kThreads) dst[f * ldw + in] = __float2bfloat16(0.f);
    }
    for (int i = threadIdx.x; i < (rows - out) * ldw; i += kThreads) dst[out * ldw + i] = __float2bfloat16(0.f);
}
// fp32 [B][width] (global) -> padded plane
__device__ __forceinline__ void stage_plane(float* dst, const float* src, int B, int width) {
    const int lda = lda_of(width);
    const bool fast = (width & 1) == 0 && (reinterpret_cast<uintptr_t>(src) & 7) == 0;
    if (fast) {
        const int pairs = width >> 1;
        const uint32_t d0 = dnb::smem_u32(dst);
        for (int i = threadIdx.x; i < B * pairs; i += kThreads) {
            const int b = i / pairs, p = i - b * pairs;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d0 + (b * lda + 2 * p) * 4),
                         "l"(src + (long long)b * width + 2 * p)
                         : "memory");
        }
    } else {
        for (int i = threadIdx.x; i < B * width; i += kThreads) {
            const int b = i / width, k = i - b * width;
            dst[b * lda + k] = src[i];
        }
    }
}
__device__ __forceinline__ void stage_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void stage_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
// zero the pad region of a plane (columns [width, lda) of every row, and rows [B, r4(B)))
__device__ __forceinline__ void zero_pads(float* plane, int B, int width) {
    const int lda = lda_of(width), Bp = r4(B), padc = lda - width;
    for (int i = threadIdx.x; i < B * padc; i += kThreads) plane[(i / padc) * lda + width + i % padc] = 0.f;
    for (int i = threadIdx.x; i < (Bp - B) * lda; i += kThreads) plane[B * lda + i] = 0.f;
}

__device__ __forceinline__ float2 bf2(const __nv_bfloat16* p) {
    const uint32_t w = *reinterpret_cast<const uint32_t*>(p);
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xFFFF0000u));
}

// nxt[b][f] = sum_k cur[b][k] * W[f][k]: thread = 4 features x 4 batch rows, k in pairs (one 32-bit weight load per
// feature, one 64-bit activation load per batch row, 32 FMAs per 8 loads).  Lanes: 8 feature tiles x 4 batch tiles.
__device__ __forceinline__ void dense_fwd(const float* cur, float* nxt, const __nv_bfloat16* Ws, int in, int out, int B) {
    const int lda = lda_of(in), ldo = lda_of(out), ldw = ldw_of(in);
    const int tiles_f = r4(out) >> 2, tiles_b = r4(B) >> 2, kp_n = (in + 1) >> 1;
    const int n = ((tiles_f + 7) >> 3) * 8 * tiles_b;
    for (int t = threadIdx.x; t < n; t += kThreads) {
        const int tf = (t / (8 * tiles_b)) * 8 + (t & 7), tb = (t >> 3) % tiles_b;
        if (tf >= tiles_f) continue;
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        const __nv_bfloat16* w = Ws + (4 * tf) * ldw;
        const float* c = cur + (4 * tb) * lda;
#pragma unroll 2
        for (int kp = 0; kp < kp_n; ++kp) {
            float2 wv[4], cv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) wv[i] = bf2(w + i * ldw + 2 * kp);
#pragma unroll
            for (int j = 0; j < 4; ++j) cv[j] = *reinterpret_cast<const float2*>(c + j * lda + 2 * kp);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(wv[i].x, cv[j].x, fmaf(wv[i].y, cv[j].y, acc[i][j]));
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            *reinterpret_cast<float4*>(nxt + (4 * tb + j) * ldo + 4 * tf) = make_float4(acc[0][j], acc[1][j], acc[2][j], acc[3][j]);
    }
}

// smem: [plane A][plane B][W buffer 0][W buffer 1 (if dbl)]
__global__ void __launch_bounds__(kThreads, 1) mlp_small_fwd_kernel(const Args a, int dbl, int wcap, int plane) {
    extern __shared__ __align__(16) float sm[];
    const int B = a.B;
    float* cur = sm;
    float* nxt = sm + plane;
    __nv_bfloat16* const w0 = reinterpret_cast<__nv_bfloat16*>(sm + 2 * plane);
    __nv_bfloat16* const w1 = dbl ? w0 + wcap : w0;
    dnb::pdl_wait();
    stage_w(w0, a.W[0], a.width[0], a.in0);
    stage_plane(cur, a.x, B, a.in0);
    stage_commit();
    zero_pads(cur, B, a.in0);
    int in = a.in0;
#pragma unroll 1
    for (int l = 0; l < a.L; ++l) {
        const int out = a.width[l];
        const bool more = l + 1 < a.L;
        if (dbl && more) {                       // prefetch the next layer's weights into the other buffer
            stage_w(((l + 1) & 1) ? w1 : w0, a.W[l + 1], a.width[l + 1], out);
            stage_commit();
            stage_wait<1>();
        } else {
            stage_wait<0>();
        }
        __syncthreads();
        dense_fwd(cur, nxt, (dbl && (l & 1)) ? w1 : w0, in, out, B);
        __syncthreads();
        if (!dbl && more) {                      // single buffer: the copy overlaps the BatchNorm phase
            stage_w(w0, a.W[l + 1], a.width[l + 1], out);
            stage_commit();
        }
        // BatchNorm over the batch (biased variance) + ReLU, one thread per feature
        const int ldo = lda_of(out);
        for (int f = threadIdx.x; f < out; f += kThreads) {
            float s = 0.f;
            for (int b = 0; b < B; ++b) s += nxt[b * ldo + f];
            const float mean = s / (float)B;
            float v = 0.f;
            for (int b = 0; b < B; ++b) {
                const float d = nxt[b * ldo + f] - mean;
                v = fmaf(d, d, v);
            }
            const float istd = rsqrtf(v / (float)B + a.eps);
            const float g = a.gamma[l][f], be = a.beta[l][f];
            a.invstd[l][f] = istd;
            for (int b = 0; b < B; ++b) {
                const float xh = (nxt[b * ldo + f] - mean) * istd;
                const float y = fmaxf(fmaf(xh, g, be), 0.f);
                nxt[b * ldo + f] = y;
                a.xhat[l][(long long)b * out + f] = xh;
                a.act[l][(long long)b * out + f] = y;
            }
        }
        __syncthreads();
        zero_pads(nxt, B, out);                  // the dense tiles wrote (finite) values into pad rows / columns
        float* t = cur; cur = nxt; nxt = t;
        in = out;
    }
    __syncthreads();
    // fc_out (with bias): one warp per (class, batch row), lanes along k
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lda = lda_of(in);
    for (int item = warp; item < a.C * B; item += kThreads / 32) {
        const int c = item % a.C, b = item / a.C;
        const __nv_bfloat16* w = a.Wfc + (long long)c * in;
        float s = 0.f;
        for (int k = lane; k < in; k += 32) s = fmaf(bf(w + k), cur[b * lda + k], s);
        s = dnb::warp_sum(s);
        if (lane == 0) a.logits[(long long)b * a.C + c] = s + a.bfc[c];
    }
}

// smem: [da][dz][dl 4 KB][W buffer 0][W 1 (if dbl)][a_prev plane 0][a_prev plane 1 (if dbl)]
__global__ void __launch_bounds__(kThreads, 1) mlp_small_bwd_kernel(const Args a, int dbl, int wcap, int plane) {
    extern __shared__ __align__(16) float sm[];
    const int B = a.B, C = a.C, acc = a.accumulate & 1, acc_fc = a.accumulate & 2;
    float* da = sm;                       // gradient w.r.t. the current layer's output (plane layout)
    float* dz = sm + plane;               // gradient w.r.t. its pre-BatchNorm values
    float* dl = sm + 2 * plane;           // [batch][class]
    __nv_bfloat16* const w0 = reinterpret_cast<__nv_bfloat16*>(dl + kMaxB * kMaxC);
    __nv_bfloat16* const w1 = dbl ? w0 + wcap : w0;
    float* const ap0 = reinterpret_cast<float*>(w0 + (dbl ? 2 : 1) * wcap);
    float* const ap1 = dbl ? ap0 + plane : ap0;
    dnb::pdl_wait();
    const int last = a.width[a.L - 1];
    auto stage_layer = [&](int l, int slot) {
        const int in = l > 0 ? a.width[l - 1] : a.in0;
        stage_w(slot ? w1 : w0, a.W[l], a.width[l], in);
        stage_plane(slot ? ap1 : ap0, l > 0 ? a.act[l - 1] : a.x, B, in);
        stage_commit();                                                 // one group per layer
    };
    stage_layer(a.L - 1, dbl ? ((a.L - 1) & 1) : 0);
    for (int i = threadIdx.x; i < B * C; i += kThreads) dl[i] = a.dlogits[i];
    __syncthreads();
    const float* aL = a.act[a.L - 1];
    const int ldl = lda_of(last);
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
        da[b * ldl + k] = s;
    }
    __syncthreads();
#pragma unroll 1
    for (int l = a.L - 1; l >= 0; --l) {
        const int out = a.width[l], in = l > 0 ? a.width[l - 1] : a.in0;
        const int ldo = lda_of(out), ldi = lda_of(in), ldw = ldw_of(in);
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
                const float dy = y > 0.f ? da[b * ldo + f] : 0.f;
                s1 += dy;
                s2 = fmaf(dy, xh, s2);
                dz[b * ldo + f] = dy;
            }
            put(a.dgamma[l] + f, s2, acc);
            put(a.dbeta[l] + f, s1, acc);
            const float k0 = g * istd / (float)B;
#pragma unroll 8
            for (int b = 0; b < B; ++b) {
                const float xh = __ldg(a.xhat[l] + (long long)b * out + f);
                dz[b * ldo + f] = k0 * ((float)B * dz[b * ldo + f] - s1 - xh * s2);
            }
        }
        zero_pads(dz, B, out);
        if (dbl && l > 0) stage_wait<1>();
        else stage_wait<0>();
        __syncthreads();
        // weight gradient dW[f][k] = sum_b dz[b][f] * a_prev[b][k]: thread = 4 features x 4 inputs, two 16-byte loads
        // and 16 FMAs per batch row; consecutive threads own consecutive input tiles (coalesced 16-byte stores)
        {
            const float* ap = slot ? ap1 : ap0;
            const int tiles_k = r4(in) >> 2, tiles_f = r4(out) >> 2;
            float* dW = a.dW[l];
            const bool vec = (in & 3) == 0 && (reinterpret_cast<uintptr_t>(dW) & 15) == 0;
            for (int t = threadIdx.x; t < tiles_k * tiles_f; t += kThreads) {
                const int tk = t % tiles_k, tf = t / tiles_k;
                float s[4][4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) s[i][j] = 0.f;
#pragma unroll 4
                for (int b = 0; b < B; ++b) {
                    const float4 z = *reinterpret_cast<const float4*>(dz + b * ldo + 4 * tf);
                    const float4 x = *reinterpret_cast<const float4*>(ap + b * ldi + 4 * tk);
                    const float zv[4] = {z.x, z.y, z.z, z.w}, xv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) s[i][j] = fmaf(zv[i], xv[j], s[i][j]);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int f = 4 * tf + i;
                    if (f >= out) continue;
                    float* o = dW + (long long)f * in + 4 * tk;
                    if (vec && !acc) {
                        *reinterpret_cast<float4*>(o) = make_float4(s[i][0], s[i][1], s[i][2], s[i][3]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (4 * tk + j < in) put(o + j, s[i][j], acc);
                    }
                }
            }
        }
        // gradient w.r.t. the previous activations da_prev[b][k] = sum_f dz[b][f] * W[f][k]: thread = 4 batch rows x
        // 2 inputs (one 32-bit weight load per feature; the dz reads are broadcasts)
        if (l > 0 || a.dx != nullptr) {
            const __nv_bfloat16* Ws = slot ? w1 : w0;
            const int kp_n = (in + 1) >> 1, tiles_b = r4(B) >> 2;
            for (int t = threadIdx.x; t < kp_n * tiles_b; t += kThreads) {
                const int kp = t % kp_n, tb = t / kp_n;
                float s[4][2];
#pragma unroll
                for (int j = 0; j < 4; ++j) s[j][0] = s[j][1] = 0.f;
                const float* z = dz + (4 * tb) * ldo;
#pragma unroll 4
                for (int f = 0; f < out; ++f) {
                    const float2 wv = bf2(Ws + f * ldw + 2 * kp);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float zv = z[j * ldo + f];
                        s[j][0] = fmaf(wv.x, zv, s[j][0]);
                        s[j][1] = fmaf(wv.y, zv, s[j][1]);
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) *reinterpret_cast<float2*>(da + (4 * tb + j) * ldi + 2 * kp) = make_float2(s[j][0], s[j][1]);
            }
        }
        __syncthreads();
        if (!dbl && l > 0) stage_layer(l - 1, 0);
    }
    if (a.dx != nullptr) {
        const int ld0 = lda_of(a.in0);
        for (int i = threadIdx.x; i < B * a.in0; i += kThreads) a.dx[i] = da[(i / a.in0) * ld0 + i % a.in0];
    }
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

// bf16 elements one padded weight buffer must hold (largest layer of THIS network) and floats of one padded plane
int weight_capacity(const Args& a) {
    int cap = 0, in = a.in0;
    for (int l = 0; l < a.L; ++l) {
        const int n = r4(a.width[l]) * ldw_of(in);
        if (n > cap) cap = n;
        in = a.width[l];
    }
    return (cap + 7) & ~7;
}
int plane_floats(const Args& a) {
    int w = a.in0;
    for (int l = 0; l < a.L; ++l)
        if (a.width[l] > w) w = a.width[l];
    return r4(a.B) * lda_of(w);
}

int launch_fwd(const Args& a, cudaStream_t st) {
    const int plane = plane_floats(a), wcap = weight_capacity(a), kWBytes = wcap * 2;
    int dbl = 1;
    int need = 2 * plane * 4 + 2 * kWBytes;
    if (need > kSmemLimit) { dbl = 0; need = 2 * plane * 4 + kWBytes; }
    DNB_REQUIRE(need <= kSmemLimit, "mlp_small forward: batch %d with these widths needs %d bytes of shared memory", a.B, need);
    static bool configured = false;
    if (!configured) {
        DNB_CUDA_CHECK(cudaFuncSetAttribute(mlp_small_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
        configured = true;
    }
    DNB_CUDA_CHECK(dnb_launch(mlp_small_fwd_kernel, dim3(1), dim3(kThreads), (size_t)need, st, 0, a, dbl, wcap, plane));
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

int launch_bwd(const Args& a, cudaStream_t st) {
    const int plane = plane_floats(a), wcap = weight_capacity(a), kWBytes = wcap * 2;
    const int fixed = 2 * plane * 4 + kMaxB * kMaxC * 4;
    int dbl = 1;
    int need = fixed + 2 * kWBytes + 2 * plane * 4;
    if (need > kSmemLimit) { dbl = 0; need = fixed + kWBytes + plane * 4; }
    DNB_REQUIRE(need <= kSmemLimit, "mlp_small backward: batch %d with these widths needs %d bytes of shared memory", a.B, need);
    static bool configured = false;
    if (!configured) {
        DNB_CUDA_CHECK(cudaFuncSetAttribute(mlp_small_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
        configured = true;
    }
    DNB_CUDA_CHECK(dnb_launch(mlp_small_bwd_kernel, dim3(1), dim3(kThreads), (size_t)need, st, 0, a, dbl, wcap, plane));
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
