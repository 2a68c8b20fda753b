// Layout / cast kernels around the GEMMs, K10 window gather, K6 softmax + cross-entropy.
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

// x fp32 [B, S, F]  ->  rows bf16 [S*B, ldr]  (row = t*B + b, "time-major")
//                   and optionally colsT bf16 [F, ldt] (colsT[f][t*B+b]) — the K-major operand of
//                   the encoder weight-gradient GEMM.  32x32 smem tile transpose.
__global__ void cast_time_major_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ rows, long long ldr,
                                       __nv_bfloat16* __restrict__ colsT, long long ldt, int B, int S, int F) {
    __shared__ float tile[32][33];
    dnb::pdl_wait();
    dnb::pdl_launch();
    const int r0 = blockIdx.y * 32, f0 = blockIdx.x * 32;
    const int R = B * S;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, f = f0 + threadIdx.x;       // r = t*B + b
        float v = 0.f;
        if (r < R && f < F) {
            const int t = r / B, b = r % B;
            v = x[((long long)b * S + t) * F + f];
            rows[(long long)r * ldr + f] = __float2bfloat16(v);
        }
        tile[i][threadIdx.x] = v;
    }
    if (colsT == nullptr) return;
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int f = f0 + i, r = r0 + threadIdx.x;
        if (f < F && r < R) colsT[(long long)f * ldt + r] = __float2bfloat16(tile[threadIdx.x][i]);
    }
}

// rows only (no transposed copy), F % 8 == 0: each thread converts 8 consecutive features — two 16-byte loads,
// one 16-byte store; consecutive threads walk along a row, so both sides are fully coalesced
__global__ void cast_rows_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ rows, long long ldr, int B, int S,
                                 int F8) {
    dnb::pdl_wait();
    dnb::pdl_launch();
    const long long total = (long long)B * S * F8;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % F8);
        const long long r = i / F8;                       // r = t*B + b
        const int t = (int)(r / B), b = (int)(r % B);
        const float4* src = reinterpret_cast<const float4*>(x + (((long long)b * S + t) * F8 + c) * 8);
        const float4 v0 = src[0], v1 = src[1];
        uint4 o;
        o.x = dnb::pack_bf16x2(v0.x, v0.y); o.y = dnb::pack_bf16x2(v0.z, v0.w);
        o.z = dnb::pack_bf16x2(v1.x, v1.y); o.w = dnb::pack_bf16x2(v1.z, v1.w);
        *reinterpret_cast<uint4*>(rows + r * ldr + (long long)c * 8) = o;
    }
}

__global__ void transpose_bf16_kernel(const __nv_bfloat16* __restrict__ in, long long ldi,
                                      __nv_bfloat16* __restrict__ out, long long ldo, int R, int C) {
    __shared__ __nv_bfloat16 tile[32][34];
    dnb::pdl_wait();
    dnb::pdl_launch();
    const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < R && c < C) ? in[(long long)r * ldi + c] : __float2bfloat16(0.f);
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (c < C && r < R) out[(long long)c * ldo + r] = tile[threadIdx.x][i];
    }
}

// 64 x 64 tiles moved as bf16 pairs (4-byte accesses on both sides, 128 bytes per warp row): the 2-byte
// version above moves half as much per instruction and is kept for odd shapes / unaligned operands.
__global__ void __launch_bounds__(256) transpose_bf16_x2_kernel(const __nv_bfloat16* __restrict__ in, long long ldi,
                                                                __nv_bfloat16* __restrict__ out, long long ldo, int R, int C) {
    __shared__ uint32_t tile[64][33];                 // [row][column pair]
    dnb::pdl_wait();
    dnb::pdl_launch();
    const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 32 x 8
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = r0 + ty + 8 * i, c = c0 + 2 * tx;
        uint32_t v = 0;
        if (r < R && c < C) v = *reinterpret_cast<const uint32_t*>(in + (long long)r * ldi + c);   // C even: pair in range
        tile[ty + 8 * i][tx] = v;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int cl = ty + 8 * i;                    // local column 0..63
        const int c = c0 + cl, r = r0 + 2 * tx;       // output row c, output columns r, r+1
        if (c < C && r < R) {
            const uint32_t a = tile[2 * tx][cl >> 1], b = tile[2 * tx + 1][cl >> 1];
            const uint32_t lo = (cl & 1) ? (a >> 16) : (a & 0xFFFFu);
            const uint32_t hi = (cl & 1) ? (b >> 16) : (b & 0xFFFFu);
            if (r + 1 < R) *reinterpret_cast<uint32_t*>(out + (long long)c * ldo + r) = lo | (hi << 16);
            else out[(long long)c * ldo + r] = __ushort_as_bfloat16((unsigned short)lo);
        }
    }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
    long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const long long stride = (long long)gridDim.x * blockDim.x * 4;
    for (; i + 3 < n; i += stride) {
        const float4 v = *reinterpret_cast<const float4*>(in + i);
        uint2 p;
        p.x = dnb::pack_bf16x2(v.x, v.y);
        p.y = dnb::pack_bf16x2(v.z, v.w);
        *reinterpret_cast<uint2*>(out + i) = p;
    }
    if (i < n && i + 3 >= n)
        for (long long k = i; k < n; ++k) out[k] = __float2bfloat16(in[k]);
}

// K10: [N, C, T] -> [N, S, C, W], window j = data[:, :, j*stride : j*stride+W]
// (reference: the Python double loop in comps/icalstm/__init__.py:29-32)
__global__ void window_gather_kernel(const float* __restrict__ in, float* __restrict__ out, int N, int C, int T,
                                     int S, int W, int stride) {
    const long long total = (long long)N * S * C * W;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int w = i % W;
        const int c = (i / W) % C;
        const int s = (i / ((long long)W * C)) % S;
        const int n = i / ((long long)W * C * S);
        out[i] = in[((long long)n * C + c) * T + s * stride + w];
    }
}

// Device-resident batches: gather B rows of a feature table (and their labels) selected by a permutation that lives
// on the device.  state = {cursor, blocks-done counter}: every launch reads rows perm[(cursor + b) % n_perm] and the
// last block advances the cursor by B, so the identical launch can be replayed from a CUDA graph and still walks
// through the epoch.  Replaces DataLoader + default_collate + a host-to-device copy per step
// (reference comps/icalstm/__init__.py:36-38, comps/fs/__init__.py:33-39).
__global__ void __launch_bounds__(256) batch_gather_kernel(const float* __restrict__ table, long long row_floats,
                                                           const long long* __restrict__ labels,
                                                           const long long* __restrict__ perm, long long n_perm,
                                                           unsigned long long* state, int B, float* __restrict__ out,
                                                           long long* __restrict__ labels_out,
                                                           long long* __restrict__ index_out) {
    dnb::pdl_wait();
    const unsigned long long cursor = *reinterpret_cast<volatile unsigned long long*>(state);
    const int b = blockIdx.y;
    const long long row = perm[(cursor + (unsigned long long)b) % (unsigned long long)n_perm];
    const float* src = table + row * row_floats;
    float* dst = out + (long long)b * row_floats;
    const long long i0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const long long stride = (long long)gridDim.x * blockDim.x * 4;
    if ((row_floats & 3) == 0 && ((reinterpret_cast<uintptr_t>(table) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
        for (long long i = i0; i < row_floats; i += stride)
            *reinterpret_cast<float4*>(dst + i) = __ldcs(reinterpret_cast<const float4*>(src + i));
    } else {
        for (long long i = i0; i < row_floats; i += stride)
            for (int k = 0; k < 4 && i + k < row_floats; ++k) dst[i + k] = src[i + k];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (labels_out && labels) labels_out[b] = labels[row];
        if (index_out) index_out[b] = row;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned long long prev = atomicAdd(state + 1, 1ull);
        if (prev == (unsigned long long)gridDim.x * gridDim.y - 1) {
            state[1] = 0ull;
            state[0] = cursor + (unsigned long long)B;
            __threadfence();
        }
    }
}

// K6: softmax + cross-entropy (mean over the batch) forward, and its gradient, in one pass.
// One warp per row; classes strided over lanes.  Small batches (the federated sites' 16-64 rows) run as ONE
// CTA that loops over the rows and reduces the loss on-chip: no memset node, no atomics.  Larger batches
// use one atomic per row into a zeroed scalar.
__global__ void softmax_ce_small_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                        float* __restrict__ prob, float* __restrict__ dlogits, float* __restrict__ loss,
                                        int B, int ncls, int log_softmax_out) {
    __shared__ float part[8];
    dnb::pdl_wait();
    dnb::pdl_launch();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const float invB = 1.f / (float)B;
    float acc = 0.f;
    for (int row = warp; row < B; row += nw) {
        const float* z = logits + (long long)row * ncls;
        float mx = -INFINITY;
        for (int c = lane; c < ncls; c += 32) mx = fmaxf(mx, z[c]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
        for (int c = lane; c < ncls; c += 32) sum += __expf(z[c] - mx);
        sum = dnb::warp_sum(sum);
        const float lse = mx + __logf(sum);
        const int y = (int)labels[row];
        for (int c = lane; c < ncls; c += 32) {
            const float p = __expf(z[c] - lse);
            if (prob) prob[(long long)row * ncls + c] = log_softmax_out ? (z[c] - lse) : p;
            if (dlogits) dlogits[(long long)row * ncls + c] = (p - (c == y ? 1.f : 0.f)) * invB;
        }
        acc += (lse - z[y]) * invB;
    }
    if (lane == 0) part[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0 && loss) {
        float s = 0.f;
        for (int w = 0; w < nw; ++w) s += part[w];
        *loss = s;
    }
}

__global__ void softmax_ce_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                  float* __restrict__ prob, float* __restrict__ dlogits, float* __restrict__ loss,
                                  int B, int ncls, int log_softmax_out) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= B) return;
    const float* z = logits + (long long)row * ncls;
    float mx = -INFINITY;
    for (int c = lane; c < ncls; c += 32) mx = fmaxf(mx, z[c]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    for (int c = lane; c < ncls; c += 32) sum += __expf(z[c] - mx);
    sum = dnb::warp_sum(sum);
    const float lse = mx + __logf(sum);
    const int y = (int)labels[row];
    const float invB = 1.f / (float)B;
    for (int c = lane; c < ncls; c += 32) {
        const float p = __expf(z[c] - lse);
        if (prob) prob[(long long)row * ncls + c] = log_softmax_out ? (z[c] - lse) : p;
        if (dlogits) dlogits[(long long)row * ncls + c] = (p - (c == y ? 1.f : 0.f)) * invB;
    }
    if (lane == 0 && loss) atomicAdd(loss, (lse - z[y]) * invB);
}

// K9: device-side metric accumulators.  One tiny launch per step appends the batch's positive-class
// scores and labels to a device buffer (ROC-AUC needs every pair), bumps the confusion counts at the
// given threshold and the running loss sum — no host sync in the training loop (the reference pays a
// loss.item() per iteration, comps/fs/__init__.py:61).  state = {cursor, tp, fp, tn, fn}; lossacc = {sum, n}.
__global__ void metrics_append_kernel(const float* __restrict__ scores_in, long long ld, int col,
                                      const long long* __restrict__ labels, const float* __restrict__ loss, int B,
                                      float* __restrict__ scores, int* __restrict__ labs, int cap,
                                      unsigned long long* __restrict__ state, float* __restrict__ lossacc, float thr) {
    const unsigned long long cur = state[0];
    __syncthreads();
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        const float s = scores_in[(long long)b * ld + col];
        const int y = (int)labels[b];
        const unsigned long long pos = cur + b;
        if (pos < (unsigned long long)cap) {
            scores[pos] = s;
            labs[pos] = y;
        }
        const int p = s >= thr;
        atomicAdd(&state[1 + (p ? (y ? 0 : 1) : (y ? 3 : 2))], 1ull);      // tp, fp, tn, fn
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        state[0] = cur + B;
        if (loss) {
            lossacc[0] += loss[0] * (float)B;
            lossacc[1] += (float)B;
        }
    }
}

}  // namespace

DNB_EXPORT int dnb_metrics_append(const float* scores_in, long long ld, int col, const long long* labels,
                                  const float* loss, int B, float* scores, int* labs, int cap,
                                  unsigned long long* state, float* lossacc, float thr, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(scores_in && labels && scores && labs && state && lossacc && B > 0 && cap > 0,
                "dnb_metrics_append: bad arguments");
    metrics_append_kernel<<<1, 128, 0, st>>>(scores_in, ld, col, labels, loss, B, scores, labs, cap, state, lossacc, thr);
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

DNB_EXPORT int dnb_cast_time_major(const float* x, void* rows, long long ldr, void* colsT, long long ldt, int B, int S,
                                   int F, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(x && rows && B > 0 && S > 0 && F > 0, "dnb_cast_time_major: bad arguments");
    if (colsT == nullptr && F % 8 == 0 && ldr % 8 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)rows & 15) == 0) {
        const long long total = (long long)B * S * (F / 8);
        const int blocks = (int)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
        DNB_CUDA_CHECK(dnb_launch(cast_rows_kernel, dim3(blocks), dim3(256), 0, st, 0, x, reinterpret_cast<__nv_bfloat16*>(rows), ldr,
                                  B, S, F / 8));
        dnb_count_launch(1);
        return 0;
    }
    dim3 grid((F + 31) / 32, (B * S + 31) / 32), block(32, 8);
    DNB_CUDA_CHECK(dnb_launch(cast_time_major_kernel, grid, block, 0, st, 0, x, reinterpret_cast<__nv_bfloat16*>(rows), ldr,
                              reinterpret_cast<__nv_bfloat16*>(colsT), ldt, B, S, F));
    dnb_count_launch(1);
    return 0;
}

DNB_EXPORT int dnb_transpose_bf16(const void* in, long long ldi, void* out, long long ldo, int R, int C, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(in && out && R > 0 && C > 0, "dnb_transpose_bf16: bad arguments");
    if (C % 2 == 0 && ldi % 2 == 0 && ldo % 2 == 0 && ((uintptr_t)in & 3) == 0 && ((uintptr_t)out & 3) == 0) {
        dim3 grid2((C + 63) / 64, (R + 63) / 64);
        DNB_CUDA_CHECK(dnb_launch(transpose_bf16_x2_kernel, grid2, dim3(256), 0, st, 0, reinterpret_cast<const __nv_bfloat16*>(in),
                                  ldi, reinterpret_cast<__nv_bfloat16*>(out), ldo, R, C));
        dnb_count_launch(1);
        return 0;
    }
    dim3 grid((C + 31) / 32, (R + 31) / 32), block(32, 8);
    DNB_CUDA_CHECK(dnb_launch(transpose_bf16_kernel, grid, block, 0, st, 0, reinterpret_cast<const __nv_bfloat16*>(in), ldi,
                              reinterpret_cast<__nv_bfloat16*>(out), ldo, R, C));
    dnb_count_launch(1);
    return 0;
}

DNB_EXPORT int dnb_cast_f32_bf16(const float* in, void* out, long long n, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(in && out && n >= 0, "dnb_cast_f32_bf16: bad arguments");
    DNB_REQUIRE(((uintptr_t)in & 15) == 0 && ((uintptr_t)out & 7) == 0, "dnb_cast_f32_bf16: unaligned buffers");
    if (n == 0) return 0;
    int blocks = dnb_ceil_div(n, 256 * 4);
    if (blocks > 148 * 8) blocks = 148 * 8;
    cast_f32_bf16_kernel<<<blocks, 256, 0, st>>>(in, reinterpret_cast<__nv_bfloat16*>(out), n);
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

DNB_EXPORT int dnb_window_gather(const float* in, float* out, int N, int C, int T, int S, int W, int stride,
                                 void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(in && out, "dnb_window_gather: null buffer");
    DNB_REQUIRE((long long)(S - 1) * stride + W <= T, "dnb_window_gather: windows exceed the time axis");
    const long long total = (long long)N * S * C * W;
    int blocks = dnb_ceil_div(total, 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    window_gather_kernel<<<blocks, 256, 0, st>>>(in, out, N, C, T, S, W, stride);
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

DNB_EXPORT int dnb_batch_gather(const float* table, long long row_floats, const long long* labels, const long long* perm,
                                long long n_perm, unsigned long long* state, int B, float* out, long long* labels_out,
                                long long* index_out, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(table && perm && state && out && row_floats > 0 && n_perm > 0 && B > 0 && B <= 65535,
                "dnb_batch_gather: bad arguments");
    int bx = (int)((row_floats / 4 + 255) / 256);
    if (bx < 1) bx = 1;
    if (bx > 64) bx = 64;
    DNB_CUDA_CHECK(dnb_launch(batch_gather_kernel, dim3(bx, B), dim3(256), 0, st, 0, table, row_floats, labels, perm, n_perm,
                              state, B, out, labels_out, index_out));
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

DNB_EXPORT int dnb_softmax_ce(const float* logits, const long long* labels, float* prob, float* dlogits, float* loss,
                              int B, int ncls, int log_softmax_out, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(logits && labels && B > 0 && ncls > 0, "dnb_softmax_ce: bad arguments");
    if (B <= 256) {
        DNB_CUDA_CHECK(dnb_launch(softmax_ce_small_kernel, dim3(1), dim3(256), 0, st, 0, logits, labels, prob, dlogits, loss, B,
                                  ncls, log_softmax_out));
    } else {
        if (loss) DNB_CUDA_CHECK(cudaMemsetAsync(loss, 0, sizeof(float), st));
        softmax_ce_kernel<<<(B + 3) / 4, 128, 0, st>>>(logits, labels, prob, dlogits, loss, B, ncls, log_softmax_out);
        DNB_LAUNCH_CHECK();
    }
    dnb_count_launch(1);
    return 0;
}
