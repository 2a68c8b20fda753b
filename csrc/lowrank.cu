// K3b — orthonormalise the columns of a tall-skinny matrix P [n, r] in place (modified
// Gram-Schmidt, fp32), the step between the two powerSGD all-reduces and inside
// the rankDAD subspace iteration (E3).  r is tiny (<= 16), so this is a latency problem: one CTA
// per matrix, block-wide reductions over n, columns kept in global/L2.  Columns whose residual
// norm collapses (rank-deficient input) are zeroed instead of amplified — same rule as the fp64
// oracle (distrib/lowrank.py: orthogonalize).
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

constexpr int kOrthThreads = 512;

__device__ __forceinline__ float block_sum(float v, float* red) {
    v = dnb::warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = (threadIdx.x < (kOrthThreads >> 5)) ? red[threadIdx.x] : 0.f;
    if (warp == 0) {
        t = dnb::warp_sum(t);
        if (lane == 0) red[0] = t;
    }
    __syncthreads();
    return red[0];
}

struct OrthBatch {
    float* P[DNB_ORTH_MAX_BATCH];
    int n[DNB_ORTH_MAX_BATCH];
    int r[DNB_ORTH_MAX_BATCH];
    int ld[DNB_ORTH_MAX_BATCH];
};

__global__ void __launch_bounds__(kOrthThreads) orth_kernel(const OrthBatch b, float eps) {
    __shared__ float red[32];
    float* P = b.P[blockIdx.x];
    const int n = b.n[blockIdx.x], r = b.r[blockIdx.x], ld = b.ld[blockIdx.x];
    float tot = 0.f;
    for (int i = threadIdx.x; i < n * r; i += kOrthThreads) {
        const float v = P[(size_t)(i / r) * ld + (i % r)];
        tot += v * v;
    }
    const float scale = sqrtf(block_sum(tot, red)) + eps;
    // Norm of every column before any projection: in fp32 the residual of a linearly dependent column
    // is ~1e-7 of its original size, not zero, so "collapsed" is judged relative to where it started.
    __shared__ float orig[DNB_ORTH_MAX_COLS];
    for (int c = 0; c < r; ++c) {
        float s0 = 0.f;
        for (int i = threadIdx.x; i < n; i += kOrthThreads) {
            const float v = P[(size_t)i * ld + c];
            s0 += v * v;
        }
        const float t = sqrtf(block_sum(s0, red));
        if (threadIdx.x == 0) orig[c] = t;
    }
    __syncthreads();
    for (int c = 0; c < r; ++c) {
        float s = 0.f;
        for (int i = threadIdx.x; i < n; i += kOrthThreads) {
            const float v = P[(size_t)i * ld + c];
            s += v * v;
        }
        const float nrm = sqrtf(block_sum(s, red));
        const bool dead = nrm <= fmaxf(eps * scale, 1e-4f * orig[c]);
        const float inv = dead ? 0.f : 1.f / nrm;
        for (int i = threadIdx.x; i < n; i += kOrthThreads) P[(size_t)i * ld + c] *= inv;
        __syncthreads();
        for (int c2 = c + 1; c2 < r; ++c2) {
            float d = 0.f;
            for (int i = threadIdx.x; i < n; i += kOrthThreads) d += P[(size_t)i * ld + c] * P[(size_t)i * ld + c2];
            d = block_sum(d, red);
            for (int i = threadIdx.x; i < n; i += kOrthThreads) P[(size_t)i * ld + c2] -= d * P[(size_t)i * ld + c];
            __syncthreads();
        }
    }
}

}  // namespace

DNB_EXPORT int dnb_orthogonalize_batch(float* const* P, const int* n, const int* r, const int* ld, int count, float eps,
                                       void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(count >= 0 && P && n && r && ld, "dnb_orthogonalize_batch: bad arguments");
    for (int off = 0; off < count; off += DNB_ORTH_MAX_BATCH) {
        OrthBatch b;
        const int m = count - off < DNB_ORTH_MAX_BATCH ? count - off : DNB_ORTH_MAX_BATCH;
        for (int i = 0; i < m; ++i) {
            DNB_REQUIRE(P[off + i] && r[off + i] >= 1 && r[off + i] <= ld[off + i] && r[off + i] <= DNB_ORTH_MAX_COLS,
                        "orthogonalize: bad matrix %d (at most %d columns)", off + i, DNB_ORTH_MAX_COLS);
            b.P[i] = P[off + i]; b.n[i] = n[off + i]; b.r[i] = r[off + i]; b.ld[i] = ld[off + i];
        }
        orth_kernel<<<m, kOrthThreads, 0, st>>>(b, eps);
        DNB_LAUNCH_CHECK();
        dnb_count_launch(1);
    }
    return 0;
}
