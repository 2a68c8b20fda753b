// K1 — dSGD aggregation fused with the optimiser (engine semantics §2.3 E2).
//
// One launch per training step on every GPU (= site).  Each GPU's gradient buffer, parameter
// buffer and signal pad live in a symmetric arena that every peer has mapped (CUDA IPC over
// NVLink/NVSwitch), so the kernel itself pulls the peers' gradients and pushes the updated
// parameters; no NCCL call and no host round-trip sit between backward and the next forward.
//
//   start barrier  (system-scope release/acquire flags, one slot per (block, peer))
//   two-shot mode : rank r owns shard r.  For its shard it sums the S sites' gradients in fixed
//                   rank order, scales by 1/(S*loss_scale) (the reference's unweighted mean over
//                   sites), applies Adam/SGD on the fp32 master values, and stores the *updated
//                   parameters* (fp32, plus an optional bf16 shadow for the tensor-core kernels)
//                   into every peer's parameter buffer.  Optimiser state is only ever touched
//                   for the local shard.
//   one-shot mode : for tiny payloads (FreeSurfer MLP: 0.24 MB) every rank reduces the whole
//                   vector itself and updates only its own copy; one NVLink traversal, no
//                   parameter broadcast.  Summation order is the same on all ranks, so all
//                   sites stay bit-identical.
//   end barrier    (peers' pushes have landed; gradient buffers may be overwritten)
//
// Wire format of the gradients: fp32, bf16 or fp16 (`precision_bits`, compspec.json:161-176).
// opt = 2 turns the optimiser off and writes the reduced mean to `out` (all-reduce of the
// powerSGD P/Q factors, aggregation micro-benchmark).
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

constexpr int kMaxRanks = DNB_MAX_RANKS;
constexpr int kMaxBlocks = DNB_SIGNAL_BLOCKS;
constexpr int kThreads = 512;

struct DsgdArgs {
    const void* grads[kMaxRanks];
    float* params[kMaxRanks];
    __nv_bfloat16* shadow[kMaxRanks];
    uint32_t* flags[kMaxRanks];
    float* m;
    float* v;
    float* out;
    long long n;
    int nranks, rank, mode, opt, use_flags;
    float lr, beta1, beta2, eps, wd, bc1, bc2, grad_scale, momentum;
    uint32_t epoch;
    uint32_t* dev_state;   // optional {completed steps, blocks-done counter, lr bits}: makes the launch replayable
    float* stage;          // optional fp32 local gradients to cast into grads[rank] in-kernel (16-bit wire)
    int zero_stage;
    // mode 2 (NVLS): multicast addresses of the same symmetric buffers.  A load through grads_mc is reduced over all
    // sites inside the NVSwitch, a store through params_mc / shadow_mc lands on every site.
    const void* grads_mc;
    float* params_mc;
    __nv_bfloat16* shadow_mc;
};

// ---- NVLS (multimem) primitives: 16 bytes per instruction -------------------------------------------------
template <int WIRE>
__device__ __forceinline__ void mc_load_reduce8(const void* mc_base, long long i, float (&x)[8]) {
    if constexpr (WIRE == 0) {
        const float* p = reinterpret_cast<const float*>(mc_base) + i;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(x[0]), "=f"(x[1]), "=f"(x[2]), "=f"(x[3]) : "l"(p) : "memory");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(x[4]), "=f"(x[5]), "=f"(x[6]), "=f"(x[7]) : "l"(p + 4) : "memory");
    } else {
        const uint16_t* p = reinterpret_cast<const uint16_t*>(mc_base) + i;
        uint32_t w[4];
        if constexpr (WIRE == 1)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                         : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p) : "memory");
        else
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                         : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p) : "memory");
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if constexpr (WIRE == 1) {
                x[2 * k] = __uint_as_float(w[k] << 16);
                x[2 * k + 1] = __uint_as_float(w[k] & 0xFFFF0000u);
            } else {
                __half2 h = *reinterpret_cast<const __half2*>(&w[k]);
                float2 f = __half22float2(h);
                x[2 * k] = f.x;
                x[2 * k + 1] = f.y;
            }
        }
    }
}
__device__ __forceinline__ void mc_store16(void* mc_addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(mc_addr), "f"(__uint_as_float(a)), "f"(__uint_as_float(b)), "f"(__uint_as_float(c)),
                 "f"(__uint_as_float(d))
                 : "memory");
}

__device__ __forceinline__ void flag_barrier(const DsgdArgs& a, int phase, uint32_t epoch) {
    // Every block pairs up with the same-index block on each peer.
    __syncthreads();
    if (threadIdx.x < a.nranks) {
        const int peer = threadIdx.x;
        const size_t slot = ((size_t)phase * kMaxBlocks + blockIdx.x) * kMaxRanks;
        dnb::st_release_sys(a.flags[peer] + slot + a.rank, epoch);
        const uint32_t* mine = a.flags[a.rank] + slot + peer;
        const long long t0 = clock64();
        while ((int32_t)(dnb::ld_acquire_sys(mine) - epoch) < 0) {
            if (clock64() - t0 > dnb::kSpinBudgetCycles) {
                printf("dnb: dSGD flag barrier timed out (rank %d waits for %d, block %d, phase %d, epoch %u)\n",
                       a.rank, peer, blockIdx.x, phase, epoch);
                __trap();
            }
        }
    }
    __syncthreads();
}

template <int WIRE>
__device__ __forceinline__ void load8(const void* base, long long i, float (&x)[8]) {
    if constexpr (WIRE == 0) {
        const float4* p = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + i);
        float4 a = p[0], b = p[1];
        x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
        x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
    } else {
        uint4 raw = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(base) + i);
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if constexpr (WIRE == 1) {
                x[2 * k] = __uint_as_float(w[k] << 16);
                x[2 * k + 1] = __uint_as_float(w[k] & 0xFFFF0000u);
            } else {
                __half2 h = *reinterpret_cast<const __half2*>(&w[k]);
                float2 f = __half22float2(h);
                x[2 * k] = f.x;
                x[2 * k + 1] = f.y;
            }
        }
    }
}

template <int WIRE>
__device__ __forceinline__ void store8_wire(void* base, long long i, const float (&x)[8]) {
    uint4 o;
    if constexpr (WIRE == 1) {
        o.x = dnb::pack_bf16x2(x[0], x[1]); o.y = dnb::pack_bf16x2(x[2], x[3]);
        o.z = dnb::pack_bf16x2(x[4], x[5]); o.w = dnb::pack_bf16x2(x[6], x[7]);
    } else {
        __half2 h0 = __floats2half2_rn(x[0], x[1]), h1 = __floats2half2_rn(x[2], x[3]);
        __half2 h2 = __floats2half2_rn(x[4], x[5]), h3 = __floats2half2_rn(x[6], x[7]);
        o.x = *reinterpret_cast<uint32_t*>(&h0); o.y = *reinterpret_cast<uint32_t*>(&h1);
        o.z = *reinterpret_cast<uint32_t*>(&h2); o.w = *reinterpret_cast<uint32_t*>(&h3);
    }
    *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(base) + i) = o;
}

__device__ __forceinline__ void store8_f32(float* base, long long i, const float (&x)[8]) {
    float4* p = reinterpret_cast<float4*>(base + i);
    p[0] = make_float4(x[0], x[1], x[2], x[3]);
    p[1] = make_float4(x[4], x[5], x[6], x[7]);
}

// NR = number of sites as a compile-time constant (1, 2, 4, 8: the peer loop unrolls exactly and the staging
// registers are sized for it) or 0 = run-time count (any 1..8, guarded loop).
template <int WIRE, int NR, bool NVLS>
__global__ void __launch_bounds__(kThreads, 1) dsgd_fused_kernel(const DsgdArgs a) {
    const int nranks = NR > 0 ? NR : a.nranks;
    // Step-dependent scalars come either from the launch arguments or — when the launch lives in a
    // CUDA graph and is replayed unchanged — from a small device-resident state block.
    dnb::pdl_wait();
    uint32_t epoch = a.epoch;
    float lr = a.lr, bc1 = a.bc1, bc2 = a.bc2;
    uint32_t t_dev = 0;
    if (a.dev_state) {
        t_dev = *reinterpret_cast<volatile uint32_t*>(a.dev_state) + 1u;
        epoch = a.epoch + t_dev;
        lr = __uint_as_float(*reinterpret_cast<volatile uint32_t*>(a.dev_state + 2));
        bc1 = 1.f - powf(a.beta1, (float)t_dev);
        bc2 = sqrtf(1.f - powf(a.beta2, (float)t_dev));
    }
    const long long stride = (long long)gridDim.x * blockDim.x * 8;
    if constexpr (WIRE != 0) {
        if (a.stage != nullptr) {
            // Cast this rank's fp32 gradients into its peer-visible wire buffer.  Block b packs, for every
            // peer's shard, exactly the elements that peer's block b will read (same indexing as the reduce
            // loop below), so the per-block flag barrier is all the ordering the peers need; the local reduce
            // re-reads only what the same thread wrote.
            void* wire = const_cast<void*>(a.grads[a.rank]);
            const int shards = (a.mode != 1 && nranks > 1) ? nranks : 1;
            const long long per = shards > 1 ? ((a.n / 8 + nranks - 1) / nranks) * 8 : a.n;
            for (int p = 0; p < shards; ++p) {
                long long plo = per * p, phi = plo + per < a.n ? plo + per : a.n;
                if (plo > phi) plo = phi;
                for (long long i = plo + ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8; i < phi; i += stride) {
                    float x[8];
                    load8<0>(a.stage, i, x);
                    store8_wire<WIRE>(wire, i, x);
                    if (a.zero_stage) {
                        const float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                        store8_f32(a.stage, i, z);
                    }
                }
            }
        }
    }
    if (a.use_flags) flag_barrier(a, 0, epoch);

    long long lo = 0, hi = a.n;
    if (a.mode != 1 && nranks > 1) {
        long long per = ((a.n / 8 + nranks - 1) / nranks) * 8;
        lo = per * a.rank;
        hi = lo + per < a.n ? lo + per : a.n;
        if (lo > hi) lo = hi;
    }
    // NVLS: the in-switch reduction of the NEXT element block is requested before this one is consumed (the round trip
    // through the switch is longer than a peer load; with one block in flight per thread the links idle)
    float gn[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const long long i_first = lo + ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (NVLS && i_first < hi) mc_load_reduce8<WIRE>(a.grads_mc, i_first, gn);
    for (long long i = i_first; i < hi; i += stride) {
        float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if constexpr (NVLS) {
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = gn[k];
            if (i + stride < hi) mc_load_reduce8<WIRE>(a.grads_mc, i + stride, gn);
        } else {
            // the loads of (up to) four peers are issued before the first add: 4 x 16-byte requests in flight per
            // thread, and the staging registers stay at 32 whatever the site count (the fixed summation order
            // 0,1,2,.. is the same on every site)
            constexpr int kChunk = 4;
            constexpr int kTop = NR > 0 ? NR : kMaxRanks;
#pragma unroll
            for (int r0 = 0; r0 < kTop; r0 += kChunk) {
                float x[kChunk][8];
#pragma unroll
                for (int r = 0; r < kChunk; ++r)
                    if (r0 + r < kTop && (NR > 0 || r0 + r < nranks)) load8<WIRE>(a.grads[r0 + r], i, x[r]);
#pragma unroll
                for (int r = 0; r < kChunk; ++r)
                    if (r0 + r < kTop && (NR > 0 || r0 + r < nranks)) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) g[k] += x[r][k];
                    }
            }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] *= a.grad_scale;

        if (a.opt == 2) {
            store8_f32(a.out, i, g);
            continue;
        }
        float p[8], m[8], v[8];
        load8<0>(a.params[a.rank], i, p);
        load8<0>(a.m, i, m);
        if (a.opt == 0) {
            load8<0>(a.v, i, v);
            // bc2 = sqrt(1 - beta2^t), as torch.optim.Adam.  The bias corrections are folded into two scalars and
            // the per-element divide / square root use the approximate MUFU forms (<= 2 ulp): with IEEE
            // division and sqrt this loop was ~790 instructions per 8 elements and the kernel instruction-bound
            // (15.6k active cycles per SMSP for 30 MB of traffic).
            const float step_size = lr / bc1, inv_bc2 = 1.f / bc2;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float gk = g[k] + a.wd * p[k];
                m[k] = a.beta1 * m[k] + (1.f - a.beta1) * gk;
                v[k] = a.beta2 * v[k] + (1.f - a.beta2) * gk * gk;
                float sq;
                asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(v[k]));
                const float denom = fmaf(sq, inv_bc2, a.eps);
                p[k] -= step_size * __fdividef(m[k], denom);
            }
            store8_f32(a.v, i, v);
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float gk = g[k] + a.wd * p[k];
                m[k] = a.momentum * m[k] + gk;
                p[k] -= lr * (a.momentum != 0.f ? m[k] : gk);
            }
        }
        store8_f32(a.m, i, m);
        uint4 sh;
        sh.x = dnb::pack_bf16x2(p[0], p[1]);
        sh.y = dnb::pack_bf16x2(p[2], p[3]);
        sh.z = dnb::pack_bf16x2(p[4], p[5]);
        sh.w = dnb::pack_bf16x2(p[6], p[7]);
        if (NVLS) {
            // one multicast store reaches every site's copy (this one included)
            mc_store16(a.params_mc + i, __float_as_uint(p[0]), __float_as_uint(p[1]), __float_as_uint(p[2]), __float_as_uint(p[3]));
            mc_store16(a.params_mc + i + 4, __float_as_uint(p[4]), __float_as_uint(p[5]), __float_as_uint(p[6]), __float_as_uint(p[7]));
            if (a.shadow_mc) mc_store16(a.shadow_mc + i, sh.x, sh.y, sh.z, sh.w);
        } else if (a.mode == 0) {
            constexpr int kTop = NR > 0 ? NR : kMaxRanks;
#pragma unroll
            for (int r = 0; r < kTop; ++r)
                if (NR > 0 || r < nranks) {
                    store8_f32(a.params[r], i, p);
                    if (a.shadow[r]) *reinterpret_cast<uint4*>(a.shadow[r] + i) = sh;
                }
        } else {
            store8_f32(a.params[a.rank], i, p);
            if (a.shadow[a.rank]) *reinterpret_cast<uint4*>(a.shadow[a.rank] + i) = sh;
        }
    }
    if (a.use_flags) flag_barrier(a, 1, epoch);
    if (a.dev_state) {
        // the last CTA to finish publishes the new step count; every CTA read the old one at entry
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned prev = atomicAdd(a.dev_state + 1, 1u);
            if (prev == gridDim.x - 1) {
                a.dev_state[1] = 0u;
                a.dev_state[0] = t_dev;
                __threadfence();
            }
        }
    }
}

}  // namespace

DNB_EXPORT int dnb_dsgd_fused_step(const DnbDsgdDesc* d, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_dsgd_fused_step: null descriptor");
    DNB_REQUIRE(d->nranks >= 1 && d->nranks <= kMaxRanks, "nranks must be in [1,%d], got %d", kMaxRanks, d->nranks);
    DNB_REQUIRE(d->rank >= 0 && d->rank < d->nranks, "bad rank %d", d->rank);
    DNB_REQUIRE(d->n % 8 == 0, "flat length must be padded to a multiple of 8 (got %lld)", d->n);
    DNB_REQUIRE(d->wire >= 0 && d->wire <= 2, "wire dtype must be 0 (fp32), 1 (bf16) or 2 (fp16)");
    DNB_REQUIRE(d->opt >= 0 && d->opt <= 2, "opt must be 0 (adam), 1 (sgd) or 2 (reduce only)");
    DsgdArgs a;
    memset(&a, 0, sizeof(a));
    for (int r = 0; r < d->nranks; ++r) {
        a.grads[r] = d->grads[r];
        a.params[r] = d->params[r];
        a.shadow[r] = reinterpret_cast<__nv_bfloat16*>(d->shadow_bf16[r]);
        a.flags[r] = d->flags[r];
        DNB_REQUIRE(a.grads[r] != nullptr, "grads[%d] is null", r);
        DNB_REQUIRE(d->opt == 2 || a.params[r] != nullptr, "params[%d] is null", r);
    }
    a.m = d->m; a.v = d->v; a.out = d->out;
    a.n = d->n; a.nranks = d->nranks; a.rank = d->rank; a.mode = d->mode; a.opt = d->opt;
    a.use_flags = d->use_flags && d->nranks > 1;
    a.lr = d->lr; a.beta1 = d->beta1; a.beta2 = d->beta2; a.eps = d->eps; a.wd = d->weight_decay;
    a.bc1 = 1.f - powf(d->beta1, (float)d->step);
    a.bc2 = sqrtf(1.f - powf(d->beta2, (float)d->step));
    a.grad_scale = d->grad_scale; a.momentum = d->momentum; a.epoch = d->epoch;
    a.dev_state = d->dev_state;
    a.stage = d->stage; a.zero_stage = d->zero_stage;
    a.grads_mc = d->grads_mc; a.params_mc = d->params_mc; a.shadow_mc = reinterpret_cast<__nv_bfloat16*>(d->shadow_mc);
    DNB_REQUIRE(d->mode >= 0 && d->mode <= 2, "mode must be 0 (two-shot P2P), 1 (one-shot P2P) or 2 (two-shot NVLS multicast)");
    if (d->mode == 2) {
        DNB_REQUIRE(d->opt != 2, "NVLS mode updates parameters; use mode 0/1 for a plain reduction");
        DNB_REQUIRE(a.grads_mc && a.params_mc, "NVLS mode needs the multicast addresses of the gradient and parameter buffers");
        DNB_REQUIRE((a.shadow_mc != nullptr) == (a.shadow[d->rank] != nullptr), "NVLS mode: shadow and shadow_mc must be given together");
        DNB_REQUIRE(d->nranks > 1 && d->use_flags, "NVLS mode is for real multi-GPU groups");
    }
    DNB_REQUIRE(a.stage == nullptr || d->wire != 0, "in-kernel gradient packing is for the 16-bit wire formats only");
    if (a.use_flags)
        for (int r = 0; r < d->nranks; ++r) DNB_REQUIRE(a.flags[r] != nullptr, "flags[%d] is null", r);

    long long span = a.n;
    if (a.mode != 1 && a.nranks > 1) span = ((a.n / 8 + a.nranks - 1) / a.nranks) * 8;
    int blocks = dnb_ceil_div(span, (long)kThreads * 8);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (blocks > sms) blocks = sms;           // one CTA per SM: all CTAs co-resident for the flag barrier
    if (blocks > kMaxBlocks) blocks = kMaxBlocks;
    if (blocks < 1) blocks = 1;
    const int nr = (d->nranks == 1 || d->nranks == 2 || d->nranks == 4 || d->nranks == 8) ? d->nranks : 0;
#define DNB_DSGD_LAUNCH(W, N)                                                                                             \
    do {                                                                                                              \
        if (d->mode == 2) DNB_CUDA_CHECK(dnb_launch(dsgd_fused_kernel<W, N, true>, dim3(blocks), dim3(kThreads), 0, stream, 0, a));   \
        else DNB_CUDA_CHECK(dnb_launch(dsgd_fused_kernel<W, N, false>, dim3(blocks), dim3(kThreads), 0, stream, 0, a));               \
    } while (0)
#define DNB_DSGD_WIRE(W)                                  \
    switch (nr) {                                         \
        case 1: DNB_DSGD_LAUNCH(W, 1); break;             \
        case 2: DNB_DSGD_LAUNCH(W, 2); break;             \
        case 4: DNB_DSGD_LAUNCH(W, 4); break;             \
        case 8: DNB_DSGD_LAUNCH(W, 8); break;             \
        default: DNB_DSGD_LAUNCH(W, 0); break;            \
    }
    switch (d->wire) {
        case 0: DNB_DSGD_WIRE(0); break;
        case 1: DNB_DSGD_WIRE(1); break;
        default: DNB_DSGD_WIRE(2); break;
    }
#undef DNB_DSGD_WIRE
#undef DNB_DSGD_LAUNCH
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}
