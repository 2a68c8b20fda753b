// K2a / K2b — device-mode rankDAD (engine semantics §2.3 E3; tunables
// dad_reduction_rank / dad_num_pow_iters, reference compspec.json:236-238; engine name
// comps/__init__.py:15).
//
// A Linear layer's weight gradient is G = Δᵀ·A (Δ = output deltas [N,out], A = input activations
// [N,in]).  G is never formed: a block power (subspace) iteration that touches only Δ and A yields
// rank-r factors,
//
//     T = A·V      W = Δᵀ·T      U = orth(W)      T' = Δ·U      V = Aᵀ·T'        (x dad_num_pow_iters)
//
// and G ≈ U·Vᵀ.  One launch handles every tapped layer of the model: a thread-block cluster per
// layer.  Each of the four products is a tall-skinny GEMM D[M,32] = X[M,K]·Y[32,K]ᵀ on the tensor
// cores: X (Δ, Δᵀ, A or Aᵀ, bf16, both orientations come out of the fused backward) is streamed by TMA
// through a shared-memory ring (it stays L2-resident across the iterations), Y is the current iterate
// split into bf16 hi+lo halves (16+16 rows, so the iterate keeps ~16 mantissa bits), the accumulator
// lives in TMEM (tcgen05.mma M=128, N=32) and is read back with tcgen05.ld.  The M tiles of a product
// are dealt round-robin to the CTAs of the cluster; the iterates travel between CTAs through L2 and a
// cluster barrier per product.  Gram-Schmidt over the r columns runs in registers (fp32), redundantly in
// every CTA, with block-wide shuffle reductions.  All TMA loads read this GPU's memory only.
//
// mode 0 (K2a, "compress"): after the last iteration the kernel's tail stores the factors (U [out,16],
//   V [in,16], fp32) plus the dense gradients (biases, norm layers) straight into slot `rank` of EVERY
//   peer's factor buffer with P2P stores over NVLink and then raises a system-scope flag on every peer:
//   the all-gather of the factors is part of the kernel.
// mode 1 (K2b, "recompress + reconstruct"): waits for all peers' flags, builds the stacked [S·16·ns, ·]
//   operands from the S received factor pairs (ns = 3 stacks hi/hi/lo against hi/lo/hi so that the bf16
//   operands carry the fp32 factors to ~16 bits), runs the same iteration on the stack (every site does
//   this redundantly and deterministically: bit-identical results, no broadcast), and finally runs the
//   reconstruction GEMM G = (1/S)·U·Vᵀ (tcgen05, M=128 x N=128 tiles, K = 48 of hi/lo terms) whose epilogue
//   writes W.grad straight into the site's flat gradient buffer; the dense gradients are averaged in
//   fixed site order next to it.  No NCCL, no cuBLAS, no torch op between backward and the optimiser.
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);
int dnb_make_tmap_bf16_2d(CUtensorMap* map, const void* base, long rows, long cols, long ld, int box_rows,
                          int box_cols);

namespace {

constexpr int R = DNB_DAD_R;              // padded rank (columns of every iterate)
constexpr int RB = 2 * R;                 // hi + lo rows of the skinny operand
constexpr int BM = 128, BK = 64;
constexpr int kThreads = 192;             // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue
constexpr int kWarps = kThreads / 32;
constexpr int kStages = 8;
constexpr int kABytes = BM * BK * 2;      // 16 KB
constexpr int kBBytes = RB * BK * 2;      // 4 KB
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kRowsPT = DNB_DAD_MAX_OUT / kThreads;   // rows of W a thread keeps in registers while orthogonalising
constexpr uint32_t kTmemCols = 128;
constexpr int kRedFloats = 2 * kWarps * R;
constexpr int kSmemBytes = kStages * kStageBytes + (2 * kStages + 8) * 8 + 16 + kRedFloats * 4 + 1024;
static_assert(kRowsPT * kThreads == DNB_DAD_MAX_OUT, "DNB_DAD_MAX_OUT must be a multiple of the block size");

struct LayerK {
    CUtensorMap tA, tAT, tD, tDT;         // big operands (box 128 x 64): A [N,in], Aᵀ [in,N], Δ [N,out], Δᵀ [out,N]
    CUtensorMap tZN, tZI, tZO, tV0;       // skinny hi/lo operands [32, K] (box 32 x 64) over rows / in / out; seeded start
    CUtensorMap tUc, tVc;                 // mode 1: [out,64] / [in,64] hi|lo stacks for the reconstruction (box 128 x 64)
    __nv_bfloat16 *ZN, *ZI, *ZO;
    long long ldzn, ldzi, ldzo;
    float *W, *U, *V;
    __nv_bfloat16 *Uc, *Vc;
    float* G;
    long long ldg;
    __nv_bfloat16 *bA, *bAT, *bD, *bDT;   // mode 1: the stacked operands the kernel builds (the memory behind tA..tDT)
    long long lda, ldat, ldd, lddt;
    long long fact_off;
    int N, in, out, pad_;
};

struct DadArgs {
    LayerK L[DNB_DAD_MAX_LAYERS];
    float* fact[DNB_MAX_RANKS];
    uint32_t* flags[DNB_MAX_RANKS];
    long long slot_floats, dense_off;
    long long span_off[DNB_DAD_MAX_SPANS];
    int span_len[DNB_DAD_MAX_SPANS];
    float* flat_grad;
    uint32_t* dev_state;
    float grad_scale;
    int n_layers, mode, iters, cluster, nsplit, nranks, rank, use_flags, n_spans, rank_cols;
};

struct PhaseIO {
    const CUtensorMap* big;
    const CUtensorMap* skinny;
    int Mtot, Ktot;
    __nv_bfloat16* Z; long long ldz;      // hi/lo transposed iterate out (may be null)
    float* F;                             // fp32 [Mtot,16] out (may be null)
    __nv_bfloat16* cat;                   // [Mtot,64] = hi | lo | hi | 0 (may be null)
};

struct Bars {
    uint64_t *full, *empty, *tfull, *tempty, *rfull, *rfree, *rtfull, *rtempty;
};

__device__ __forceinline__ void split_bf16(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
    hi = __float2bfloat16(x);
    lo = __float2bfloat16(x - __bfloat162float(hi));
}
__device__ __forceinline__ uint32_t pack2(__nv_bfloat16 a, __nv_bfloat16 b) {
    return (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(b) << 16);
}

// One tall-skinny product, executed by the whole CTA (warp-specialised), M tiles dealt round-robin over the cluster.
__device__ __forceinline__ void run_phase(const PhaseIO& io, uint8_t* ring, const Bars& b, uint32_t tmem_base, int warp,
                                          int lane, int crank, int C, uint32_t& ld_cnt, uint32_t& tile_cnt) {
    const int n_tiles = (io.Mtot + BM - 1) / BM;
    const int n_kt = (io.Ktot + BK - 1) / BK;
    const int my_tiles = crank < n_tiles ? (n_tiles - crank + C - 1) / C : 0;
    if (warp == 0) {
        if (dnb::elect_one()) {
            uint32_t lc = ld_cnt;
            for (int t = crank; t < n_tiles; t += C) {
                for (int kt = 0; kt < n_kt; ++kt, ++lc) {
                    const uint32_t s = lc % kStages, ph = (lc / kStages) & 1;
                    dnb::mbar_wait(&b.empty[s], ph ^ 1);
                    uint8_t* dst = ring + s * kStageBytes;
                    dnb::mbar_expect_tx(&b.full[s], kStageBytes);
                    dnb::tma_load_2d(dst, io.big, &b.full[s], kt * BK, t * BM);
                    dnb::tma_load_2d(dst + kABytes, io.skinny, &b.full[s], kt * BK, 0);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (dnb::elect_one()) {
            constexpr uint32_t idesc = dnb::umma_idesc_bf16(BM, RB);
            uint32_t lc = ld_cnt, tc = tile_cnt;
            for (int t = crank; t < n_tiles; t += C, ++tc) {
                const uint32_t buf = tc & 1;
                dnb::mbar_wait(&b.tempty[buf], ((tc >> 1) & 1) ^ 1);
                dnb::tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * RB;
                for (int kt = 0; kt < n_kt; ++kt, ++lc) {
                    const uint32_t s = lc % kStages, ph = (lc / kStages) & 1;
                    dnb::mbar_wait(&b.full[s], ph);
                    dnb::tc_fence_after();
                    const uint32_t a_addr = dnb::smem_u32(ring + s * kStageBytes);
                    const uint64_t da = dnb::umma_desc_k_sw128(a_addr);
                    const uint64_t db = dnb::umma_desc_k_sw128(a_addr + kABytes);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)
                        dnb::umma_bf16_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (kt | k) != 0);
                    dnb::umma_commit(&b.empty[s]);
                }
                dnb::umma_commit(&b.tfull[buf]);
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3;
        uint32_t tc = tile_cnt;
        for (int t = crank; t < n_tiles; t += C, ++tc) {
            const uint32_t buf = tc & 1;
            dnb::mbar_wait(&b.tfull[buf], (tc >> 1) & 1);
            dnb::tc_fence_after();
            uint32_t raw[RB];
            dnb::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * RB, raw);
            dnb::tmem_ld_wait();
            dnb::tc_fence_before();
            __syncwarp();
            if (lane == 0) dnb::mbar_arrive(&b.tempty[buf]);
            const int m = t * BM + q * 32 + lane;
            if (m < io.Mtot) {
                float v[R];
#pragma unroll
                for (int j = 0; j < R; ++j) v[j] = __uint_as_float(raw[j]) + __uint_as_float(raw[R + j]);
                __nv_bfloat16 hi[R], lo[R];
#pragma unroll
                for (int j = 0; j < R; ++j) split_bf16(v[j], hi[j], lo[j]);
                if (io.Z) {
#pragma unroll
                    for (int j = 0; j < R; ++j) {
                        io.Z[(long long)j * io.ldz + m] = hi[j];
                        io.Z[(long long)(R + j) * io.ldz + m] = lo[j];
                    }
                }
                if (io.F) {
                    float4* f = reinterpret_cast<float4*>(io.F + (long long)m * R);
#pragma unroll
                    for (int j = 0; j < R; j += 4) f[j / 4] = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                }
                if (io.cat) {
                    uint4* c = reinterpret_cast<uint4*>(io.cat + (long long)m * 64);
                    uint32_t h[8], l[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        h[j] = pack2(hi[2 * j], hi[2 * j + 1]);
                        l[j] = pack2(lo[2 * j], lo[2 * j + 1]);
                    }
                    c[0] = make_uint4(h[0], h[1], h[2], h[3]); c[1] = make_uint4(h[4], h[5], h[6], h[7]);
                    c[2] = make_uint4(l[0], l[1], l[2], l[3]); c[3] = make_uint4(l[4], l[5], l[6], l[7]);
                    c[4] = c[0]; c[5] = c[1];
                    c[6] = make_uint4(0, 0, 0, 0); c[7] = c[6];
                }
            }
        }
    }
    ld_cnt += (uint32_t)(my_tiles * n_kt);
    tile_cnt += (uint32_t)my_tiles;
}

// every thread ends up with the block-wide sums of its 16 values; fixed summation order
__device__ __forceinline__ void block_reduce16(float (&acc)[R], float* red, int buf) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int c = 0; c < R; ++c) acc[c] = dnb::warp_sum(acc[c]);
    float* mine = red + (buf * kWarps + warp) * R;
    if (lane == 0) {
#pragma unroll
        for (int c = 0; c < R; ++c) mine[c] = acc[c];
    }
    __syncthreads();
    const float* base = red + buf * kWarps * R;
#pragma unroll
    for (int c = 0; c < R; ++c) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) s += base[w * R + c];
        acc[c] = s;
    }
}

// U = orth(W): modified Gram-Schmidt over the columns, the rows spread over the block's registers.  Same
// dead-column rule as distrib/lowrank.orthogonalize and csrc/lowrank.cu.  Every CTA of the cluster (and, in
// mode 1, every site) computes the same bits.
__device__ __forceinline__ void orth_step(const LayerK& L, float* red, bool last, bool writer, bool want_cat,
                                          int rank_cols) {
    const int tid = threadIdx.x;
    const int out = L.out;
    float w[kRowsPT][R];
#pragma unroll
    for (int i = 0; i < kRowsPT; ++i) {
        const int o = tid + i * kThreads;
        if (o < out) {
            const float4* src = reinterpret_cast<const float4*>(L.W + (long long)o * R);
#pragma unroll
            for (int j = 0; j < R; j += 4) {
                const float4 x = __ldcg(src + j / 4);
                w[i][j] = x.x; w[i][j + 1] = x.y; w[i][j + 2] = x.z; w[i][j + 3] = x.w;
            }
        } else {
#pragma unroll
            for (int j = 0; j < R; ++j) w[i][j] = 0.f;
        }
    }
    float acc[R];
#pragma unroll
    for (int c = 0; c < R; ++c) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < kRowsPT; ++i) s += w[i][c] * w[i][c];
        acc[c] = s;
    }
    block_reduce16(acc, red, 0);
    float orig[R], tot = 0.f;
#pragma unroll
    for (int c = 0; c < R; ++c) {
        tot += acc[c];
        orig[c] = sqrtf(acc[c]);
    }
    const float eps = 1e-8f;
    const float scale = sqrtf(tot) + eps;
#pragma unroll
    for (int c = 0; c < R; ++c) {
        if (c < rank_cols) {      // uniform
#pragma unroll
            for (int j = 0; j < R; ++j) {
                float s = 0.f;
                if (j >= c) {
#pragma unroll
                    for (int i = 0; i < kRowsPT; ++i) s += w[i][c] * w[i][j];
                }
                acc[j] = s;
            }
            block_reduce16(acc, red, (c + 1) & 1);
            const float nrm = sqrtf(acc[c]);
            const bool dead = nrm <= fmaxf(eps * scale, 1e-4f * orig[c]);
            const float inv = dead ? 0.f : 1.f / nrm;
#pragma unroll
            for (int i = 0; i < kRowsPT; ++i) {
                const float wc = w[i][c] * inv;
                w[i][c] = wc;
#pragma unroll
                for (int j = c + 1; j < R; ++j) w[i][j] -= wc * (acc[j] * inv);
            }
        } else {
#pragma unroll
            for (int i = 0; i < kRowsPT; ++i) w[i][c] = 0.f;
        }
    }
#pragma unroll
    for (int i = 0; i < kRowsPT; ++i) {
        const int o = tid + i * kThreads;
        if (o < out) {
            __nv_bfloat16 hi[R], lo[R];
#pragma unroll
            for (int j = 0; j < R; ++j) split_bf16(w[i][j], hi[j], lo[j]);
#pragma unroll
            for (int j = 0; j < R; ++j) {
                L.ZO[(long long)j * L.ldzo + o] = hi[j];
                L.ZO[(long long)(R + j) * L.ldzo + o] = lo[j];
            }
            if (last && writer) {
                float4* f = reinterpret_cast<float4*>(L.U + (long long)o * R);
#pragma unroll
                for (int j = 0; j < R; j += 4) f[j / 4] = make_float4(w[i][j], w[i][j + 1], w[i][j + 2], w[i][j + 3]);
                if (want_cat) {
                    uint4* c = reinterpret_cast<uint4*>(L.Uc + (long long)o * 64);
                    uint32_t h[8], l[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        h[j] = pack2(hi[2 * j], hi[2 * j + 1]);
                        l[j] = pack2(lo[2 * j], lo[2 * j + 1]);
                    }
                    c[0] = make_uint4(h[0], h[1], h[2], h[3]); c[1] = make_uint4(h[4], h[5], h[6], h[7]);
                    c[2] = c[0]; c[3] = c[1];
                    c[4] = make_uint4(l[0], l[1], l[2], l[3]); c[5] = make_uint4(l[4], l[5], l[6], l[7]);
                    c[6] = make_uint4(0, 0, 0, 0); c[7] = c[6];
                }
            }
        }
    }
}

__device__ __forceinline__ void phase_barrier() {
    // iterate / factor stores of this CTA -> visible to the other CTAs' TMA (async proxy) and generic loads
    dnb::fence_proxy_async();
    __syncthreads();
    dnb::cluster_arrive();
    dnb::cluster_wait();
    dnb::fence_proxy_async();
}

// mode 1 prologue: stack the S received factor pairs into the bf16 operands of the recompression.
// Row (blk * S + s) * 16 + j of the stack = part_blk of row j of site s's factor; D side parts = hi,hi,lo
// and A side parts = hi,lo,hi, so that the bf16 product of the stacks equals the fp32 product of the factors
// up to the lo·lo term.
__device__ __forceinline__ void build_stack(const DadArgs& a, const LayerK& L, const float* fact_base, int crank, int C,
                                            bool d_side) {
    const int dim = d_side ? L.out : L.in;
    __nv_bfloat16* rowm = d_side ? L.bD : L.bA;       // [ncat, ld]
    __nv_bfloat16* colm = d_side ? L.bDT : L.bAT;     // [dim, ldt]
    const long long ld = d_side ? L.ldd : L.lda, ldt = d_side ? L.lddt : L.ldat;
    const long long foff = L.fact_off + (d_side ? 0 : (long long)L.out * R);
    const int S = a.nranks, ns = a.nsplit;
    const int gtid = crank * kThreads + threadIdx.x, gn = C * kThreads;
    // (1) feature-major copy: one work item = (feature o, site s): 16 floats in, ns x 32 bytes out
    for (int it = gtid; it < dim * S; it += gn) {
        const int o = it / S, s = it % S;
        const float4* src = reinterpret_cast<const float4*>(fact_base + (long long)s * a.slot_floats + foff + (long long)o * R);
        float x[R];
#pragma unroll
        for (int j = 0; j < R; j += 4) {
            const float4 t = __ldcg(src + j / 4);
            x[j] = t.x; x[j + 1] = t.y; x[j + 2] = t.z; x[j + 3] = t.w;
        }
        uint32_t h[8], l[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            __nv_bfloat16 h0, l0, h1, l1;
            split_bf16(x[2 * j], h0, l0);
            split_bf16(x[2 * j + 1], h1, l1);
            h[j] = pack2(h0, h1);
            l[j] = pack2(l0, l1);
        }
        for (int blk = 0; blk < ns; ++blk) {
            const bool use_lo = d_side ? (blk == 2) : (blk == 1);
            uint4* dst = reinterpret_cast<uint4*>(colm + (long long)o * ldt + (long long)(blk * S + s) * R);
            dst[0] = use_lo ? make_uint4(l[0], l[1], l[2], l[3]) : make_uint4(h[0], h[1], h[2], h[3]);
            dst[1] = use_lo ? make_uint4(l[4], l[5], l[6], l[7]) : make_uint4(h[4], h[5], h[6], h[7]);
        }
    }
    // (2) row-major copy: one work item = (site s, factor row j, 8 consecutive features)
    const int chunks = (int)(ld / 8);
    for (int it = gtid; it < S * R * chunks; it += gn) {
        const int ch = it % chunks, j = (it / chunks) % R, s = it / (chunks * R);
        const float* src = fact_base + (long long)s * a.slot_floats + foff + j;
        uint32_t h[4], l[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int o0 = ch * 8 + 2 * e;
            const float x0 = o0 < dim ? __ldcg(src + (long long)o0 * R) : 0.f;
            const float x1 = o0 + 1 < dim ? __ldcg(src + (long long)(o0 + 1) * R) : 0.f;
            __nv_bfloat16 h0, l0, h1, l1;
            split_bf16(x0, h0, l0);
            split_bf16(x1, h1, l1);
            h[e] = pack2(h0, h1);
            l[e] = pack2(l0, l1);
        }
        for (int blk = 0; blk < ns; ++blk) {
            const bool use_lo = d_side ? (blk == 2) : (blk == 1);
            *reinterpret_cast<uint4*>(rowm + (long long)((blk * S + s) * R + j) * ld + ch * 8) =
                use_lo ? make_uint4(l[0], l[1], l[2], l[3]) : make_uint4(h[0], h[1], h[2], h[3]);
        }
    }
}

__global__ void __launch_bounds__(kThreads, 1) dad_kernel(const __grid_constant__ DadArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (dnb::smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* ring = smem;
    uint64_t* bar0 = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
    Bars b;
    b.full = bar0; b.empty = bar0 + kStages; b.tfull = bar0 + 2 * kStages; b.tempty = b.tfull + 2;
    b.rfull = b.tempty + 2; b.rfree = b.rfull + 1; b.rtfull = b.rfree + 1; b.rtempty = b.rtfull + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar0 + 2 * kStages + 8);
    float* red = reinterpret_cast<float*>(tmem_slot + 4);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int C = a.cluster;
    const int crank = (int)dnb::cluster_ctarank();
    const int layer = blockIdx.x / C;
    const LayerK& L = a.L[layer];

    if (warp == 0 && lane == 0) {
        dnb::tma_prefetch_desc(&L.tA); dnb::tma_prefetch_desc(&L.tAT);
        dnb::tma_prefetch_desc(&L.tD); dnb::tma_prefetch_desc(&L.tDT);
        dnb::tma_prefetch_desc(&L.tZN); dnb::tma_prefetch_desc(&L.tZI);
        dnb::tma_prefetch_desc(&L.tZO); dnb::tma_prefetch_desc(&L.tV0);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < kStages; ++s) {
                dnb::mbar_init(&b.full[s], 1);
                dnb::mbar_init(&b.empty[s], 1);
            }
            dnb::mbar_init(&b.tfull[0], 1); dnb::mbar_init(&b.tfull[1], 1);
            dnb::mbar_init(&b.tempty[0], 4); dnb::mbar_init(&b.tempty[1], 4);
            dnb::mbar_init(b.rfull, 1); dnb::mbar_init(b.rfree, 1);
            dnb::mbar_init(b.rtfull, 1); dnb::mbar_init(b.rtempty, 4);
            dnb::fence_mbar_init();
        }
        __syncwarp();
        dnb::tmem_alloc(tmem_slot, kTmemCols);
    }
    dnb::tc_fence_before();
    __syncthreads();
    dnb::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    dnb::pdl_wait();

    // exchange generation: mode 0 opens generation e = done + 1, mode 1 consumes the generation mode 0 closed
    uint32_t epoch = 1;
    if (a.dev_state) epoch = *reinterpret_cast<volatile uint32_t*>(a.dev_state) + (a.mode == 0 ? 1u : 0u);
    const long long half = (long long)a.nranks * a.slot_floats;       // one generation of the double-buffered slots
    const long long gen_off = (long long)(epoch & 1u) * half;

    if (a.mode == 1) {
        if (a.use_flags) {
            if (threadIdx.x < a.nranks) {
                const uint32_t* f = a.flags[a.rank] + threadIdx.x;
                const long long t0 = clock64();
                while ((int32_t)(dnb::ld_acquire_sys(f) - epoch) < 0) {
                    if (clock64() - t0 > dnb::kSpinBudgetCycles) {
                        printf("dnb: rankDAD factor flag timed out (rank %d waits for %d, generation %u)\n", a.rank,
                               (int)threadIdx.x, epoch);
                        __trap();
                    }
                }
            }
            __syncthreads();
        }
        const float* fb = a.fact[a.rank] + gen_off;
        build_stack(a, L, fb, crank, C, true);
        build_stack(a, L, fb, crank, C, false);
        if (layer == 0) {
            // dense gradients (biases, norm layers): mean over the sites in fixed order, into the flat gradient buffer
            long long pos = a.dense_off;
            for (int sp = 0; sp < a.n_spans; ++sp) {
                const int len = a.span_len[sp];
                for (int i = crank * kThreads + threadIdx.x; i < len; i += C * kThreads) {
                    float s = 0.f;
                    for (int r = 0; r < a.nranks; ++r) s += __ldcg(fb + (long long)r * a.slot_floats + pos + i);
                    a.flat_grad[a.span_off[sp] + i] = s * a.grad_scale;
                }
                pos += len;
            }
        }
    }
    phase_barrier();

    uint32_t ld_cnt = 0, tile_cnt = 0;
    for (int it = 0; it < a.iters; ++it) {
        const bool last = it == a.iters - 1;
        PhaseIO io;
        // T = A·V
        io.big = &L.tA; io.skinny = it == 0 ? &L.tV0 : &L.tZI; io.Mtot = L.N; io.Ktot = L.in;
        io.Z = L.ZN; io.ldz = L.ldzn; io.F = nullptr; io.cat = nullptr;
        run_phase(io, ring, b, tmem_base, warp, lane, crank, C, ld_cnt, tile_cnt);
        phase_barrier();
        // W = Δᵀ·T
        io.big = &L.tDT; io.skinny = &L.tZN; io.Mtot = L.out; io.Ktot = L.N;
        io.Z = nullptr; io.F = L.W;
        run_phase(io, ring, b, tmem_base, warp, lane, crank, C, ld_cnt, tile_cnt);
        phase_barrier();
        // U = orth(W)   (redundantly in every CTA: no extra barrier)
        orth_step(L, red, last, crank == 0, a.mode == 1, a.rank_cols);
        dnb::fence_proxy_async();
        __syncthreads();
        // T' = Δ·U
        io.big = &L.tD; io.skinny = &L.tZO; io.Mtot = L.N; io.Ktot = L.out;
        io.Z = L.ZN; io.ldz = L.ldzn; io.F = nullptr;
        run_phase(io, ring, b, tmem_base, warp, lane, crank, C, ld_cnt, tile_cnt);
        phase_barrier();
        // V = Aᵀ·T'
        io.big = &L.tAT; io.skinny = &L.tZN; io.Mtot = L.in; io.Ktot = L.N;
        io.Z = L.ZI; io.ldz = L.ldzi; io.F = last ? L.V : nullptr; io.cat = (last && a.mode == 1) ? L.Vc : nullptr;
        run_phase(io, ring, b, tmem_base, warp, lane, crank, C, ld_cnt, tile_cnt);
        phase_barrier();
    }

    if (a.mode == 0) {
        // ---- fused all-gather: this layer's factors (and, from layer 0's cluster, the dense gradients) go
        // into slot `rank` of every site's factor buffer, over NVLink for the remote ones
        const int gtid = crank * kThreads + threadIdx.x, gn = C * kThreads;
        const long long slot = gen_off + (long long)a.rank * a.slot_floats;
        const int nu4 = L.out * (R / 4), n4 = (L.out + L.in) * (R / 4);
        const float4* U4 = reinterpret_cast<const float4*>(L.U);
        const float4* V4 = reinterpret_cast<const float4*>(L.V);
        for (int i = gtid; i < n4; i += gn) {
            const float4 x = i < nu4 ? __ldcg(U4 + i) : __ldcg(V4 + (i - nu4));
            for (int p = 0; p < a.nranks; ++p)
                reinterpret_cast<float4*>(a.fact[p] + slot + L.fact_off)[i] = x;
        }
        if (layer == 0) {
            long long pos = a.dense_off;
            for (int sp = 0; sp < a.n_spans; ++sp) {
                const int len = a.span_len[sp];
                for (int i = gtid; i < len; i += gn) {
                    const float x = a.flat_grad[a.span_off[sp] + i];
                    for (int p = 0; p < a.nranks; ++p) a.fact[p][slot + pos + i] = x;
                }
                pos += len;
            }
        }
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0 && a.dev_state) {
            const unsigned prev = atomicAdd(a.dev_state + 1, 1u);
            if (prev == gridDim.x - 1) {
                // last CTA of the grid: every CTA's stores are ordered before this point
                a.dev_state[1] = 0u;
                a.dev_state[0] = epoch;
                __threadfence_system();
                if (a.use_flags)
                    for (int p = 0; p < a.nranks; ++p) dnb::st_release_sys(a.flags[p] + a.rank, epoch);
            }
        }
    } else {
        // ---- reconstruction: G[out,in] = grad_scale * U·Vᵀ from the hi/lo stacks (K = 48 of 64), 128 x 128 tiles
        const int n_mo = (L.out + BM - 1) / BM, n_ni = (L.in + BM - 1) / BM;
        const int tiles = n_mo * n_ni;
        uint32_t li = 0;
        for (int tix = crank; tix < tiles; tix += C, ++li) {
            const int mo = tix / n_ni, ni = tix % n_ni;
            if (warp == 0) {
                if (dnb::elect_one()) {
                    dnb::mbar_wait(b.rfree, (li & 1) ^ 1);
                    dnb::mbar_expect_tx(b.rfull, 2 * kABytes);
                    dnb::tma_load_2d(ring, &L.tUc, b.rfull, 0, mo * BM);
                    dnb::tma_load_2d(ring + kABytes, &L.tVc, b.rfull, 0, ni * BM);
                }
                __syncwarp();
            } else if (warp == 1) {
                if (dnb::elect_one()) {
                    constexpr uint32_t idesc = dnb::umma_idesc_bf16(BM, BM);
                    dnb::mbar_wait(b.rfull, li & 1);
                    dnb::mbar_wait(b.rtempty, (li & 1) ^ 1);
                    dnb::tc_fence_after();
                    const uint32_t a_addr = dnb::smem_u32(ring);
                    const uint64_t da = dnb::umma_desc_k_sw128(a_addr);
                    const uint64_t db = dnb::umma_desc_k_sw128(a_addr + kABytes);
#pragma unroll
                    for (int k = 0; k < 3; ++k) dnb::umma_bf16_ss(tmem_base, da + 2 * k, db + 2 * k, idesc, k != 0);
                    dnb::umma_commit(b.rfree);
                    dnb::umma_commit(b.rtfull);
                }
                __syncwarp();
            } else {
                const int q = warp & 3;
                dnb::mbar_wait(b.rtfull, li & 1);
                dnb::tc_fence_after();
                const int o = mo * BM + q * 32 + lane;
#pragma unroll 1
                for (int c = 0; c < BM / 32; ++c) {
                    uint32_t raw[32];
                    dnb::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * 32), raw);
                    dnb::tmem_ld_wait();
                    const int i0 = ni * BM + c * 32;
                    if (o < L.out && i0 < L.in) {
                        float* g = L.G + (long long)o * L.ldg + i0;
                        if (i0 + 32 <= L.in && (reinterpret_cast<uintptr_t>(g) & 15) == 0) {
#pragma unroll
                            for (int j = 0; j < 32; j += 4)
                                *reinterpret_cast<float4*>(g + j) =
                                    make_float4(__uint_as_float(raw[j]) * a.grad_scale, __uint_as_float(raw[j + 1]) * a.grad_scale,
                                                __uint_as_float(raw[j + 2]) * a.grad_scale, __uint_as_float(raw[j + 3]) * a.grad_scale);
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (i0 + j < L.in) g[j] = __uint_as_float(raw[j]) * a.grad_scale;
                        }
                    }
                }
                dnb::tc_fence_before();
                __syncwarp();
                if (lane == 0) dnb::mbar_arrive(b.rtempty);
            }
        }
    }
    if (threadIdx.x == 0) dnb::pdl_launch();

    dnb::tc_fence_before();
    __syncthreads();
    // no CTA leaves while the cluster's barrier protocol is still in use by a slower CTA
    dnb::cluster_arrive();
    dnb::cluster_wait();
    if (warp == 1) {
        dnb::tc_fence_after();
        dnb::tmem_dealloc(tmem_base, kTmemCols);
    }
}

// fp32 [N, d] -> bf16 [N, ldr] and bf16 [d, ldt] (both orientations of a (delta, activation) matrix whose producer
// only has fp32, e.g. the classifier-head pairs: a few thousand elements each)
struct PackArgs {
    const float* src[DNB_DAD_MAX_PACK];
    __nv_bfloat16* rows[DNB_DAD_MAX_PACK];
    __nv_bfloat16* colsT[DNB_DAD_MAX_PACK];
    long long lds[DNB_DAD_MAX_PACK], ldr[DNB_DAD_MAX_PACK], ldt[DNB_DAD_MAX_PACK];
    int N[DNB_DAD_MAX_PACK], d[DNB_DAD_MAX_PACK];
};

__global__ void __launch_bounds__(256) dad_pack_kernel(const PackArgs p) {
    dnb::pdl_wait();
    const int k = blockIdx.y;
    const float* src = nullptr;
    __nv_bfloat16 *rows = nullptr, *colsT = nullptr;
    long long lds = 0, ldr = 0, ldt = 0;
    int N = 0, d = 0;
#pragma unroll
    for (int i = 0; i < DNB_DAD_MAX_PACK; ++i)      // static indexing: kernel arguments stay in the constant bank
        if (i == k) {
            src = p.src[i]; rows = p.rows[i]; colsT = p.colsT[i];
            lds = p.lds[i]; ldr = p.ldr[i]; ldt = p.ldt[i]; N = p.N[i]; d = p.d[i];
        }
    const long long total = (long long)N * d;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(i / d), c = (int)(i % d);
        const __nv_bfloat16 v = __float2bfloat16(src[(long long)n * lds + c]);
        if (rows) rows[(long long)n * ldr + c] = v;
        if (colsT) colsT[(long long)c * ldt + n] = v;
    }
}

struct MapCache {
    DnbDadDesc key;
    DadArgs args;
    bool valid = false;
};
thread_local MapCache g_cache[2];

}  // namespace

DNB_EXPORT int dnb_dad_pack(const DnbDadPackDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d && d->count >= 1 && d->count <= DNB_DAD_MAX_PACK, "dnb_dad_pack: count must be in [1,%d]", DNB_DAD_MAX_PACK);
    PackArgs p;
    memset(&p, 0, sizeof(p));
    long long mx = 0;
    for (int k = 0; k < d->count; ++k) {
        DNB_REQUIRE(d->src[k] && (d->rows[k] || d->colsT[k]) && d->N[k] > 0 && d->d[k] > 0, "dnb_dad_pack: bad item %d", k);
        p.src[k] = d->src[k];
        p.rows[k] = reinterpret_cast<__nv_bfloat16*>(d->rows[k]);
        p.colsT[k] = reinterpret_cast<__nv_bfloat16*>(d->colsT[k]);
        p.lds[k] = d->lds[k]; p.ldr[k] = d->ldr[k]; p.ldt[k] = d->ldt[k];
        p.N[k] = d->N[k]; p.d[k] = d->d[k];
        const long long t = (long long)d->N[k] * d->d[k];
        if (t > mx) mx = t;
    }
    int bx = (int)((mx + 255) / 256);
    if (bx > 64) bx = 64;
    DNB_CUDA_CHECK(dnb_launch(dad_pack_kernel, dim3(bx, d->count), dim3(256), 0, st, 0, p));
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}

DNB_EXPORT int dnb_dad_step(const DnbDadDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_dad_step: null descriptor");
    DNB_REQUIRE(d->n_layers >= 1 && d->n_layers <= DNB_DAD_MAX_LAYERS, "n_layers must be in [1,%d], got %d",
                DNB_DAD_MAX_LAYERS, d->n_layers);
    DNB_REQUIRE(d->mode == 0 || d->mode == 1, "mode must be 0 (compress) or 1 (recompress+reconstruct)");
    DNB_REQUIRE(d->iters >= 1, "iters must be >= 1");
    DNB_REQUIRE(d->cluster >= 1 && d->cluster <= 8, "cluster size must be in [1,8], got %d", d->cluster);
    DNB_REQUIRE(d->nranks >= 1 && d->nranks <= DNB_MAX_RANKS && d->rank >= 0 && d->rank < d->nranks, "bad rank %d / %d",
                d->rank, d->nranks);
    DNB_REQUIRE(d->nsplit == 1 || d->nsplit == 3, "nsplit must be 1 or 3");
    DNB_REQUIRE(d->rank_cols >= 1 && d->rank_cols <= DNB_DAD_R, "rank must be in [1,%d], got %d", DNB_DAD_R, d->rank_cols);
    DNB_REQUIRE(d->n_spans >= 0 && d->n_spans <= DNB_DAD_MAX_SPANS, "at most %d dense spans", DNB_DAD_MAX_SPANS);
    DNB_REQUIRE(d->slot_floats % 4 == 0 && d->dense_off % 4 == 0, "slot layout must be 16-byte aligned");
    for (int r = 0; r < d->nranks; ++r) DNB_REQUIRE(d->fact[r] != nullptr, "fact[%d] is null", r);
    const bool flags = d->use_flags && d->nranks > 1;
    if (flags) {
        DNB_REQUIRE(d->dev_state != nullptr, "cross-GPU exchange needs dev_state");
        for (int r = 0; r < d->nranks; ++r) DNB_REQUIRE(d->flags[r] != nullptr, "flags[%d] is null", r);
    }
    MapCache& mc = g_cache[d->mode];
    if (!(mc.valid && memcmp(&mc.key, d, sizeof(*d)) == 0)) {
        mc.valid = false;
        DadArgs& a = mc.args;
        memset(&a, 0, sizeof(a));
        for (int i = 0; i < d->n_layers; ++i) {
            const DnbDadLayerDesc& s = d->layer[i];
            LayerK& L = a.L[i];
            DNB_REQUIRE(s.N > 0 && s.in > 0 && s.out > 0, "layer %d: empty shape", i);
            DNB_REQUIRE(s.out <= DNB_DAD_MAX_OUT, "layer %d: out = %d exceeds the in-register orthogonalisation limit %d", i,
                        s.out, DNB_DAD_MAX_OUT);
            DNB_REQUIRE(s.A && s.AT && s.D && s.DT && s.ZN && s.ZI && s.ZO && s.V0 && s.W && s.U && s.V,
                        "layer %d: null operand / workspace pointer", i);
            DNB_REQUIRE(s.fact_off % 4 == 0, "layer %d: factor offset must be a multiple of 4 floats", i);
            DNB_REQUIRE(s.fact_off + (long long)(s.in + s.out) * DNB_DAD_R <= d->slot_floats, "layer %d: factors exceed the slot", i);
            int rc = 0;
            rc |= dnb_make_tmap_bf16_2d(&L.tA, s.A, s.N, s.in, s.lda, BM, BK);
            rc |= dnb_make_tmap_bf16_2d(&L.tAT, s.AT, s.in, s.N, s.ldat, BM, BK);
            rc |= dnb_make_tmap_bf16_2d(&L.tD, s.D, s.N, s.out, s.ldd, BM, BK);
            rc |= dnb_make_tmap_bf16_2d(&L.tDT, s.DT, s.out, s.N, s.lddt, BM, BK);
            rc |= dnb_make_tmap_bf16_2d(&L.tZN, s.ZN, RB, s.N, s.ldzn, RB, BK);
            rc |= dnb_make_tmap_bf16_2d(&L.tZI, s.ZI, RB, s.in, s.ldzi, RB, BK);
            rc |= dnb_make_tmap_bf16_2d(&L.tZO, s.ZO, RB, s.out, s.ldzo, RB, BK);
            rc |= dnb_make_tmap_bf16_2d(&L.tV0, s.V0, RB, s.in, s.ldzi, RB, BK);
            if (d->mode == 1) {
                DNB_REQUIRE(s.Uc && s.Vc && s.G, "layer %d: reconstruction needs Uc, Vc and G", i);
                DNB_REQUIRE(s.N == d->nsplit * d->nranks * DNB_DAD_R, "layer %d: stacked rows must be nsplit*nranks*%d", i,
                            DNB_DAD_R);
                rc |= dnb_make_tmap_bf16_2d(&L.tUc, s.Uc, s.out, 64, 64, BM, BK);
                rc |= dnb_make_tmap_bf16_2d(&L.tVc, s.Vc, s.in, 64, 64, BM, BK);
            }
            if (rc) return rc;
            L.ZN = (__nv_bfloat16*)s.ZN; L.ZI = (__nv_bfloat16*)s.ZI; L.ZO = (__nv_bfloat16*)s.ZO;
            L.ldzn = s.ldzn; L.ldzi = s.ldzi; L.ldzo = s.ldzo;
            L.W = s.W; L.U = s.U; L.V = s.V;
            L.Uc = (__nv_bfloat16*)s.Uc; L.Vc = (__nv_bfloat16*)s.Vc;
            L.G = s.G; L.ldg = s.ldg;
            L.bA = (__nv_bfloat16*)s.A; L.bAT = (__nv_bfloat16*)s.AT; L.bD = (__nv_bfloat16*)s.D; L.bDT = (__nv_bfloat16*)s.DT;
            L.lda = s.lda; L.ldat = s.ldat; L.ldd = s.ldd; L.lddt = s.lddt;
            L.fact_off = s.fact_off;
            L.N = s.N; L.in = s.in; L.out = s.out;
        }
        long long dense_total = 0;
        for (int k = 0; k < d->n_spans; ++k) {
            a.span_off[k] = d->span_off[k];
            a.span_len[k] = d->span_len[k];
            dense_total += d->span_len[k];
        }
        DNB_REQUIRE(d->n_spans == 0 || (d->flat_grad && d->dense_off + dense_total <= d->slot_floats),
                    "dense gradients do not fit the slot");
        for (int r = 0; r < d->nranks; ++r) {
            a.fact[r] = d->fact[r];
            a.flags[r] = d->flags[r];
        }
        a.slot_floats = d->slot_floats; a.dense_off = d->dense_off;
        a.flat_grad = d->flat_grad; a.dev_state = d->dev_state; a.grad_scale = d->grad_scale;
        a.n_layers = d->n_layers; a.mode = d->mode; a.iters = d->iters; a.cluster = d->cluster; a.nsplit = d->nsplit;
        a.nranks = d->nranks; a.rank = d->rank; a.use_flags = flags ? 1 : 0; a.n_spans = d->n_spans;
        a.rank_cols = d->rank_cols;
        mc.key = *d;
        mc.valid = true;
    }
    static bool configured = false;
    if (!configured) {
        DNB_CUDA_CHECK(cudaFuncSetAttribute(dad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        configured = true;
    }
    DNB_CUDA_CHECK(dnb_launch(dad_kernel, dim3(d->n_layers * d->cluster), dim3(kThreads), kSmemBytes, st, d->cluster, mc.args));
    DNB_LAUNCH_CHECK();
    dnb_count_launch(1);
    return 0;
}
