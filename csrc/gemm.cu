// tcgen05 + TMA GEMM with fused epilogues (K4 wide layers, K5a LSTM input
// projection, K7 encoder, the weight/data-gradient GEMMs of their backward passes, and the
// K3 powerSGD products).
//
//     D[M,N] = alpha * A[M,K] · B[N,K]ᵀ   (A, B bf16 row-major = "K-major"; fp32 accumulate)
//
// Structure (one 128 x BN output tile per CTA, optional split-K over grid.z):
//   warp 0      TMA producer: cp.async.bulk.tensor 2-D boxes (128x64 of A, BNx64 of B, 128-byte
//               swizzle) into a STAGES-deep shared-memory ring, completion on "full" mbarriers;
//   warp 1      MMA issuer: one elected thread issues tcgen05.mma.kind::f16 (M=128, N=BN, K=16)
//               four per k-tile, accumulating in TMEM; tcgen05.commit releases the ring slot
//               ("empty" mbarrier) and finally signals the epilogue;
//   warps 2-5   epilogue: tcgen05.ld the accumulator (one row per thread, 32 columns per load),
//               apply row/column bias, ReLU, ReLU-mask multiply, optional column sums (bias
//               gradient), then store fp32 or bf16, and optionally a transposed bf16 copy so the
//               next GEMM can consume it K-major without a separate transpose pass.
// Ragged M/N/K are handled by TMA out-of-bounds zero fill plus guarded stores.
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);
int dnb_make_tmap_bf16_2d(CUtensorMap* map, const void* base, long rows, long cols, long ld, int box_rows,
                          int box_cols);

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kGemmThreads = 192;

struct Epi {
    void* out; long long ldo; int out_bf16;
    __nv_bfloat16* outT; long long ldt;
    const float* bias_n; const float* bias_m;
    const __nv_bfloat16* mask; long long ldmask;
    float* colsum;
    float alpha; int relu; int atomic_acc;
    // BatchNorm-over-columns mode (rows = features, columns = the whole batch, N <= BN):
    const float* bn_gamma; const float* bn_beta; float bn_eps;
    float* bn_xhat; long long ldxh;      // [M, N] normalised values (saved for backward)
    float* bn_invstd;                    // [M]
};

template <int BN>
struct SmemLayout {
    static constexpr int kStages = (BN <= 64) ? 6 : (BN <= 128 ? 4 : 3);
    static constexpr int kABytes = BM * BK * 2;
    static constexpr int kBBytes = BN * BK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kBarOffset = kStages * kStageBytes;
    static constexpr int kTotal = kBarOffset + (2 * kStages + 1) * 8 + 16 + 1024;  // + alignment slack
};

template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Epi epi,
                 int M, int N, int K, int k_tiles_per_split, int STAGES) {
    using L = SmemLayout<BN>;
    constexpr uint32_t kTmemCols = BN < 32 ? 32 : BN;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (dnb::smem_u32(smem_raw) & 1023u)) & 1023u);
    // STAGES (<= L::kStages) is chosen by the host: a short K loop does not need the full ring, and a
    // smaller footprint lets two CTAs share an SM when the tile count slightly exceeds one wave
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * L::kStageBytes);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform for the compiler
    const int lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN;
    const int m0 = blockIdx.y * BM;
    const int total_k_tiles = (K + BK - 1) / BK;
    const int kt_begin = blockIdx.z * k_tiles_per_split;
    int kt_end = kt_begin + k_tiles_per_split;
    if (kt_end > total_k_tiles) kt_end = total_k_tiles;
    const int num_kt = kt_end - kt_begin;   // >= 1 by construction of the grid

    if (warp == 0 && lane == 0) {
        dnb::tma_prefetch_desc(&tmA);
        dnb::tma_prefetch_desc(&tmB);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < STAGES; ++s) {
                dnb::mbar_init(&full_bar[s], 1);
                dnb::mbar_init(&empty_bar[s], 1);
            }
            dnb::mbar_init(tmem_full, 1);
            dnb::fence_mbar_init();
        }
        __syncwarp();
        dnb::tmem_alloc(tmem_slot, kTmemCols);
    }
    dnb::tc_fence_before();
    __syncthreads();
    dnb::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    dnb::pdl_wait();      // prologue done; operands of the preceding kernel are complete and visible from here on

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        // (elect.sync under warp-uniform control flow keeps descriptors in uniform registers)
        if (dnb::elect_one()) {
            for (int i = 0; i < num_kt; ++i) {
                const int s = i % STAGES;
                const uint32_t ph = (i / STAGES) & 1;
                dnb::mbar_wait(&empty_bar[s], ph ^ 1);
                uint8_t* a_dst = smem + s * L::kStageBytes;
                uint8_t* b_dst = a_dst + L::kABytes;
                dnb::mbar_expect_tx(&full_bar[s], L::kStageBytes);
                const int kc = (kt_begin + i) * BK;
                dnb::tma_load_2d(a_dst, &tmA, &full_bar[s], kc, m0);
                dnb::tma_load_2d(b_dst, &tmB, &full_bar[s], kc, n0);
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (dnb::elect_one()) {
            constexpr uint32_t idesc = dnb::umma_idesc_bf16(BM, BN);
            for (int i = 0; i < num_kt; ++i) {
                const int s = i % STAGES;
                const uint32_t ph = (i / STAGES) & 1;
                dnb::mbar_wait(&full_bar[s], ph);
                dnb::tc_fence_after();
                const uint32_t a_addr = dnb::smem_u32(smem + s * L::kStageBytes);
                const uint64_t da = dnb::umma_desc_k_sw128(a_addr);
                const uint64_t db = dnb::umma_desc_k_sw128(a_addr + L::kABytes);
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                    // +32 bytes per K=16 step inside the 128-byte swizzle atom (address field is >>4)
                    dnb::umma_bf16_ss(tmem_base, da + 2 * k, db + 2 * k, idesc, (i | k) != 0);
                }
                dnb::umma_commit(&empty_bar[s]);
            }
            dnb::umma_commit(tmem_full);
        }
    } else {
        // ------------------------------------------------------------------ epilogue (4 warps)
        const int q = warp & 3;                       // TMEM lane quadrant this warp may read
        const int m = m0 + q * 32 + lane;
        dnb::mbar_wait(tmem_full, 0);
        dnb::tc_fence_after();
        if (warp == 2 && lane == 0) dnb::pdl_launch();     // only the epilogue is left: let the next kernel start up
        constexpr int CH = BN < 32 ? BN : 32;
        // ---- optional BatchNorm over the columns of each row (batch statistics, biased variance):
        // every thread owns one row with all N <= BN batch columns in TMEM, so mean/variance need no
        // cross-thread reduction; TMEM is simply read three times (sum, centred squares, normalise).
        const bool bn = epi.bn_gamma != nullptr;
        float bn_mean = 0.f, bn_istd = 1.f, bn_g = 1.f, bn_b = 0.f;
        const float row_bias = (epi.bias_m && m < M) ? epi.bias_m[m] : 0.f;
        if (bn) {
            float s1 = 0.f;
#pragma unroll 1
            for (int c = 0; c < BN / CH; ++c) {
                uint32_t raw[CH];
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * CH);
                if constexpr (CH == 32) dnb::tmem_ld_32x32(taddr, raw);
                else dnb::tmem_ld_32x16(taddr, raw);
                dnb::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < CH; ++j)
                    if (n0 + c * CH + j < N) s1 += __uint_as_float(raw[j]) * epi.alpha + row_bias;
            }
            bn_mean = s1 / (float)N;
            float s2 = 0.f;
#pragma unroll 1
            for (int c = 0; c < BN / CH; ++c) {
                uint32_t raw[CH];
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * CH);
                if constexpr (CH == 32) dnb::tmem_ld_32x32(taddr, raw);
                else dnb::tmem_ld_32x16(taddr, raw);
                dnb::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < CH; ++j)
                    if (n0 + c * CH + j < N) {
                        const float dlt = __uint_as_float(raw[j]) * epi.alpha + row_bias - bn_mean;
                        s2 += dlt * dlt;
                    }
            }
            bn_istd = rsqrtf(s2 / (float)N + epi.bn_eps);
            if (m < M) {
                bn_g = epi.bn_gamma[m];
                bn_b = epi.bn_beta ? epi.bn_beta[m] : 0.f;
                if (epi.bn_invstd) epi.bn_invstd[m] = bn_istd;
            }
        }
#pragma unroll 1
        for (int c = 0; c < BN / CH; ++c) {
            uint32_t raw[CH];
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * CH);
            if constexpr (CH == 32) dnb::tmem_ld_32x32(taddr, raw);
            else dnb::tmem_ld_32x16(taddr, raw);
            dnb::tmem_ld_wait();
            const int nb = n0 + c * CH;
            if (nb >= N) continue;   // warp-uniform
            float v[CH];
            const float bm = row_bias;
#pragma unroll
            for (int j = 0; j < CH; ++j) {
                float x = __uint_as_float(raw[j]) * epi.alpha + bm;
                const int n = nb + j;
                if (epi.bias_n && n < N) x += epi.bias_n[n];
                if (bn) {
                    const float xh = (x - bn_mean) * bn_istd;
                    if (epi.bn_xhat && m < M && n < N) epi.bn_xhat[(long long)m * epi.ldxh + n] = xh;
                    x = xh * bn_g + bn_b;
                }
                if (epi.relu) x = fmaxf(x, 0.f);
                v[j] = x;
            }
            if (epi.mask && m < M) {
                const __nv_bfloat16* mrow = epi.mask + (long long)m * epi.ldmask + nb;
#pragma unroll
                for (int j = 0; j < CH; ++j)
                    if (nb + j < N && !(__bfloat162float(mrow[j]) > 0.f)) v[j] = 0.f;
            }
            if (epi.colsum) {
#pragma unroll
                for (int j = 0; j < CH; ++j) {
                    float s = dnb::warp_sum(m < M ? v[j] : 0.f);
                    if (lane == 0 && nb + j < N) atomicAdd(epi.colsum + nb + j, s);
                }
            }
            if (m < M) {
                const bool full = (nb + CH <= N);
                if (epi.out) {
                    if (epi.out_bf16) {
                        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(epi.out) + (long long)m * epi.ldo + nb;
                        if (full && (epi.ldo % 8 == 0) && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
                            for (int j = 0; j < CH; j += 8) {
                                uint4 pk;
                                pk.x = dnb::pack_bf16x2(v[j], v[j + 1]);
                                pk.y = dnb::pack_bf16x2(v[j + 2], v[j + 3]);
                                pk.z = dnb::pack_bf16x2(v[j + 4], v[j + 5]);
                                pk.w = dnb::pack_bf16x2(v[j + 6], v[j + 7]);
                                *reinterpret_cast<uint4*>(o + j) = pk;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < CH; ++j)
                                if (nb + j < N) o[j] = __float2bfloat16(v[j]);
                        }
                    } else {
                        float* o = reinterpret_cast<float*>(epi.out) + (long long)m * epi.ldo + nb;
                        if (epi.atomic_acc) {
#pragma unroll
                            for (int j = 0; j < CH; ++j)
                                if (nb + j < N) atomicAdd(o + j, v[j]);
                        } else if (full && (epi.ldo % 4 == 0) && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
                            for (int j = 0; j < CH; j += 4)
                                *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                        } else {
#pragma unroll
                            for (int j = 0; j < CH; ++j)
                                if (nb + j < N) o[j] = v[j];
                        }
                    }
                }
                if (epi.outT) {
                    // lanes hold consecutive m: each store instruction writes 64 contiguous bytes
#pragma unroll
                    for (int j = 0; j < CH; ++j)
                        if (nb + j < N) epi.outT[(long long)(nb + j) * epi.ldt + m] = __float2bfloat16(v[j]);
                }
            }
        }
    }

    dnb::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        dnb::tc_fence_after();
        dnb::tmem_dealloc(tmem_base, kTmemCols);
    }
}

template <int BN>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const Epi& e, int M, int N, int K, int splits,
           cudaStream_t st) {
    using L = SmemLayout<BN>;
    static bool configured = false;
    if (!configured) {
        DNB_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
        configured = true;
    }
    const int total_kt = (K + BK - 1) / BK;
    int per = (total_kt + splits - 1) / splits;
    int real_splits = (total_kt + per - 1) / per;
    dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, real_splits);
    int stages = per < L::kStages ? per : L::kStages;
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const long ctas = (long)grid.x * grid.y * grid.z;
    if (ctas > sms && ctas <= 2L * sms && BN <= 128) {
        // a bit more than one wave: shrink the ring so that two CTAs fit on an SM (<= ~110 KB each, TMEM
        // 2 x BN columns) and the whole grid is resident at once instead of running a nearly empty second wave
        const int fit = (108 * 1024) / L::kStageBytes;
        if (fit >= 2 && stages > fit) stages = fit;
    }
    if (stages < 1) stages = 1;
    const size_t smem = (size_t)stages * L::kStageBytes + (2 * L::kStages + 1) * 8 + 16 + 1024;
    DNB_CUDA_CHECK(dnb_launch(gemm_bf16_kernel<BN>, grid, dim3(kGemmThreads), smem, st, 0, ta, tb, e, M, N, K, per, stages));
    dnb_count_launch(1);
    return 0;
}

}  // namespace

DNB_EXPORT int dnb_gemm_bf16(const DnbGemmDesc* g, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(g != nullptr, "dnb_gemm_bf16: null descriptor");
    DNB_REQUIRE(g->M > 0 && g->N > 0 && g->K > 0, "dnb_gemm_bf16: empty problem %dx%dx%d", g->M, g->N, g->K);
    DNB_REQUIRE(g->A && g->B, "dnb_gemm_bf16: null operand");
    DNB_REQUIRE(g->out || g->outT || g->colsum, "dnb_gemm_bf16: no output requested");
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int mt = (g->M + BM - 1) / BM;
    const int total_kt = (g->K + BK - 1) / BK;
    int bn = g->block_n;
    if (bn == 0) {
        // pick the tile width that minimises (waves) x (per-tile time ~ fixed cost + BN-proportional work)
        double best = 1e30;
        for (int cand : {256, 128, 64, 32}) {
            if (cand > 32 && g->N <= cand / 2) continue;          // mostly padding
            if (cand == 256 && total_kt < 16) continue;            // short K: the 256-column epilogue dominates
            const long tiles = (long)mt * ((g->N + cand - 1) / cand);
            const long waves = (tiles + sms - 1) / sms;
            const double cost = (double)waves * (6.0 + (double)total_kt * (0.35 + cand / 128.0 * 0.5) + cand / 64.0);
            if (cost < best) { best = cost; bn = cand; }
        }
        if (g->N <= 16) bn = 16;
    }
    int splits = g->split_k;
    const bool can_split = g->out && !g->out_bf16 && !g->outT && !g->relu && !g->mask && !g->bias_n && !g->bias_m &&
                           !g->colsum && g->accumulate;
    if (splits == 0 || (splits == 1 && can_split && g->split_k_auto)) {
        // auto: fill the machine when there are few output tiles and a long K loop (atomic accumulate)
        splits = 1;
        if (can_split) {
            const long tiles = (long)mt * ((g->N + bn - 1) / bn);
            if (tiles < sms) {
                splits = (int)(sms / tiles);
                if (splits > total_kt / 3) splits = total_kt / 3;
                if (splits < 1) splits = 1;
            }
        }
    }
    if (splits < 1) splits = 1;
    if (splits > 1) {
        DNB_REQUIRE(g->out && !g->out_bf16 && !g->outT && !g->relu && !g->mask && !g->bias_n && !g->bias_m &&
                        !g->colsum,
                    "split-K needs a plain fp32 output (it accumulates with atomics)");
    }
    Epi e;
    e.out = g->out; e.ldo = g->ldo; e.out_bf16 = g->out_bf16;
    e.outT = reinterpret_cast<__nv_bfloat16*>(g->outT); e.ldt = g->ldt;
    e.bias_n = g->bias_n; e.bias_m = g->bias_m;
    e.mask = reinterpret_cast<const __nv_bfloat16*>(g->mask); e.ldmask = g->ldmask;
    e.colsum = g->colsum;
    e.alpha = g->alpha; e.relu = g->relu;
    e.atomic_acc = (splits > 1 || g->accumulate) ? 1 : 0;
    e.bn_gamma = g->bn_gamma; e.bn_beta = g->bn_beta; e.bn_eps = g->bn_eps;
    e.bn_xhat = g->bn_xhat; e.ldxh = g->ldxh; e.bn_invstd = g->bn_invstd;
    if (g->bn_gamma) {
        DNB_REQUIRE(splits == 1 && !g->accumulate && !g->bias_n, "BatchNorm epilogue: no split-K / accumulate / column bias");
        if (bn < g->N) {      // the whole batch must sit in one tile
            bn = g->N <= 16 ? 16 : g->N <= 32 ? 32 : g->N <= 64 ? 64 : g->N <= 128 ? 128 : 256;
        }
        DNB_REQUIRE(g->N <= 256, "BatchNorm epilogue needs the whole batch in one tile (N <= 256, got %d)", g->N);
    }
    CUtensorMap ta, tb;
    int rc = dnb_make_tmap_bf16_2d(&ta, g->A, g->M, g->K, g->lda, BM, BK);
    if (rc) return rc;
    rc = dnb_make_tmap_bf16_2d(&tb, g->B, g->N, g->K, g->ldb, bn, BK);
    if (rc) return rc;
    switch (bn) {
        case 16: return launch<16>(ta, tb, e, g->M, g->N, g->K, splits, st);
        case 32: return launch<32>(ta, tb, e, g->M, g->N, g->K, splits, st);
        case 64: return launch<64>(ta, tb, e, g->M, g->N, g->K, splits, st);
        case 128: return launch<128>(ta, tb, e, g->M, g->N, g->K, splits, st);
        case 256: return launch<256>(ta, tb, e, g->M, g->N, g->K, splits, st);
        default: DNB_REQUIRE(false, "block_n must be 0 (auto), 16, 32, 64, 128 or 256 (got %d)", bn);
    }
    return 0;
}
