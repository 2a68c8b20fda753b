// K5b / K5c — persistent LSTM recurrence (forward and BPTT) on tcgen05 with cluster DSMEM.
//
// Math: the reference LSTMCell (comps/icalstm/models.py:20-45) for both directions of the
// bidirectional wrapper (:59-66), quirks included: gate order [i|f|o|g] along 4H, i/f/o pass
// through sigmoid TWICE, g through tanh; the reverse direction consumes the time-flipped
// input and its outputs stay in processing order.
//
// The input projection i2h(x_t) for all timesteps is one big GEMM done beforehand (K5a,
// csrc/gemm.cu); these kernels run only the serial part.  One thread-block cluster handles
// one (direction, batch tile of 16).  CTA c of the H/32 CTAs in the cluster owns hidden units
// [32c, 32c+32): all four gates of those units, their cell state, and the matching slice of
// W_hh, resident in shared memory for the whole sequence.
//
// forward, per step (all-gather design):
//   tcgen05.mma  D[128 gate rows, 16 batch] = W_hh[own rows, 0:H] · h_{t-1}[batch, 0:H]ᵀ
//                (A = weight slice, K-major SW128 smem; B = h_{t-1} in smem, every CTA holds
//                the full vector; accumulator in TMEM)
//   epilogue     tcgen05.ld → + i2h(x_t) + b_hh → σ(σ(·)) / tanh → cell update in registers →
//                h_t slice (bf16) is stored straight into EVERY cluster CTA's next-step B operand
//                through distributed shared memory → barrier.cluster.
// backward, per step (reduce-scatter design):
//   tcgen05.mma  partial dh[H units, 16 batch] = W_hhᵀ[0:H, own gate rows] · dP_t[batch, own rows]ᵀ
//                (K = this CTA's 128 gate rows, so the B operand is purely local)
//   exchange     each warp ships its 32-unit block of the partial to the CTA that owns those
//                units (DSMEM) → barrier.cluster → owner sums the H/32 partials, adds the
//                upstream gradient, back-propagates through the cell and the double sigmoid,
//                emits dP_{t-1} (bf16, to global for the weight-gradient GEMMs and to its local B
//                operand for the next step).
// Activations use tanh.approx (one MUFU op per sigmoid/tanh).
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

constexpr int NB = 16;           // batch tile = MMA N
constexpr int U = 32;            // hidden units per CTA
constexpr int kThreads = 128;
constexpr int kMaxCtas = 8;

struct FwdArgs {
    float* PX;                       // [S*B, D*4H] fp32, time-major rows (row = time*B + b)
    const __nv_bfloat16* Whh[2];     // [4H, H]
    const float* bhh[2];             // [4H]
    float* C;                        // [D][S][B][H] fp32 cell state per processing step
    __nv_bfloat16* Hprev;            // [S*B, D*H] bf16: h_{t-1} aligned with the time row of step t
    float* Hmean;                    // [B, D*H]
    float* hT;                       // [B, D*H]
    float* cT;                       // [B, D*H]
    int B, S, H, D;
};

struct BwdArgs {
    const float* P;                  // [S*B, D*4H] full pre-activations saved by the forward
    const float* C;                  // [D][S][B][H]
    const __nv_bfloat16* Whh[2];
    const float* dHmean;             // [B, D*H] gradient w.r.t. the time-mean of h
    __nv_bfloat16* dP;               // [S*B, D*4H] bf16 out
    float* db_a[2];                  // bias gradients [4H] (i2h and h2h biases get the same values)
    float* db_b[2];
    int accumulate;
    int B, S, H, D;
};

// byte offset of element (row, k) inside a K-major SWIZZLE_128B operand made of 64-wide k-blocks
__device__ __forceinline__ uint32_t sw128_off(int rows_per_block, int row, int k) {
    const int kb = k >> 6, kk = k & 63;
    return (uint32_t)(kb * rows_per_block * 128 + (row >> 3) * 1024 + (row & 7) * 128 +
                      ((((kk >> 3) ^ (row & 7)) & 7) << 4) + (kk & 7) * 2);
}

__device__ __forceinline__ float dsig(float y) { return y * (1.f - y); }

// ------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(kThreads, 1) lstm_fwd_kernel(const FwdArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (dnb::smem_u32(smem_raw) & 1023u)) & 1023u);
    const int H = a.H, KB = (H + 63) / 64;           // k-blocks of the recurrent GEMM
    const int a_bytes = KB * 128 * 128;              // weight slice: 128 rows x (KB*64) k
    const int h_bytes = KB * NB * 128;               // one h buffer
    uint8_t* sA = smem;
    uint8_t* sH = smem + a_bytes;                    // two buffers
    float* sAct = reinterpret_cast<float*>(sH + 2 * h_bytes);   // [4][NB][U]
    uint64_t* mma_bar = reinterpret_cast<uint64_t*>(sAct + 4 * NB * U);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mma_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ncta = dnb::cluster_nctarank();
    const int cta = dnb::cluster_ctarank();
    const int cid = blockIdx.x / ncta;
    const int dir = cid % a.D, b0 = (cid / a.D) * NB;
    const int ldp = a.D * 4 * H, ldh = a.D * H;

    // ---- one-time: weight slice -> swizzled smem, zero h_{-1}
    const __nv_bfloat16* W = a.Whh[dir];
    const int chunks_per_row = H / 8;
    for (int idx = tid; idx < 128 * chunks_per_row; idx += kThreads) {
        const int r = idx / chunks_per_row, ch = idx % chunks_per_row;
        const int grow = (r >> 5) * H + cta * U + (r & 31);
        const uint4 v = *reinterpret_cast<const uint4*>(W + (size_t)grow * H + ch * 8);
        *reinterpret_cast<uint4*>(sA + sw128_off(128, r, ch * 8)) = v;
    }
    if (H & 63) {   // zero the unused tail of the last k-block so stale smem never feeds the MMA
        for (int idx = tid; idx < 128 * (KB * 8 - chunks_per_row); idx += kThreads) {
            const int r = idx / (KB * 8 - chunks_per_row), ch = chunks_per_row + idx % (KB * 8 - chunks_per_row);
            *reinterpret_cast<uint4*>(sA + sw128_off(128, r, ch * 8)) = make_uint4(0, 0, 0, 0);
        }
    }
    for (int idx = tid; idx < 2 * h_bytes / 16; idx += kThreads)
        reinterpret_cast<uint4*>(sH)[idx] = make_uint4(0, 0, 0, 0);
    if (warp == 0) {
        if (lane == 0) {
            dnb::mbar_init(mma_bar, 1);
            dnb::fence_mbar_init();
        }
        __syncwarp();
        dnb::tmem_alloc(tmem_slot, 32);
    }
    dnb::fence_proxy_async();
    dnb::tc_fence_before();
    __syncthreads();
    dnb::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    dnb::cluster_sync_all();       // every CTA's buffers are initialised before anyone writes remotely

    // epilogue-1 role: thread = gate row (q = warp, u = lane)
    const int grow = warp * H + cta * U + lane;
    const float bias = a.bhh[dir][grow];
    // epilogue-2 role: thread = (batch b2, 4 consecutive units)
    const int b2 = tid >> 3, ug = tid & 7;
    const bool b2_ok = (b0 + b2) < a.B;
    float c[4] = {0.f, 0.f, 0.f, 0.f}, hsum[4] = {0.f, 0.f, 0.f, 0.f}, hlast[4] = {0.f, 0.f, 0.f, 0.f};

    // h_{-1} = 0 in the saved sequence
    {
        const int row0 = (dir ? a.S - 1 : 0) * a.B + b0 + b2;
        if (b2_ok) *reinterpret_cast<uint2*>(a.Hprev + (size_t)row0 * ldh + dir * H + cta * U + ug * 4) = make_uint2(0, 0);
    }

    const uint32_t idesc = dnb::umma_idesc_bf16(128, NB);
    const uint32_t sA_addr = dnb::smem_u32(sA);

    for (int t = 0; t < a.S; ++t) {
        const int trow = (dir ? a.S - 1 - t : t) * a.B + b0;
        // i2h(x_t) for this thread's gate row, all 16 batch entries (coalesced across the warp)
        float px[NB];
#pragma unroll
        for (int b = 0; b < NB; ++b)
            px[b] = (b0 + b < a.B) ? a.PX[(size_t)(trow + b) * ldp + dir * 4 * H + grow] : 0.f;

        if (tid == 0) {
            dnb::fence_proxy_async();
            dnb::tc_fence_after();
            const uint32_t sB_addr = dnb::smem_u32(sH + (t & 1) * h_bytes);
            for (int k = 0; k < H / 16; ++k) {
                const uint64_t da = dnb::umma_desc_k_sw128(sA_addr + (k >> 2) * (128 * 128)) + 2 * (k & 3);
                const uint64_t db = dnb::umma_desc_k_sw128(sB_addr + (k >> 2) * (NB * 128)) + 2 * (k & 3);
                dnb::umma_bf16_ss(tmem_base, da, db, idesc, k != 0);
            }
            dnb::umma_commit(mma_bar);
        }
        dnb::mbar_wait(mma_bar, t & 1);
        dnb::tc_fence_after();
        uint32_t acc[NB];
        dnb::tmem_ld_32x16(tmem_base + ((uint32_t)(warp * 32) << 16), acc);
        dnb::tmem_ld_wait();
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            const float p = __uint_as_float(acc[b]) + px[b] + bias;
            if (b0 + b < a.B) a.PX[(size_t)(trow + b) * ldp + dir * 4 * H + grow] = p;   // saved for BPTT
            const float act = (warp < 3) ? dnb::sigmoid_fast(dnb::sigmoid_fast(p)) : dnb::tanh_fast(p);
            sAct[(warp * NB + b) * U + lane] = act;
        }
        dnb::tc_fence_before();
        __syncthreads();

        // cell update: (b2, units ug*4 .. ug*4+3)
        const float4 gi = *reinterpret_cast<const float4*>(sAct + (0 * NB + b2) * U + ug * 4);
        const float4 gf = *reinterpret_cast<const float4*>(sAct + (1 * NB + b2) * U + ug * 4);
        const float4 go = *reinterpret_cast<const float4*>(sAct + (2 * NB + b2) * U + ug * 4);
        const float4 gg = *reinterpret_cast<const float4*>(sAct + (3 * NB + b2) * U + ug * 4);
        const float iv[4] = {gi.x, gi.y, gi.z, gi.w}, fv[4] = {gf.x, gf.y, gf.z, gf.w};
        const float ov[4] = {go.x, go.y, go.z, go.w}, gv[4] = {gg.x, gg.y, gg.z, gg.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            c[j] = fv[j] * c[j] + iv[j] * gv[j];
            hlast[j] = ov[j] * dnb::tanh_fast(c[j]);
            hsum[j] += hlast[j];
        }
        if (b2_ok)
            *reinterpret_cast<float4*>(a.C + (((size_t)dir * a.S + t) * a.B + b0 + b2) * H + cta * U + ug * 4) =
                make_float4(c[0], c[1], c[2], c[3]);
        uint2 hb;
        hb.x = dnb::pack_bf16x2(hlast[0], hlast[1]);
        hb.y = dnb::pack_bf16x2(hlast[2], hlast[3]);
        if (t + 1 < a.S) {
            if (b2_ok) {
                const int nrow = (dir ? a.S - 2 - t : t + 1) * a.B + b0 + b2;
                *reinterpret_cast<uint2*>(a.Hprev + (size_t)nrow * ldh + dir * H + cta * U + ug * 4) = hb;
            }
            // pair two lanes into one 16-byte swizzle chunk and push it into every CTA's next B operand
            uint2 other;
            other.x = __shfl_xor_sync(0xffffffffu, hb.x, 1);
            other.y = __shfl_xor_sync(0xffffffffu, hb.y, 1);
            if ((ug & 1) == 0) {
                const uint4 chunk = make_uint4(hb.x, hb.y, other.x, other.y);
                const int k = cta * U + ug * 4;
                const uint32_t off = dnb::smem_u32(sH + ((t + 1) & 1) * h_bytes) + sw128_off(NB, b2, k);
                for (int peer = 0; peer < ncta; ++peer) dnb::st_cluster_v4(dnb::mapa(off, peer), chunk);
            }
        }
        dnb::fence_proxy_async();
        dnb::cluster_arrive();
        dnb::cluster_wait();
    }

    if (b2_ok) {
        const size_t o = (size_t)(b0 + b2) * ldh + dir * H + cta * U + ug * 4;
        const float inv = 1.f / (float)a.S;
        *reinterpret_cast<float4*>(a.Hmean + o) = make_float4(hsum[0] * inv, hsum[1] * inv, hsum[2] * inv, hsum[3] * inv);
        if (a.hT) *reinterpret_cast<float4*>(a.hT + o) = make_float4(hlast[0], hlast[1], hlast[2], hlast[3]);
        if (a.cT) *reinterpret_cast<float4*>(a.cT + o) = make_float4(c[0], c[1], c[2], c[3]);
    }
    dnb::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        dnb::tc_fence_after();
        dnb::tmem_dealloc(tmem_base, 32);
    }
}

// ------------------------------------------------------------------------------------ backward
__global__ void __launch_bounds__(kThreads, 1) lstm_bwd_kernel(const BwdArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (dnb::smem_u32(smem_raw) & 1023u)) & 1023u);
    const int H = a.H;
    const int MT = (H + 127) / 128;                      // M tiles over hidden units
    const int tile_bytes = 2 * 128 * 128;                // [128 rows][K = 128 own gate rows] = 2 k-blocks
    uint8_t* sA = smem;                                  // MT tiles
    uint8_t* sB = smem + MT * tile_bytes;                // dP_t slice: [NB rows][128 k] = 2 k-blocks x NB x 128 B
    float* sRed = reinterpret_cast<float*>(sB + 2 * NB * 128);   // [2][ncta][U][NB]
    const int ncta = dnb::cluster_nctarank();
    float* sDb = sRed + 2 * ncta * U * NB;               // [128] bias-gradient reduction
    uint64_t* mma_bar = reinterpret_cast<uint64_t*>(sDb + 128);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mma_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int cta = dnb::cluster_ctarank();
    const int cid = blockIdx.x / ncta;
    const int dir = cid % a.D, b0 = (cid / a.D) * NB;
    const int ldp = a.D * 4 * H, ldh = a.D * H;

    // ---- one-time: A[unit k', own gate row j] = W_hh[j, k']  (transposed while loading)
    for (int idx = tid; idx < MT * tile_bytes / 16; idx += kThreads)
        reinterpret_cast<uint4*>(sA)[idx] = make_uint4(0, 0, 0, 0);
    for (int idx = tid; idx < (2 * NB * 128) / 16; idx += kThreads)
        reinterpret_cast<uint4*>(sB)[idx] = make_uint4(0, 0, 0, 0);
    sDb[tid] = 0.f;
    __syncthreads();
    const __nv_bfloat16* W = a.Whh[dir];
    for (int idx = tid; idx < 128 * H; idx += kThreads) {
        const int jl = idx / H, kp = idx % H;            // jl: local gate row (q*32+u), kp: hidden unit
        const int grow = (jl >> 5) * H + cta * U + (jl & 31);
        const __nv_bfloat16 v = W[(size_t)grow * H + kp];
        const int mt = kp >> 7, r = kp & 127;
        *reinterpret_cast<__nv_bfloat16*>(sA + mt * tile_bytes + sw128_off(128, r, jl)) = v;
    }
    if (warp == 0) {
        if (lane == 0) {
            dnb::mbar_init(mma_bar, 1);
            dnb::fence_mbar_init();
        }
        __syncwarp();
        dnb::tmem_alloc(tmem_slot, 32);
    }
    dnb::fence_proxy_async();
    dnb::tc_fence_before();
    __syncthreads();
    dnb::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    dnb::cluster_sync_all();

    const int b2 = tid >> 3, ug = tid & 7;
    const bool b_ok = (b0 + b2) < a.B;
    const int ucol = cta * U + ug * 4;                  // first of this thread's 4 hidden units
    float dc[4] = {0.f, 0.f, 0.f, 0.f};
    float dh_out[4] = {0.f, 0.f, 0.f, 0.f};
    if (b_ok) {
        const float4 v = *reinterpret_cast<const float4*>(a.dHmean + (size_t)(b0 + b2) * ldh + dir * H + ucol);
        const float inv = 1.f / (float)a.S;
        dh_out[0] = v.x * inv; dh_out[1] = v.y * inv; dh_out[2] = v.z * inv; dh_out[3] = v.w * inv;
    }
    float dbacc[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < 4; ++j) dbacc[q][j] = 0.f;

    const uint32_t idesc = dnb::umma_idesc_bf16(128, NB);
    const uint32_t sA_addr = dnb::smem_u32(sA), sB_addr = dnb::smem_u32(sB);

    for (int t = a.S - 1; t >= 0; --t) {
        const int trow = (dir ? a.S - 1 - t : t) * a.B + b0 + b2;
        // saved forward state of my 4 cells (independent of the recurrence -> issue loads first)
        float4 pq[4], cc, cp;
#pragma unroll
        for (int q = 0; q < 4; ++q)
            pq[q] = b_ok ? *reinterpret_cast<const float4*>(a.P + (size_t)trow * ldp + dir * 4 * H + q * H + ucol)
                         : make_float4(0.f, 0.f, 0.f, 0.f);
        cc = b_ok ? *reinterpret_cast<const float4*>(a.C + (((size_t)dir * a.S + t) * a.B + b0 + b2) * H + ucol)
                  : make_float4(0.f, 0.f, 0.f, 0.f);
        cp = (b_ok && t > 0)
                 ? *reinterpret_cast<const float4*>(a.C + (((size_t)dir * a.S + t - 1) * a.B + b0 + b2) * H + ucol)
                 : make_float4(0.f, 0.f, 0.f, 0.f);

        float dh[4] = {dh_out[0], dh_out[1], dh_out[2], dh_out[3]};
        if (t < a.S - 1) {
            const int n = a.S - 2 - t;                  // index of this MMA round
            if (tid == 0) {
                dnb::fence_proxy_async();
                dnb::tc_fence_after();
                for (int mt = 0; mt < MT; ++mt)
                    for (int k = 0; k < 8; ++k) {
                        const uint64_t da =
                            dnb::umma_desc_k_sw128(sA_addr + mt * tile_bytes + (k >> 2) * (128 * 128)) + 2 * (k & 3);
                        const uint64_t db = dnb::umma_desc_k_sw128(sB_addr + (k >> 2) * (NB * 128)) + 2 * (k & 3);
                        dnb::umma_bf16_ss(tmem_base + mt * NB, da, db, idesc, k != 0);
                    }
                dnb::umma_commit(mma_bar);
            }
            dnb::mbar_wait(mma_bar, n & 1);
            dnb::tc_fence_after();
            float* red_par = sRed + (size_t)(n & 1) * ncta * U * NB;
            for (int mt = 0; mt < MT; ++mt) {
                const int dest = mt * 4 + warp;         // CTA owning units [32*dest, 32*dest+32)
                uint32_t v[NB];
                dnb::tmem_ld_32x16(tmem_base + ((uint32_t)(warp * 32) << 16) + mt * NB, v);
                dnb::tmem_ld_wait();
                if (dest < ncta) {
                    const uint32_t local = dnb::smem_u32(red_par + ((size_t)cta * U + lane) * NB);
                    const uint32_t remote = dnb::mapa(local, dest);
#pragma unroll
                    for (int x = 0; x < NB; x += 4)
                        dnb::st_cluster_v4(remote + x * 4, make_uint4(v[x], v[x + 1], v[x + 2], v[x + 3]));
                }
            }
            dnb::tc_fence_before();
            dnb::cluster_arrive();
            dnb::cluster_wait();
            for (int src = 0; src < ncta; ++src) {
#pragma unroll
                for (int j = 0; j < 4; ++j) dh[j] += red_par[((size_t)src * U + ug * 4 + j) * NB + b2];
            }
        }

        const float pi[4] = {pq[0].x, pq[0].y, pq[0].z, pq[0].w}, pf[4] = {pq[1].x, pq[1].y, pq[1].z, pq[1].w};
        const float po[4] = {pq[2].x, pq[2].y, pq[2].z, pq[2].w}, pg[4] = {pq[3].x, pq[3].y, pq[3].z, pq[3].w};
        const float cv[4] = {cc.x, cc.y, cc.z, cc.w}, cpv[4] = {cp.x, cp.y, cp.z, cp.w};
        float dp[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float si = dnb::sigmoid_fast(pi[j]), sf = dnb::sigmoid_fast(pf[j]), so = dnb::sigmoid_fast(po[j]);
            const float i = dnb::sigmoid_fast(si), f = dnb::sigmoid_fast(sf), o = dnb::sigmoid_fast(so);
            const float g = dnb::tanh_fast(pg[j]);
            const float tc = dnb::tanh_fast(cv[j]);
            const float dct = dc[j] + dh[j] * o * (1.f - tc * tc);
            const float d_o = dh[j] * tc;
            dp[0][j] = dct * g * dsig(i) * dsig(si);
            dp[1][j] = dct * cpv[j] * dsig(f) * dsig(sf);
            dp[2][j] = d_o * dsig(o) * dsig(so);
            dp[3][j] = dct * i * (1.f - g * g);
            dc[j] = dct * f;
            if (!b_ok) { dp[0][j] = dp[1][j] = dp[2][j] = dp[3][j] = 0.f; dc[j] = 0.f; }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint2 pk;
            pk.x = dnb::pack_bf16x2(dp[q][0], dp[q][1]);
            pk.y = dnb::pack_bf16x2(dp[q][2], dp[q][3]);
            if (b_ok) *reinterpret_cast<uint2*>(a.dP + (size_t)trow * ldp + dir * 4 * H + q * H + ucol) = pk;
            // local B operand for the next (earlier) step: k = q*32 + ug*4 + j
            *reinterpret_cast<uint2*>(sB + sw128_off(NB, b2, q * U + ug * 4)) = pk;
#pragma unroll
            for (int j = 0; j < 4; ++j) dbacc[q][j] += dp[q][j];
        }
        dnb::fence_proxy_async();
        __syncthreads();
    }

    // bias gradients: sum over batch (threads sharing ug) — once per kernel
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < 4; ++j) atomicAdd(&sDb[q * U + ug * 4 + j], dbacc[q][j]);
    __syncthreads();
    {
        const int gr = (tid >> 5) * H + cta * U + (tid & 31);
        const float v = sDb[tid];
        // batch tiles of the same direction add into the same vector
        if (a.accumulate || gridDim.x > (unsigned)(ncta * a.D)) {
            atomicAdd(a.db_a[dir] + gr, v);
            if (a.db_b[dir]) atomicAdd(a.db_b[dir] + gr, v);
        } else {
            a.db_a[dir][gr] = v;
            if (a.db_b[dir]) a.db_b[dir][gr] = v;
        }
    }
    dnb::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        dnb::tc_fence_after();
        dnb::tmem_dealloc(tmem_base, 32);
    }
    dnb::cluster_sync_all();
}

template <typename Args>
int launch_cluster(void (*kernel)(const Args), const Args& args, int grid, int ncta, size_t smem, cudaStream_t st) {
    DNB_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid, 1, 1);
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ncta;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    DNB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kernel, args));
    dnb_count_launch(1);
    return 0;
}

int check_shape(int B, int S, int H, int D) {
    DNB_REQUIRE(B >= 1 && S >= 1, "lstm: empty batch/sequence");
    DNB_REQUIRE(D == 1 || D == 2, "lstm: directions must be 1 or 2");
    DNB_REQUIRE(H % 32 == 0 && H >= 32 && H <= 32 * kMaxCtas,
                "lstm: per-direction hidden size must be a multiple of 32 in [32, %d] (got %d)", 32 * kMaxCtas, H);
    return 0;
}

}  // namespace

DNB_EXPORT int dnb_lstm_forward(const DnbLstmFwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_lstm_forward: null descriptor");
    if (int rc = check_shape(d->B, d->S, d->H, d->D)) return rc;
    FwdArgs a;
    a.PX = d->PX;
    for (int i = 0; i < 2; ++i) {
        a.Whh[i] = reinterpret_cast<const __nv_bfloat16*>(d->Whh[i]);
        a.bhh[i] = d->bhh[i];
    }
    a.C = d->C; a.Hprev = reinterpret_cast<__nv_bfloat16*>(d->Hprev); a.Hmean = d->Hmean; a.hT = d->hT; a.cT = d->cT;
    a.B = d->B; a.S = d->S; a.H = d->H; a.D = d->D;
    for (int i = 0; i < d->D; ++i)
        DNB_REQUIRE(a.Whh[i] && a.bhh[i] && ((uintptr_t)a.Whh[i] & 15) == 0, "lstm fwd: W_hh[%d] missing or not 16B aligned", i);
    DNB_REQUIRE(a.PX && a.C && a.Hprev && a.Hmean, "lstm fwd: null buffer");
    const int ncta = d->H / U, KB = (d->H + 63) / 64;
    const size_t smem = (size_t)KB * 128 * 128 + 2 * (size_t)KB * NB * 128 + 4 * NB * U * 4 + 64 + 1024;
    const int tiles = (d->B + NB - 1) / NB;
    return launch_cluster(lstm_fwd_kernel, a, ncta * d->D * tiles, ncta, smem, st);
}

DNB_EXPORT int dnb_lstm_backward(const DnbLstmBwdDesc* d, void* stream_) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
    DNB_REQUIRE(d != nullptr, "dnb_lstm_backward: null descriptor");
    if (int rc = check_shape(d->B, d->S, d->H, d->D)) return rc;
    BwdArgs a;
    a.P = d->P; a.C = d->C; a.dHmean = d->dHmean; a.dP = reinterpret_cast<__nv_bfloat16*>(d->dP);
    for (int i = 0; i < 2; ++i) {
        a.Whh[i] = reinterpret_cast<const __nv_bfloat16*>(d->Whh[i]);
        a.db_a[i] = d->db_ih[i];
        a.db_b[i] = d->db_hh[i];
    }
    a.accumulate = d->accumulate;
    a.B = d->B; a.S = d->S; a.H = d->H; a.D = d->D;
    for (int i = 0; i < d->D; ++i) DNB_REQUIRE(a.Whh[i] && a.db_a[i], "lstm bwd: W_hh/db[%d] missing", i);
    DNB_REQUIRE(a.P && a.C && a.dHmean && a.dP, "lstm bwd: null buffer");
    const int ncta = d->H / U, MT = (d->H + 127) / 128;
    const int tiles = (d->B + NB - 1) / NB;
    if (tiles > 1 && !d->accumulate) {
        // several batch tiles add into the same bias-gradient vectors
        for (int i = 0; i < d->D; ++i) {
            DNB_CUDA_CHECK(cudaMemsetAsync(a.db_a[i], 0, sizeof(float) * 4 * d->H, st));
            if (a.db_b[i]) DNB_CUDA_CHECK(cudaMemsetAsync(a.db_b[i], 0, sizeof(float) * 4 * d->H, st));
        }
    }
    const size_t smem = (size_t)MT * 2 * 128 * 128 + 2 * NB * 128 + (size_t)2 * ncta * U * NB * 4 + 128 * 4 + 64 + 1024;
    return launch_cluster(lstm_bwd_kernel, a, ncta * d->D * tiles, ncta, smem, st);
}
