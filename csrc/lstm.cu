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
//                (A = weight slice, K-major SW128 smem; B = h_{t-1}, K-major un-swizzled smem, every
//                CTA holds the full vector; accumulator in TMEM; issued under elect.sync)
//   epilogue     tcgen05.ld → + i2h(x_t) + b_hh → σ(σ(·)) / tanh → cell update in registers →
//                the CTA's h_t slice (bf16, two contiguous 512-byte runs in the operand layout) is
//                staged and pushed into EVERY cluster CTA's next-step B operand with cp.async.bulk
//                (shared::cta → shared::cluster); each copy does one complete_tx on the receiver's
//                mbarrier, whose completion releases the receiver's next MMA.  No cluster barrier and
//                no proxy fence between CTAs on the critical path.
// backward, per step (reduce-scatter design):
//   tcgen05.mma  partial dh[H units, 16 batch] = W_hhᵀ[0:H, own gate rows] · dP_t[batch, own rows]ᵀ
//                (K = this CTA's 128 gate rows, so the B operand is purely local)
//   exchange     each warp stages its 32-unit block of the partial (2 KB) and bulk-copies it to the
//                CTA that owns those units → that CTA's mbarrier completes when all H/32 blocks have
//                landed → owner sums them, adds the upstream gradient, back-propagates through the cell
//                and the double sigmoid, emits dP_{t-1} (bf16: to global memory for the weight-gradient
//                GEMMs, and to its local B operand for the next step).
// Global loads for step t±1 and the global stores of step t∓1 are software-pipelined around the MMA.
// Activations use tanh.approx (one MUFU op per sigmoid/tanh).
#include "common.cuh"
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

constexpr int NB = 16;           // batch tile = MMA N
constexpr int U = 32;            // hidden units per CTA
constexpr int kThreads = 128;
constexpr int kMaxCtas = 8;
constexpr int kTmemCols = 256;   // tensor memory per CTA: accumulators in columns [0, 32), the A operand from kTmemA on
constexpr int kTmemA = 32;       // up to 128 columns = 256 k-elements per lane (fwd: H <= 256; bwd: 2 tiles x 128 gate rows)

struct FwdArgs {
    float* PX;                       // [S*B, D*4H] fp32, time-major rows (row = time*B + b)
    const __nv_bfloat16* Whh[2];     // [4H, H]
    const float* bhh[2];             // [4H]
    float* C;                        // [D][S][B][H] fp32 cell state per processing step
    __nv_bfloat16* Hprev;            // [S*B, D*H] bf16: h_{t-1} aligned with the time row of step t
    float* Hmean;                    // [B, D*H]
    float* hT;                       // [B, D*H]
    float* cT;                       // [B, D*H]
    long long* dbg;                  // optional: clock64 stamps of one mid-sequence step (profiling aid)
    int B, S, H, D;
};

struct BwdArgs {
    const float* P;                  // [S*B, D*4H] full pre-activations saved by the forward
    const float* C;                  // [D][S][B][H]
    const __nv_bfloat16* Whh[2];
    const __nv_bfloat16* WhhT[2];    // optional [H, 4H]
    const float* dHmean;             // [B, D*H] gradient w.r.t. the time-mean of h
    __nv_bfloat16* dP;               // [S*B, D*4H] bf16 out
    float* db_a[2];                  // bias gradients [4H] (i2h and h2h biases get the same values)
    float* db_b[2];
    long long* dbg;
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

// Profiling aid: per-phase cycle counts of one thread (block 0, thread 0), averaged over the middle half
// of the sequence in registers and written once at the end (dbg[i] = mean clock at stamp i relative to
// stamp 0, so consecutive differences are mean phase lengths).
#define DNB_STAMP(i)                                                  \
    do {                                                              \
        if (stamp) {                                                  \
            const long long now_ = clock64();                         \
            if ((i) > 0) stamp_acc[(i)] += now_ - stamp_last;         \
            stamp_last = now_;                                        \
        }                                                             \
    } while (0)
#define DNB_STAMP_DECL long long stamp_acc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, stamp_last = 0; int stamp_n = 0
#define DNB_STAMP_FLUSH                                               \
    do {                                                              \
        if (a.dbg != nullptr && blockIdx.x == 0 && tid == 0 && stamp_n > 0) { \
            long long cum_ = 0;                                       \
            _Pragma("unroll") for (int i_ = 0; i_ < 12; ++i_) {                         \
                cum_ += stamp_acc[i_];                                \
                a.dbg[i_] = cum_ / stamp_n;                           \
            }                                                         \
        }                                                             \
    } while (0)

// descriptor address-field increments (units of 16 bytes) of k-step `k` inside operands made of
// 64-wide k-blocks: B blocks are NB rows (2 KB) apart
__device__ __forceinline__ constexpr uint64_t koff_b(int k) { return (uint64_t)((k >> 2) * (NB * 128 / 16) + (k & 3) * 2); }
// forward B operand (h_{t-1}) is stored without swizzle: one K=16 step = two 128-byte core matrices
__device__ __forceinline__ constexpr uint64_t koff_h(int k) { return (uint64_t)(k * (256 / 16)); }

// ------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(kThreads, 1) lstm_fwd_kernel(const FwdArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (dnb::smem_u32(smem_raw) & 1023u)) & 1023u);
    const int H = a.H;
    // h_{t-1} as the MMA B operand: K-major, NO swizzle — 8-row x 16-byte core matrices, consecutive
    // k-chunks 128 B apart, the two 8-row groups h_sbo apart.  In this layout the 32 hidden units a CTA
    // produces are two contiguous 512-byte runs (one per row group), so a CTA hands its slice to a peer
    // with two bulk copies instead of 64 per-thread 16-byte stores.
    const int h_sbo = (H / 8) * 128;                 // bytes between the two 8-row groups
    const int h_bytes = NB * H * 2;                  // one h buffer (= 2 * h_sbo)
    uint8_t* sH = smem;                              // two buffers
    float* sAct = reinterpret_cast<float*>(sH + 2 * h_bytes);   // [4][NB][U]
    uint8_t* sStage = reinterpret_cast<uint8_t*>(sAct + 4 * NB * U);   // [2 parity][2 row groups][512 B] own h slice
    uint64_t* mma_bar = reinterpret_cast<uint64_t*>(sStage + 2 * 1024);
    uint64_t* hbar = mma_bar + 1;                    // [2]: "h_{t-1} has fully arrived in buffer p"
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(hbar + 2);

    const int tid = threadIdx.x, lane = tid & 31;
    const bool mark = a.dbg != nullptr && blockIdx.x == 0 && tid == 0;
    if (mark) a.dbg[12] = clock64();
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);     // provably warp-uniform for the compiler
    const int ncta = dnb::cluster_nctarank();
    const int cta = dnb::cluster_ctarank();
    const int cid = blockIdx.x / ncta;
    const int dir = cid % a.D, b0 = (cid / a.D) * NB;
    const size_t ldp = (size_t)a.D * 4 * H, ldh = (size_t)a.D * H;
    const int nvalid = min(NB, a.B - b0);            // real batch rows in this tile
    const uint32_t h_tx_bytes = (uint32_t)(NB * H * 2);

    // ---- one-time: zero h_{-1}; the weight slice goes to tensor memory once the allocation exists
    for (int idx = tid; idx < 2 * h_bytes / 16; idx += kThreads)
        reinterpret_cast<uint4*>(sH)[idx] = make_uint4(0, 0, 0, 0);
    if (warp == 0) {
        if (lane == 0) {
            dnb::mbar_init(mma_bar, 1);
            dnb::mbar_init(&hbar[0], 1);
            dnb::mbar_init(&hbar[1], 1);
            dnb::fence_mbar_init();
            // arm the first phase of both buffers: step 1 reads buffer 1, step 2 reads buffer 0
            if (a.S > 1) dnb::mbar_expect_tx(&hbar[1], h_tx_bytes);
            if (a.S > 2) dnb::mbar_expect_tx(&hbar[0], h_tx_bytes);
        }
        __syncwarp();
        dnb::tmem_alloc(tmem_slot, kTmemCols);
    }
    dnb::fence_proxy_async();
    dnb::tc_fence_before();
    __syncthreads();
    dnb::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // epilogue-1 role: thread = gate row (q = warp, u = lane)
    const int grow = warp * H + cta * U + lane;
    // A operand = this CTA's 128 gate rows of W_hh, resident in TENSOR MEMORY for the whole sequence:
    // lane = gate row (the same row this thread owns in the epilogue), 32-bit column j = k-elements
    // 2j, 2j+1.  With N = 16 an smem A operand makes every K=16 instruction re-read a 4 KB tile from
    // shared memory (>= 32 cycles at 128 B/clk — the recurrent GEMM was bound by exactly that); from
    // TMEM the same instruction issues at the tensor pipe's own rate.
    {
        const uint4* wrow = reinterpret_cast<const uint4*>(a.Whh[dir] + (size_t)grow * H);
        const uint32_t a_lane = tmem_base + ((uint32_t)(warp * 32) << 16) + kTmemA;
        for (int c0 = 0; c0 < H / 16; c0 += 4) {
            uint4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u)                       // 8 independent 16-byte loads in flight
                if (c0 * 2 + u < H / 8) v[u] = wrow[c0 * 2 + u];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (c0 + u < H / 16) {
                    const uint32_t r[8] = {v[2 * u].x, v[2 * u].y, v[2 * u].z, v[2 * u].w,
                                           v[2 * u + 1].x, v[2 * u + 1].y, v[2 * u + 1].z, v[2 * u + 1].w};
                    dnb::tmem_st_32x8(a_lane + (c0 + u) * 8, r);
                }
        }
        dnb::tmem_st_wait();
    }
    dnb::tc_fence_before();
    dnb::cluster_sync_all();       // every CTA's buffers and barriers exist before anyone writes remotely
    dnb::tc_fence_after();
    const float bias = a.bhh[dir][grow];
    // epilogue-2 role: thread = (batch b2, 4 consecutive units)
    const int b2 = tid >> 3, ug = tid & 7;
    const bool b2_ok = b2 < nvalid;
    float c[4] = {0.f, 0.f, 0.f, 0.f}, hsum[4] = {0.f, 0.f, 0.f, 0.f}, hlast[4] = {0.f, 0.f, 0.f, 0.f};

    // pointers walk the sequence in processing order (reverse direction: time runs backwards)
    const long long row_step = dir ? -(long long)a.B : (long long)a.B;
    const long long row_first = (long long)(dir ? a.S - 1 : 0) * a.B + b0;
    float* px_ptr = a.PX + (size_t)row_first * ldp + (size_t)dir * 4 * H + grow;          // + b*ldp
    const long long px_step = row_step * (long long)ldp;
    float* c_ptr = a.C + (((size_t)dir * a.S) * a.B + b0 + b2) * H + cta * U + ug * 4;     // step stride B*H
    const size_t c_step = (size_t)a.B * H;
    __nv_bfloat16* hp_ptr = a.Hprev + (size_t)(row_first + b2) * ldh + (size_t)dir * H + cta * U + ug * 4;
    const long long hp_step = row_step * (long long)ldh;

    if (b2_ok) *reinterpret_cast<uint2*>(hp_ptr) = make_uint2(0, 0);      // h_{-1} = 0 in the saved sequence

    const uint32_t idesc = dnb::umma_idesc_bf16(128, NB);
    // broadcast from lane 0 so the compiler keeps descriptors / TMEM address in uniform registers
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t sH_u = __shfl_sync(0xffffffffu, dnb::smem_u32(sH), 0);
    const uint64_t db_base0 = dnb::umma_desc_k_noswizzle(sH_u, 128, h_sbo);
    const uint64_t db_base1 = dnb::umma_desc_k_noswizzle(sH_u + h_bytes, 128, h_sbo);
    const int ksteps = H / 16;
    // this thread's 8 bytes (4 units of batch row b2) inside the staged slice: [row group][k-chunk][row][16 B]
    const uint32_t stage_off = (uint32_t)((b2 >> 3) * 512 + (ug >> 1) * 128 + (b2 & 7) * 16 + (ug & 1) * 8);
    const uint32_t sStage_u = __shfl_sync(0xffffffffu, dnb::smem_u32(sStage), 0);
    const uint32_t hbar_u = __shfl_sync(0xffffffffu, dnb::smem_u32(&hbar[0]), 0);
    const uint32_t slice_off = (uint32_t)(cta * 4 * 128);     // where this CTA's 4 k-chunks sit inside a row group

    // Software pipeline: i2h(x) for step t+1 is fetched during step t; the global stores of step t
    // (saved pre-activations, cell state, h for the weight-gradient GEMM) are issued at the start of
    // step t+1 while the MMA runs.  Cross-CTA hand-off of h_t is st.async + mbarrier complete_tx:
    // no fences, no cluster barrier on the critical path.
    float px_next[NB];
#pragma unroll
    for (int b = 0; b < NB; ++b) px_next[b] = (b < nvalid) ? px_ptr[(size_t)b * ldp] : 0.f;
    float pst[NB];
    uint2 hb_prev = make_uint2(0, 0);

    DNB_STAMP_DECL;
    if (mark) a.dbg[13] = clock64();
    for (int t = 0; t < a.S; ++t) {
        const bool stamp = a.dbg != nullptr && blockIdx.x == 0 && tid == 0 && t >= a.S / 4 && t < a.S - a.S / 4;
        stamp_n += stamp ? 1 : 0;
        DNB_STAMP(0);
        float px[NB];
#pragma unroll
        for (int b = 0; b < NB; ++b) px[b] = px_next[b];

        // ---- first, while h_{t-1} is still in flight between the CTAs: deferred stores of step t-1 and
        // the prefetch of step t+1 (the issuing warp does this BEFORE it waits for h, so neither the
        // exchange latency nor the MMA ever waits for these ~35 memory instructions)
        if (t > 0) {
            float* pprev = px_ptr - px_step;
#pragma unroll
            for (int b = 0; b < NB; ++b)
                if (b < nvalid) pprev[(size_t)b * ldp] = pst[b];
            if (b2_ok) {
                *reinterpret_cast<float4*>(c_ptr - c_step) = make_float4(c[0], c[1], c[2], c[3]);
                *reinterpret_cast<uint2*>(hp_ptr) = hb_prev;          // h_{t-1} at the time row of step t
            }
        }
        if (t + 1 < a.S) {
            const float* pn = px_ptr + px_step;
#pragma unroll
            for (int b = 0; b < NB; ++b) px_next[b] = (b < nvalid) ? pn[(size_t)b * ldp] : 0.f;
        }
        DNB_STAMP(1);
        if (warp == 0) {
            // Warp-uniform control flow + elect.sync: the descriptors stay in uniform registers and
            // each tcgen05.mma is a single UTCHMMA (a divergent `tid == 0` branch makes ptxas wrap
            // every MMA in a lane-election loop, ~60 cycles apiece).
            if (t > 0) dnb::mbar_wait(&hbar[t & 1], ((t - 1) >> 1) & 1);   // h_{t-1} complete (complete_tx, as for TMA)
            dnb::tc_fence_after();
            DNB_STAMP(2);
            const uint64_t db_base = (t & 1) ? db_base1 : db_base0;
            if (dnb::elect_one()) {
                if (t > 0 && t + 2 < a.S) dnb::mbar_expect_tx(&hbar[t & 1], h_tx_bytes);   // arm for step t+2
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    if (k < ksteps)
                        dnb::umma_bf16_ts(tmem_u, tmem_u + kTmemA + k * 8, db_base + koff_h(k), idesc, k != 0);
                dnb::umma_commit(mma_bar);
            }
            __syncwarp();
        }
        DNB_STAMP(3);
        dnb::mbar_wait(mma_bar, t & 1);
        dnb::tc_fence_after();
        DNB_STAMP(4);
        uint32_t acc[NB];
        dnb::tmem_ld_32x16(tmem_base + ((uint32_t)(warp * 32) << 16), acc);
        dnb::tmem_ld_wait();
        DNB_STAMP(5);
        // 16 independent activation chains per thread; the gate type is warp-uniform, so the branch
        // sits outside the unrolled loop and the MUFU ops of different batch columns overlap
#pragma unroll
        for (int b = 0; b < NB; ++b) pst[b] = __uint_as_float(acc[b]) + px[b] + bias;   // saved for BPTT
        float* act_row = sAct + warp * NB * U + lane;
        if (warp < 3) {
#pragma unroll
            for (int b = 0; b < NB; ++b) act_row[b * U] = dnb::sigmoid01_poly(dnb::sigmoid_fast(pst[b]));
        } else {
#pragma unroll
            for (int b = 0; b < NB; ++b) act_row[b * U] = dnb::tanh_fast(pst[b]);
        }
        DNB_STAMP(6);
        dnb::tc_fence_before();
        __syncthreads();
        DNB_STAMP(7);

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
        uint2 hb;
        hb.x = dnb::pack_bf16x2(hlast[0], hlast[1]);
        hb.y = dnb::pack_bf16x2(hlast[2], hlast[3]);
        hb_prev = hb;
        DNB_STAMP(8);
        if (t + 1 < a.S) {
            // stage the slice, then (one elected thread) two 512-byte bulk copies per peer: data lands in
            // every CTA's next B operand and counts on that CTA's "h arrived" barrier (async proxy on both
            // ends: no cluster barrier, and one mbarrier update per copy instead of one per 16 bytes)
            const int par = (t + 1) & 1;
            *reinterpret_cast<uint2*>(sStage + par * 1024 + stage_off) = hb;
            dnb::fence_proxy_async_smem();
            __syncthreads();
            DNB_STAMP(9);
            if (dnb::elect_one()) {                       // one thread per warp; warp w serves peers w, w+4
                const uint32_t src = sStage_u + par * 1024;
                const uint32_t dst = sH_u + par * h_bytes + slice_off;
                const uint32_t bar = hbar_u + par * 8;
#pragma unroll
                for (int pp = 0; pp < kMaxCtas / 4; ++pp) {
                    const int peer = warp + 4 * pp;
                    if (peer < ncta) {
                        const uint32_t rbar = dnb::mapa(bar, peer);
                        dnb::bulk_copy_smem_to_cluster(dnb::mapa(dst, peer), src, 512, rbar);
                        dnb::bulk_copy_smem_to_cluster(dnb::mapa(dst + h_sbo, peer), src + 512, 512, rbar);
                    }
                }
            }
        }
        DNB_STAMP(10);
        px_ptr += px_step;
        c_ptr += c_step;
        hp_ptr += hp_step;
        DNB_STAMP(11);
    }
    if (mark) a.dbg[14] = clock64();
    DNB_STAMP_FLUSH;
    // flush the deferred stores of the last step
    {
        float* pprev = px_ptr - px_step;
#pragma unroll
        for (int b = 0; b < NB; ++b)
            if (b < nvalid) pprev[(size_t)b * ldp] = pst[b];
        if (b2_ok) *reinterpret_cast<float4*>(c_ptr - c_step) = make_float4(c[0], c[1], c[2], c[3]);
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
        dnb::tmem_dealloc(tmem_base, kTmemCols);
    }
    dnb::cluster_sync_all();       // nobody leaves while a peer may still address its shared memory
    if (mark) a.dbg[15] = clock64();
}

// ------------------------------------------------------------------------------------ backward
__global__ void __launch_bounds__(kThreads, 1) lstm_bwd_kernel(const BwdArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (dnb::smem_u32(smem_raw) & 1023u)) & 1023u);
    const int H = a.H;
    const int MT = (H + 127) / 128;                      // M tiles over hidden units
    const int w_pitch = H * 2 + 16;                      // staging row pitch (+16 B: conflict-free 16-byte stores)
    uint8_t* sW = smem;                                  // one-time staging of the weight slice [128 gate rows][H]
    uint8_t* sB = smem + 128 * w_pitch;                  // dP_t slice: [NB rows][128 k] = 2 k-blocks x NB x 128 B
    float* sRed = reinterpret_cast<float*>(sB + 2 * NB * 128);   // [2][ncta][U][NB]
    const int ncta = dnb::cluster_nctarank();
    float* sDb = sRed + 2 * ncta * U * NB;               // [128] bias-gradient reduction
    float* sStg = sDb + 128;                             // [2 parity][4 warps][2 tiles][U*NB] outgoing partial blocks
    uint64_t* mma_bar = reinterpret_cast<uint64_t*>(sStg + 2 * 4 * 2 * U * NB);
    uint64_t* rbar = mma_bar + 1;                        // [2]: "all partial dh blocks of round n arrived"
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(rbar + 2);

    const int tid = threadIdx.x, lane = tid & 31;
    const bool mark = a.dbg != nullptr && blockIdx.x == 0 && tid == 0;
    if (mark) a.dbg[12] = clock64();
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int cta = dnb::cluster_ctarank();
    const int cid = blockIdx.x / ncta;
    const int dir = cid % a.D, b0 = (cid / a.D) * NB;
    const size_t ldp = (size_t)a.D * 4 * H, ldh = (size_t)a.D * H;
    const int nvalid = min(NB, a.B - b0);
    const uint32_t red_tx_bytes = (uint32_t)(ncta * U * NB * 4);
    const int rounds = a.S - 1;

    // ---- one-time: this CTA's 128 gate rows of W_hh -> shared memory (row-major, coalesced 16-byte
    // loads); transposed into tensor memory below, once the allocation exists
    for (int idx = tid; idx < (2 * NB * 128) / 16; idx += kThreads)
        reinterpret_cast<uint4*>(sB)[idx] = make_uint4(0, 0, 0, 0);
    sDb[tid] = 0.f;
    const __nv_bfloat16* W = a.Whh[dir];
    const __nv_bfloat16* WT = a.WhhT[dir];
    if (WT == nullptr) {
        const int cpr = H / 8;                            // 16-byte chunks per row
        for (int base = 0; base < 128 * cpr; base += kThreads * 8) {
            uint4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int idx = base + u * kThreads + tid;
                if (idx < 128 * cpr) {
                    const int jl = idx / cpr, ch = idx % cpr;
                    v[u] = *reinterpret_cast<const uint4*>(W + (size_t)((jl >> 5) * H + cta * U + (jl & 31)) * H + ch * 8);
                }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int idx = base + u * kThreads + tid;
                if (idx < 128 * cpr) *reinterpret_cast<uint4*>(sW + (idx / cpr) * w_pitch + (idx % cpr) * 16) = v[u];
            }
        }
    }
    if (warp == 0) {
        if (lane == 0) {
            dnb::mbar_init(mma_bar, 1);
            dnb::mbar_init(&rbar[0], 1);
            dnb::mbar_init(&rbar[1], 1);
            dnb::fence_mbar_init();
            if (rounds > 0) dnb::mbar_expect_tx(&rbar[0], red_tx_bytes);
            if (rounds > 1) dnb::mbar_expect_tx(&rbar[1], red_tx_bytes);
        }
        __syncwarp();
        dnb::tmem_alloc(tmem_slot, kTmemCols);
    }
    dnb::fence_proxy_async();
    dnb::tc_fence_before();
    __syncthreads();
    dnb::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // A operand in TENSOR MEMORY (see the forward kernel): tile mt, lane m = hidden unit mt*128 + m,
    // 32-bit column j2 = W_hh[own gate row 2*j2, unit] | W_hh[own gate row 2*j2+1, unit] << 16 — i.e.
    // W_hh^T restricted to this CTA's gate rows, K = 128.  Thread m reads column `unit` of the staged
    // slice (consecutive lanes -> consecutive 2-byte elements: conflict-free).
    for (int mt = 0; mt < MT; ++mt) {
        const int unit = mt * 128 + tid;
        const uint32_t a_lane = tmem_base + ((uint32_t)(warp * 32) << 16) + kTmemA + mt * 64;
        if (WT != nullptr) {
            // pre-transposed weights: row `unit` of W_hh^T, the four 32-element runs of this CTA's gate rows
            uint4 v[16];
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int x = 0; x < 4; ++x)
                    v[q * 4 + x] = (unit < H) ? *reinterpret_cast<const uint4*>(WT + (size_t)unit * 4 * H + q * H + cta * U + x * 8)
                                              : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint32_t r[8] = {v[2 * c].x, v[2 * c].y, v[2 * c].z, v[2 * c].w,
                                       v[2 * c + 1].x, v[2 * c + 1].y, v[2 * c + 1].z, v[2 * c + 1].w};
                dnb::tmem_st_32x8(a_lane + c * 8, r);
            }
            continue;
        }
#pragma unroll 2
        for (int c = 0; c < 8; ++c) {
            uint32_t r[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int j = (c * 8 + e) * 2;
                uint32_t lo = 0, hi = 0;
                if (unit < H) {
                    lo = *reinterpret_cast<const uint16_t*>(sW + j * w_pitch + unit * 2);
                    hi = *reinterpret_cast<const uint16_t*>(sW + (j + 1) * w_pitch + unit * 2);
                }
                r[e] = lo | (hi << 16);
            }
            dnb::tmem_st_32x8(a_lane + c * 8, r);
        }
    }
    dnb::tmem_st_wait();
    dnb::tc_fence_before();
    dnb::cluster_sync_all();
    dnb::tc_fence_after();

    const int b2 = tid >> 3, ug = tid & 7;
    const bool b_ok = b2 < nvalid;
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
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t db_base = dnb::umma_desc_k_sw128(__shfl_sync(0xffffffffu, dnb::smem_u32(sB), 0));
    // where this warp's partial-dh blocks go: tile mt of warp w belongs to CTA mt*4 + w
    const uint32_t red_blk[2] = {dnb::smem_u32(sRed + (size_t)cta * U * NB),
                                 dnb::smem_u32(sRed + (size_t)ncta * U * NB + (size_t)cta * U * NB)};
    const uint32_t rbar_local[2] = {dnb::smem_u32(&rbar[0]), dnb::smem_u32(&rbar[1])};
    const uint32_t sB_off[4] = {sw128_off(NB, b2, 0 * U + ug * 4), sw128_off(NB, b2, 1 * U + ug * 4),
                                sw128_off(NB, b2, 2 * U + ug * 4), sw128_off(NB, b2, 3 * U + ug * 4)};

    // pointers walk the sequence backwards in processing order
    const long long row_step = dir ? (long long)a.B : -(long long)a.B;      // towards step t-1
    const long long row_last = (long long)(dir ? 0 : a.S - 1) * a.B + b0 + b2;   // time row of step S-1
    const float* p_ptr = a.P + (size_t)row_last * ldp + (size_t)dir * 4 * H + ucol;       // + q*H
    const long long p_step = row_step * (long long)ldp;
    __nv_bfloat16* dp_ptr = a.dP + (size_t)row_last * ldp + (size_t)dir * 4 * H + ucol;
    const float* c_ptr = a.C + (((size_t)dir * a.S + a.S - 1) * a.B + b0 + b2) * H + ucol;
    const size_t c_step = (size_t)a.B * H;

    // Software pipeline (see forward): the saved state of step t-1 is fetched while step t is
    // processed; dP_t goes to global memory one step late; the activation recompute (MUFU-heavy,
    // independent of dh) overlaps the MMA and the DSMEM reduction.
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 pq_n[4], cc_n, cp_n;
#pragma unroll
    for (int q = 0; q < 4; ++q) pq_n[q] = b_ok ? *reinterpret_cast<const float4*>(p_ptr + q * H) : z4;
    cc_n = b_ok ? *reinterpret_cast<const float4*>(c_ptr) : z4;
    cp_n = (b_ok && a.S > 1) ? *reinterpret_cast<const float4*>(c_ptr - c_step) : z4;
    uint2 dp_store[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) dp_store[q] = make_uint2(0, 0);

    DNB_STAMP_DECL;
    if (mark) a.dbg[13] = clock64();
    for (int t = a.S - 1; t >= 0; --t) {
        const bool stamp = a.dbg != nullptr && blockIdx.x == 0 && tid == 0 && t >= a.S / 4 && t < a.S - a.S / 4;
        stamp_n += stamp ? 1 : 0;
        DNB_STAMP(0);
        float4 pq[4], cc = cc_n, cp = cp_n;
#pragma unroll
        for (int q = 0; q < 4; ++q) pq[q] = pq_n[q];

        const int n = a.S - 2 - t;                  // index of this MMA round (valid when t < S-1)
        if (t < a.S - 1 && warp == 0) {
            dnb::tc_fence_after();
            if (dnb::elect_one()) {
#pragma unroll
                for (int mt = 0; mt < 2; ++mt)
                    if (mt < MT) {
#pragma unroll
                        for (int k = 0; k < 8; ++k)
                            dnb::umma_bf16_ts(tmem_u + mt * NB, tmem_u + kTmemA + mt * 64 + k * 8,
                                              db_base + koff_b(k), idesc, k != 0);
                    }
                dnb::umma_commit(mma_bar);
            }
            __syncwarp();
        }
        DNB_STAMP(1);
        // ---- while the MMA runs
        if (t < a.S - 1 && b_ok) {                   // dP of step t+1 -> global
            __nv_bfloat16* dprev = dp_ptr - p_step;
#pragma unroll
            for (int q = 0; q < 4; ++q) *reinterpret_cast<uint2*>(dprev + q * H) = dp_store[q];
        }
        if (t > 0) {
            const float* pn = p_ptr + p_step;
#pragma unroll
            for (int q = 0; q < 4; ++q) pq_n[q] = b_ok ? *reinterpret_cast<const float4*>(pn + q * H) : z4;
            cc_n = b_ok ? *reinterpret_cast<const float4*>(c_ptr - c_step) : z4;
            cp_n = (b_ok && t > 1) ? *reinterpret_cast<const float4*>(c_ptr - 2 * c_step) : z4;
        }

        const float pi[4] = {pq[0].x, pq[0].y, pq[0].z, pq[0].w}, pf[4] = {pq[1].x, pq[1].y, pq[1].z, pq[1].w};
        const float po[4] = {pq[2].x, pq[2].y, pq[2].z, pq[2].w}, pg[4] = {pq[3].x, pq[3].y, pq[3].z, pq[3].w};
        const float cv[4] = {cc.x, cc.y, cc.z, cc.w}, cpv[4] = {cp.x, cp.y, cp.z, cp.w};
        // local Jacobian factors of the cell (recomputed from the saved pre-activations)
        float k_i[4], k_f[4], k_g[4], k_h[4], k_tc[4], k_fg[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float si = dnb::sigmoid_fast(pi[j]), sf = dnb::sigmoid_fast(pf[j]), so = dnb::sigmoid_fast(po[j]);
            const float i = dnb::sigmoid01_poly(si), f = dnb::sigmoid01_poly(sf), o = dnb::sigmoid01_poly(so);
            const float g = dnb::tanh_fast(pg[j]);
            const float tc = dnb::tanh_fast(cv[j]);
            k_h[j] = o * (1.f - tc * tc);            // dh -> dc
            k_tc[j] = tc * dsig(o) * dsig(so);       // dp_o = dh  * k_tc
            k_i[j] = g * dsig(i) * dsig(si);         // dp_i = dct * k_i
            k_f[j] = cpv[j] * dsig(f) * dsig(sf);    // dp_f = dct * k_f
            k_g[j] = i * (1.f - g * g);              // dp_g = dct * k_g
            k_fg[j] = f;                             // dc_{t-1} = dct * f
        }

        // pin the recompute HERE (in the shadow of the MMA and of the exchange): without a use before the
        // waits the compiler sinks it behind them, onto the critical path
#pragma unroll
        for (int j = 0; j < 4; ++j)
            asm volatile("" ::"f"(k_i[j]), "f"(k_f[j]), "f"(k_g[j]), "f"(k_h[j]), "f"(k_tc[j]), "f"(k_fg[j]));
        float dh[4] = {dh_out[0], dh_out[1], dh_out[2], dh_out[3]};
        DNB_STAMP(2);
        if (t < a.S - 1) {
            dnb::mbar_wait(mma_bar, n & 1);
            dnb::tc_fence_after();
            DNB_STAMP(3);
            // Each warp owns a 32-unit x 16-batch block of the partial dh per tile.  It is staged in
            // shared memory (16-byte chunks XOR-swizzled by unit so the owner's column reads spread over
            // banks) and shipped to the owning CTA with ONE bulk copy = one complete_tx on its barrier
            // (per-thread st.async cost one mbarrier update per 16 bytes, 768 per CTA and step).
            uint32_t v[2][NB];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
                if (mt < MT) dnb::tmem_ld_32x16(tmem_base + ((uint32_t)(warp * 32) << 16) + mt * NB, v[mt]);
            dnb::tmem_ld_wait();
            DNB_STAMP(4);
            float* stg = sStg + (((n & 1) * 4 + warp) * 2) * U * NB;     // this warp's two blocks of round n
            const int swz = (lane >> 2) & 3;
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
                if (mt < MT && mt * 4 + warp < ncta) {     // tile mt of warp w belongs to CTA mt*4 + w
#pragma unroll
                    for (int x = 0; x < 4; ++x)
                        *reinterpret_cast<uint4*>(stg + mt * U * NB + lane * NB + ((x ^ swz) << 2)) =
                            make_uint4(v[mt][4 * x], v[mt][4 * x + 1], v[mt][4 * x + 2], v[mt][4 * x + 3]);
                }
            dnb::fence_proxy_async_smem();                 // one fence for both blocks
            __syncwarp();
            DNB_STAMP(5);
            if (dnb::elect_one()) {
#pragma unroll
                for (int mt = 0; mt < 2; ++mt)
                    if (mt < MT && mt * 4 + warp < ncta)
                        dnb::bulk_copy_smem_to_cluster(dnb::mapa(red_blk[n & 1], mt * 4 + warp),
                                                       dnb::smem_u32(stg + mt * U * NB), U * NB * 4,
                                                       dnb::mapa(rbar_local[n & 1], mt * 4 + warp));
            }
            DNB_STAMP(6);
            dnb::tc_fence_before();
            dnb::mbar_wait(&rbar[n & 1], (n >> 1) & 1);      // all partial blocks for my units arrived
            DNB_STAMP(7);
            const float* red_par = sRed + (size_t)(n & 1) * ncta * U * NB + (ug * 4) * NB +
                                   ((((b2 >> 2) ^ (ug & 3)) << 2) | (b2 & 3));
            float part[kMaxCtas][4];
#pragma unroll
            for (int src = 0; src < kMaxCtas; ++src)       // constant offsets; all loads issued before the adds
#pragma unroll
                for (int j = 0; j < 4; ++j) part[src][j] = (src < ncta) ? red_par[(src * U + j) * NB] : 0.f;
#pragma unroll
            for (int src = 0; src < kMaxCtas; ++src)
#pragma unroll
                for (int j = 0; j < 4; ++j) dh[j] += part[src][j];
        }
        DNB_STAMP(8);
        float dp[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float dct = dc[j] + dh[j] * k_h[j];
            dp[0][j] = dct * k_i[j];
            dp[1][j] = dct * k_f[j];
            dp[2][j] = dh[j] * k_tc[j];
            dp[3][j] = dct * k_g[j];
            dc[j] = dct * k_fg[j];
            if (!b_ok) { dp[0][j] = dp[1][j] = dp[2][j] = dp[3][j] = 0.f; dc[j] = 0.f; }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint2 pk;
            pk.x = dnb::pack_bf16x2(dp[q][0], dp[q][1]);
            pk.y = dnb::pack_bf16x2(dp[q][2], dp[q][3]);
            dp_store[q] = pk;
            *reinterpret_cast<uint2*>(sB + sB_off[q]) = pk;      // local B operand of the next (earlier) step
#pragma unroll
            for (int j = 0; j < 4; ++j) dbacc[q][j] += dp[q][j];
        }
        DNB_STAMP(9);
        p_ptr += p_step;
        dp_ptr += p_step;
        c_ptr -= c_step;
        dnb::fence_proxy_async_smem();               // generic-proxy writes of sB -> visible to the MMA
        DNB_STAMP(10);
        __syncthreads();
        if (tid == 0 && t < a.S - 1 && n + 2 < rounds) dnb::mbar_expect_tx(&rbar[n & 1], red_tx_bytes);   // re-arm
        DNB_STAMP(11);
    }
    if (mark) a.dbg[14] = clock64();
    DNB_STAMP_FLUSH;
    if (b_ok) {                                      // dP of step 0
        __nv_bfloat16* dprev = dp_ptr - p_step;
#pragma unroll
        for (int q = 0; q < 4; ++q) *reinterpret_cast<uint2*>(dprev + q * H) = dp_store[q];
    }

    // bias gradients: sum over the batch rows — lanes that share ug inside a warp by shuffles (rows 4w..4w+3),
    // then the four warps through shared memory (sRed is free: every incoming round has been consumed).
    // (16-way contended shared atomics here cost ~10 us per launch.)
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float v = dbacc[q][j];
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            if (lane < 8) sRed[warp * 128 + q * U + ug * 4 + j] = v;
        }
    __syncthreads();
    sDb[tid] = sRed[tid] + sRed[128 + tid] + sRed[256 + tid] + sRed[384 + tid];
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
        dnb::tmem_dealloc(tmem_base, kTmemCols);
    }
    dnb::cluster_sync_all();
    if (mark) a.dbg[15] = clock64();
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
    a.dbg = d->dbg;
    a.B = d->B; a.S = d->S; a.H = d->H; a.D = d->D;
    for (int i = 0; i < d->D; ++i)
        DNB_REQUIRE(a.Whh[i] && a.bhh[i] && ((uintptr_t)a.Whh[i] & 15) == 0, "lstm fwd: W_hh[%d] missing or not 16B aligned", i);
    DNB_REQUIRE(a.PX && a.C && a.Hprev && a.Hmean, "lstm fwd: null buffer");
    const int ncta = d->H / U;
    const size_t smem = 2 * (size_t)NB * d->H * 2 + 4 * NB * U * 4 + 2 * 1024 + 128 + 1024;
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
        a.WhhT[i] = reinterpret_cast<const __nv_bfloat16*>(d->WhhT[i]);
        a.db_a[i] = d->db_ih[i];
        a.db_b[i] = d->db_hh[i];
    }
    a.accumulate = d->accumulate;
    a.dbg = d->dbg;
    a.B = d->B; a.S = d->S; a.H = d->H; a.D = d->D;
    for (int i = 0; i < d->D; ++i) DNB_REQUIRE(a.Whh[i] && a.db_a[i], "lstm bwd: W_hh/db[%d] missing", i);
    DNB_REQUIRE(a.P && a.C && a.dHmean && a.dP, "lstm bwd: null buffer");
    const int ncta = d->H / U;
    const int tiles = (d->B + NB - 1) / NB;
    if (tiles > 1 && !d->accumulate) {
        // several batch tiles add into the same bias-gradient vectors
        for (int i = 0; i < d->D; ++i) {
            DNB_CUDA_CHECK(cudaMemsetAsync(a.db_a[i], 0, sizeof(float) * 4 * d->H, st));
            if (a.db_b[i]) DNB_CUDA_CHECK(cudaMemsetAsync(a.db_b[i], 0, sizeof(float) * 4 * d->H, st));
        }
    }
    const size_t smem = (size_t)128 * (d->H * 2 + 16) + 2 * NB * 128 + (size_t)2 * ncta * U * NB * 4 + 128 * 4 + (size_t)2 * 4 * 2 * U * NB * 4 + 128 + 1024;
    return launch_cluster(lstm_bwd_kernel, a, ncta * d->D * tiles, ncta, smem, st);
}
