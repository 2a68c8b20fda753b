// K5b / K5c — persistent LSTM recurrence (forward and BPTT) on tcgen05 with cluster DSMEM.
//
// Math: the reference LSTMCell (comps/icalstm/models.py:20-45) for both directions of the
// bidirectional wrapper (:59-66), quirks included: gate order [i|f|o|g] along 4H, i/f/o pass
// through sigmoid TWICE, g through tanh; the reverse direction consumes the time-flipped
// input and its outputs stay in processing order.
//
// The input projection i2h(x_t) for all timesteps is one big GEMM done beforehand (K5a,
// csrc/gemm.cu); these kernels run only the serial part.  One thread-block cluster handles
// one (direction, batch sub-tile).  CTA c of the H/32 CTAs in the cluster owns hidden units
// [32c, 32c+32): all four gates of those units, their cell state, and the matching slice of
// W_hh, resident in TENSOR MEMORY for the whole sequence (tcgen05.mma with the A operand in TMEM:
// with N = 16 an smem A operand is re-read from shared memory by every K=16 instruction, 4 KB at
// 128 B/clk, which bounded the recurrent GEMM).
//
// forward, per step (all-gather design):
//   tcgen05.mma  D[128 gate rows, 16 batch] = W_hh[own rows, 0:H] · h_{t-1}[batch, 0:H]ᵀ
//                (A = weight slice in TMEM; B = h_{t-1}, K-major un-swizzled smem, every CTA holds the
//                full vector; accumulator in TMEM; issued under elect.sync)
//   epilogue     tcgen05.ld → + i2h(x_t) + b_hh → σ(σ(·)) / tanh → cell update in registers →
//                the CTA's h_t slice (bf16, contiguous 512-byte runs in the operand layout) is
//                staged and pushed into EVERY cluster CTA's next-step B operand with cp.async.bulk
//                (shared::cta → shared::cluster); each copy does one complete_tx on the receiver's
//                mbarrier, whose completion releases the receiver's next MMA.  No cluster barrier and
//                no proxy fence between CTAs on the critical path.
// backward, per step (reduce-scatter design):
//   tcgen05.mma  partial dh[H units, 16 batch] = W_hhᵀ[0:H, own gate rows] · dP_t[batch, own rows]ᵀ
//                (K = this CTA's 128 gate rows, so the B operand is purely local)
//   exchange     each warp stages its 32-unit block of the partial (bf16) and bulk-copies it to the
//                CTA that owns those units → that CTA's mbarrier completes when all H/32 blocks have
//                landed → owner sums them in fp32, adds the upstream gradient, back-propagates through the
//                cell and the double sigmoid, emits dP_{t-1} (bf16: to global memory for the weight-gradient
//                GEMMs, and to its local B operand for the next step).
// Global loads for step t±1 and the global stores of step t∓1 are software-pipelined around the MMA.
//
// Batch sub-tiling: the tensor-core tile is always N = 16 batch columns, but a cluster may own only
// NBE = 8 of them (the others stay zero and are never computed on or exchanged).  With few batch rows
// the kernels are bound by per-step latency — per-thread epilogue work, and the ~17-21 B/clk per SM of
// the SM-to-SM network that carries h_t / the partial dh — both proportional to NBE, while most of the
// 148 SMs idle.  Halving NBE doubles the number of clusters and halves both terms.
//
// Activations use tanh.approx (one MUFU op per sigmoid/tanh); the second sigmoid of the reference's
// double-sigmoid gates runs on the FMA pipe (polynomial, its argument is in [0,1]).
#include "common.cuh"
#include <cstdlib>
#include "../include/dinunet_b200.h"

void dnb_count_launch(int n);

namespace {

constexpr int NB = 16;           // MMA N (batch columns of the operand tile)
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
    long long* dbg;                  // optional: mean clock64 stamps per phase (profiling aid)
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

// per-thread roles of the cell update for a batch sub-tile of NBE rows: every thread owns UPT consecutive
// hidden units of one batch row (NBE x 32 units over 128 threads)
template <int NBE> struct Roles {
    static constexpr int UPT = NBE / 4;          // units per thread: 4 (NBE 16), 2 (NBE 8) or 1 (NBE 4)
    static constexpr int TPR = U / UPT;          // threads per batch row: 8, 16 or 32
};

// small fixed-size vector moves (N = 1, 2 or 4)
template <int N> __device__ __forceinline__ void ldv(float (&d)[N], const float* p) {
    if constexpr (N == 4) {
        const float4 v = *reinterpret_cast<const float4*>(p);
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    } else if constexpr (N == 2) {
        const float2 v = *reinterpret_cast<const float2*>(p);
        d[0] = v.x; d[1] = v.y;
    } else {
        d[0] = *p;
    }
}
template <int N> __device__ __forceinline__ void stv(float* p, const float (&s)[N]) {
    if constexpr (N == 4) *reinterpret_cast<float4*>(p) = make_float4(s[0], s[1], s[2], s[3]);
    else if constexpr (N == 2) *reinterpret_cast<float2*>(p) = make_float2(s[0], s[1]);
    else *p = s[0];
}
template <int N> struct BfPack;     // N bf16 values as one 2-, 4- or 8-byte word
template <> struct BfPack<4> { uint2 w; };
template <> struct BfPack<2> { uint32_t w; };
template <> struct BfPack<1> { uint16_t w; };
template <int N> __device__ __forceinline__ BfPack<N> pack_bf(const float (&s)[N]) {
    BfPack<N> r;
    if constexpr (N == 4) { r.w.x = dnb::pack_bf16x2(s[0], s[1]); r.w.y = dnb::pack_bf16x2(s[2], s[3]); }
    else if constexpr (N == 2) r.w = dnb::pack_bf16x2(s[0], s[1]);
    else r.w = __bfloat16_as_ushort(__float2bfloat16_rn(s[0]));
    return r;
}
template <int N> __device__ __forceinline__ BfPack<N> zero_bf() {
    BfPack<N> r;
    if constexpr (N == 4) r.w = make_uint2(0, 0); else r.w = 0;
    return r;
}
template <int N> __device__ __forceinline__ void st_bf(void* p, const BfPack<N>& v) {
    if constexpr (N == 4) *reinterpret_cast<uint2*>(p) = v.w;
    else if constexpr (N == 2) *reinterpret_cast<uint32_t*>(p) = v.w;
    else *reinterpret_cast<uint16_t*>(p) = v.w;
}
template <int N> __device__ __forceinline__ BfPack<N> ld_bf(const void* p) {
    BfPack<N> r;
    if constexpr (N == 4) r.w = *reinterpret_cast<const uint2*>(p);
    else if constexpr (N == 2) r.w = *reinterpret_cast<const uint32_t*>(p);
    else r.w = *reinterpret_cast<const uint16_t*>(p);
    return r;
}
// d[j] += float(v[j])
template <int N> __device__ __forceinline__ void add_bf(float (&d)[N], const BfPack<N>& v) {
    if constexpr (N == 4) {
        d[0] += __uint_as_float(v.w.x << 16); d[1] += __uint_as_float(v.w.x & 0xFFFF0000u);
        d[2] += __uint_as_float(v.w.y << 16); d[3] += __uint_as_float(v.w.y & 0xFFFF0000u);
    } else if constexpr (N == 2) {
        d[0] += __uint_as_float(v.w << 16); d[1] += __uint_as_float(v.w & 0xFFFF0000u);
    } else {
        d[0] += __uint_as_float((uint32_t)v.w << 16);
    }
}
template <int N> __device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&v)[N]) {
    if constexpr (N == 16) dnb::tmem_ld_32x16(taddr, v);
    else if constexpr (N == 8) dnb::tmem_ld_32x8(taddr, v);
    else dnb::tmem_ld_32x4(taddr, v);
}

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

// One recurrent GEMM = KS tcgen05.mma of K = 16.  KS is a template parameter (dispatched by a switch on
// H/16): with a run-time bound every instruction of the unrolled loop is predicated and gets its own
// descriptor / instruction-descriptor register pairs set up on the uniform datapath — ~10 uniform
// instructions per MMA, serial, 300 cycles per step before the first MMA result could even be waited for
// (and a switch inside the kernel costs an indirect branch per step, which is worse).
// Hence KS is a template parameter of the kernel for the common sizes (KS = 0: run-time bound).
template <int KS>
__device__ __forceinline__ void issue_recurrent_mma(uint32_t tmem_d, uint32_t tmem_a, uint64_t db, uint32_t idesc, int ksteps) {
    if constexpr (KS > 0) {
#pragma unroll
        for (int k = 0; k < KS; ++k) dnb::umma_bf16_ts(tmem_d, tmem_a + k * 8, db + koff_h(k), idesc, k != 0);
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k)
            if (k < ksteps) dnb::umma_bf16_ts(tmem_d, tmem_a + k * 8, db + koff_h(k), idesc, k != 0);
    }
}

// ------------------------------------------------------------------------------------ forward
template <int NBE, int KS>
__global__ void __launch_bounds__(kThreads, 1) lstm_fwd_kernel(const FwdArgs a) {
    constexpr int UPT = Roles<NBE>::UPT, TPR = Roles<NBE>::TPR, NG = (NBE + 7) / 8;   // NG = 8-row groups exchanged
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (dnb::smem_u32(smem_raw) & 1023u)) & 1023u);
    const int H = a.H;
    // h_{t-1} as the MMA B operand: K-major, NO swizzle — 8-row x 16-byte core matrices, consecutive
    // k-chunks 128 B apart, the two 8-row groups h_sbo apart.  In this layout the 32 hidden units a CTA
    // produces are one contiguous 512-byte run per row group, so a CTA hands its slice to a peer with
    // NG bulk copies instead of per-thread 16-byte stores.  Row groups >= NG stay zero.
    const int h_sbo = (H / 8) * 128;                 // bytes between the two 8-row groups
    const int h_bytes = NB * H * 2;                  // one h buffer (= 2 * h_sbo)
    uint8_t* sH = smem;                              // two buffers
    float* sAct = reinterpret_cast<float*>(sH + 2 * h_bytes);   // [4][NBE][U]
    uint8_t* sStage = reinterpret_cast<uint8_t*>(sAct + 4 * NBE * U);   // [2 parity][NG row groups][512 B] own h slice
    uint64_t* mma_bar = reinterpret_cast<uint64_t*>(sStage + 2 * NG * 512);
    uint64_t* hbar = mma_bar + 1;                    // [2]: "h_{t-1} has fully arrived in buffer p"
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(hbar + 2);

    const int tid = threadIdx.x, lane = tid & 31;
    const bool mark = a.dbg != nullptr && blockIdx.x == 0 && tid == 0;
    if (mark) a.dbg[12] = clock64();
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);     // provably warp-uniform for the compiler
    const int ncta = dnb::cluster_nctarank();
    const int cta = dnb::cluster_ctarank();
    const int cid = blockIdx.x / ncta;
    const int dir = cid % a.D, b0 = (cid / a.D) * NBE;
    const size_t ldp = (size_t)a.D * 4 * H, ldh = (size_t)a.D * H;
    const int nvalid = min(NBE, a.B - b0);           // real batch rows in this tile
    const uint32_t h_tx_bytes = (uint32_t)(NG * 8 * H * 2);   // whole 8-row groups travel (rows >= NBE are zeros)

    // ---- one-time: zero h_{-1}; the weight slice goes to tensor memory once the allocation exists
    for (int idx = tid; idx < 2 * h_bytes / 16; idx += kThreads)
        reinterpret_cast<uint4*>(sH)[idx] = make_uint4(0, 0, 0, 0);
    for (int idx = tid; idx < 2 * NG * 512 / 16; idx += kThreads)
        reinterpret_cast<uint4*>(sStage)[idx] = make_uint4(0, 0, 0, 0);
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

    dnb::pdl_wait();               // on-chip prologue done; global memory from here on
    // epilogue-1 role: thread = gate row (q = warp, u = lane)
    const int grow = warp * H + cta * U + lane;
    // A operand: lane = gate row (the same row this thread owns in the epilogue), 32-bit column j =
    // k-elements 2j, 2j+1
    {
        const uint4* wrow = reinterpret_cast<const uint4*>((dir ? a.Whh[1] : a.Whh[0]) + (size_t)grow * H);
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
    const float bias = (dir ? a.bhh[1] : a.bhh[0])[grow];   // (no dynamic indexing of kernel arguments: it copies them to local memory)
    // epilogue-2 role: thread = (batch row b2, UPT consecutive units from u0)
    const int b2 = tid / TPR, u0 = (tid % TPR) * UPT;
    const bool b2_ok = b2 < nvalid;
    float c[UPT], hsum[UPT], hlast[UPT];
#pragma unroll
    for (int j = 0; j < UPT; ++j) c[j] = hsum[j] = hlast[j] = 0.f;

    // pointers walk the sequence in processing order (reverse direction: time runs backwards)
    const long long row_step = dir ? -(long long)a.B : (long long)a.B;
    const long long row_first = (long long)(dir ? a.S - 1 : 0) * a.B + b0;
    float* px_ptr = a.PX + (size_t)row_first * ldp + (size_t)dir * 4 * H + grow;          // + b*ldp
    const long long px_step = row_step * (long long)ldp;
    float* c_ptr = a.C + (((size_t)dir * a.S) * a.B + b0 + b2) * H + cta * U + u0;         // step stride B*H
    const size_t c_step = (size_t)a.B * H;
    __nv_bfloat16* hp_ptr = a.Hprev + (size_t)(row_first + b2) * ldh + (size_t)dir * H + cta * U + u0;
    const long long hp_step = row_step * (long long)ldh;

    if (b2_ok) st_bf<UPT>(hp_ptr, zero_bf<UPT>());      // h_{-1} = 0 in the saved sequence

    const uint32_t idesc = dnb::umma_idesc_bf16(128, NB);
    // broadcast from lane 0 so the compiler keeps descriptors / TMEM address in uniform registers
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t sH_u = __shfl_sync(0xffffffffu, dnb::smem_u32(sH), 0);
    const uint64_t db_base0 = dnb::umma_desc_k_noswizzle(sH_u, 128, h_sbo);
    const uint64_t db_base1 = dnb::umma_desc_k_noswizzle(sH_u + h_bytes, 128, h_sbo);
    const int ksteps = H / 16;
    // this thread's UPT units of batch row b2 inside the staged slice: [row group][k-chunk][row][16 B]
    const uint32_t stage_off = (uint32_t)((b2 >> 3) * 512 + (u0 >> 3) * 128 + (b2 & 7) * 16 + (u0 & 7) * 2);
    const uint32_t sStage_u = __shfl_sync(0xffffffffu, dnb::smem_u32(sStage), 0);
    const uint32_t hbar_u = __shfl_sync(0xffffffffu, dnb::smem_u32(&hbar[0]), 0);
    const uint32_t slice_off = (uint32_t)(cta * 4 * 128);     // where this CTA's 4 k-chunks sit inside a row group

    // Software pipeline: i2h(x) for step t+1 is fetched during step t; the global stores of step t
    // (saved pre-activations, cell state, h for the weight-gradient GEMM) are issued at the start of
    // step t+1, while h_t is still in flight between the CTAs.
    float px_next[NBE];
#pragma unroll
    for (int b = 0; b < NBE; ++b) px_next[b] = (b < nvalid) ? px_ptr[(size_t)b * ldp] : 0.f;
    float pst[NBE];
    BfPack<UPT> hb_prev = zero_bf<UPT>();

    DNB_STAMP_DECL;
    if (mark) a.dbg[13] = clock64();
    for (int t = 0; t < a.S; ++t) {
        const bool stamp = a.dbg != nullptr && blockIdx.x == 0 && tid == 0 && t >= a.S / 4 && t < a.S - a.S / 4;
        stamp_n += stamp ? 1 : 0;
        DNB_STAMP(0);
        if (t == a.S - 1 && tid == 0) dnb::pdl_launch();   // last step: the next kernel may start its prologue
        float px[NBE];
#pragma unroll
        for (int b = 0; b < NBE; ++b) px[b] = px_next[b];

        // ---- first, while h_{t-1} is still in flight between the CTAs: deferred stores of step t-1 and
        // the prefetch of step t+1 (the issuing warp does this BEFORE it waits for h, so neither the
        // exchange latency nor the MMA ever waits for these memory instructions)
        if (t > 0) {
            float* pprev = px_ptr - px_step;
#pragma unroll
            for (int b = 0; b < NBE; ++b)
                if (b < nvalid) pprev[(size_t)b * ldp] = pst[b];
            if (b2_ok) {
                stv<UPT>(c_ptr - c_step, c);
                st_bf<UPT>(hp_ptr, hb_prev);              // h_{t-1} at the time row of step t
            }
        }
        if (t + 1 < a.S) {
            const float* pn = px_ptr + px_step;
#pragma unroll
            for (int b = 0; b < NBE; ++b) px_next[b] = (b < nvalid) ? pn[(size_t)b * ldp] : 0.f;
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
                issue_recurrent_mma<KS>(tmem_u, tmem_u + kTmemA, db_base, idesc, ksteps);
                dnb::umma_commit(mma_bar);
            }
            __syncwarp();
        }
        DNB_STAMP(3);
        dnb::mbar_wait(mma_bar, t & 1);
        dnb::tc_fence_after();
        DNB_STAMP(4);
        uint32_t acc[NBE];
        tmem_ld_cols<NBE>(tmem_base + ((uint32_t)(warp * 32) << 16), acc);
        dnb::tmem_ld_wait();
        DNB_STAMP(5);
        // NBE independent activation chains per thread; the gate type is warp-uniform, so the branch
        // sits outside the unrolled loop and the MUFU ops of different batch columns overlap
#pragma unroll
        for (int b = 0; b < NBE; ++b) pst[b] = __uint_as_float(acc[b]) + px[b] + bias;   // saved for BPTT
        float* act_row = sAct + warp * NBE * U + lane;
        if (warp < 3) {
#pragma unroll
            for (int b = 0; b < NBE; ++b) act_row[b * U] = dnb::sigmoid01_poly(dnb::sigmoid_fast(pst[b]));
        } else {
#pragma unroll
            for (int b = 0; b < NBE; ++b) act_row[b * U] = dnb::tanh_fast(pst[b]);
        }
        DNB_STAMP(6);
        dnb::tc_fence_before();
        __syncthreads();
        DNB_STAMP(7);

        // cell update: (b2, units u0 .. u0+UPT-1)
        float iv[UPT], fv[UPT], ov[UPT], gv[UPT];
        ldv<UPT>(iv, sAct + (0 * NBE + b2) * U + u0);
        ldv<UPT>(fv, sAct + (1 * NBE + b2) * U + u0);
        ldv<UPT>(ov, sAct + (2 * NBE + b2) * U + u0);
        ldv<UPT>(gv, sAct + (3 * NBE + b2) * U + u0);
#pragma unroll
        for (int j = 0; j < UPT; ++j) {
            c[j] = fv[j] * c[j] + iv[j] * gv[j];
            hlast[j] = ov[j] * dnb::tanh_fast(c[j]);
            hsum[j] += hlast[j];
        }
        const BfPack<UPT> hb = pack_bf<UPT>(hlast);
        hb_prev = hb;
        DNB_STAMP(8);
        if (t + 1 < a.S) {
            // stage the slice, then (one elected thread per warp) one 512-byte bulk copy per row group and
            // peer: data lands in every CTA's next B operand and counts on that CTA's "h arrived" barrier
            // (async proxy on both ends: no cluster barrier, one mbarrier update per copy)
            const int par = (t + 1) & 1;
            st_bf<UPT>(sStage + par * NG * 512 + stage_off, hb);
            dnb::fence_proxy_async_smem();
            __syncthreads();
            DNB_STAMP(9);
            if (dnb::elect_one()) {                       // warp w serves peers w, w+4
                const uint32_t src = sStage_u + par * NG * 512;
                const uint32_t dst = sH_u + par * h_bytes + slice_off;
                const uint32_t bar = hbar_u + par * 8;
#pragma unroll
                for (int pp = 0; pp < kMaxCtas / 4; ++pp) {
                    const int peer = warp + 4 * pp;
                    if (peer < ncta) {
                        const uint32_t rbar = dnb::mapa(bar, peer);
#pragma unroll
                        for (int g = 0; g < NG; ++g)
                            dnb::bulk_copy_smem_to_cluster(dnb::mapa(dst + g * h_sbo, peer), src + g * 512, 512, rbar);
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
        for (int b = 0; b < NBE; ++b)
            if (b < nvalid) pprev[(size_t)b * ldp] = pst[b];
        if (b2_ok) stv<UPT>(c_ptr - c_step, c);
    }

    if (b2_ok) {
        const size_t o = (size_t)(b0 + b2) * ldh + dir * H + cta * U + u0;
        const float inv = 1.f / (float)a.S;
        float hm[UPT];
#pragma unroll
        for (int j = 0; j < UPT; ++j) hm[j] = hsum[j] * inv;
        stv<UPT>(a.Hmean + o, hm);
        if (a.hT) stv<UPT>(a.hT + o, hlast);
        if (a.cT) stv<UPT>(a.cT + o, c);
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
template <int NBE>
__global__ void __launch_bounds__(kThreads, 1) lstm_bwd_kernel(const BwdArgs a) {
    constexpr int UPT = Roles<NBE>::UPT, TPR = Roles<NBE>::TPR;
    constexpr int BLK = NBE * U;                         // elements of one partial-dh block: [batch row][unit]
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (dnb::smem_u32(smem_raw) & 1023u)) & 1023u);
    const int H = a.H;
    const int MT = (H + 127) / 128;                      // M tiles over hidden units
    const int w_pitch = H * 2 + 16;                      // staging row pitch (+16 B: conflict-free 16-byte stores)
    const int ncta = dnb::cluster_nctarank();
    uint8_t* sW = smem;                                  // one-time staging of the weight slice [128 gate rows][H]
    uint8_t* sB = smem + 128 * w_pitch;                  // dP_t slice: [NB rows][128 k] = 2 k-blocks x NB x 128 B
    __nv_bfloat16* sRed = reinterpret_cast<__nv_bfloat16*>(sB + 2 * NB * 128);   // [2 parity][ncta sources][BLK] incoming
    __nv_bfloat16* sStg = sRed + 2 * ncta * BLK;         // [2 parity][4 warps][2 tiles][BLK] outgoing
    float* sDb = reinterpret_cast<float*>(sStg + 2 * 4 * 2 * BLK);   // [4 warps + 1][128] bias-gradient reduction
    uint64_t* mma_bar = reinterpret_cast<uint64_t*>(sDb + 5 * 128);
    uint64_t* rbar = mma_bar + 1;                        // [2]: "all partial dh blocks of round n arrived"
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(rbar + 2);

    const int tid = threadIdx.x, lane = tid & 31;
    const bool mark = a.dbg != nullptr && blockIdx.x == 0 && tid == 0;
    if (mark) a.dbg[12] = clock64();
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int cta = dnb::cluster_ctarank();
    const int cid = blockIdx.x / ncta;
    const int dir = cid % a.D, b0 = (cid / a.D) * NBE;
    const size_t ldp = (size_t)a.D * 4 * H, ldh = (size_t)a.D * H;
    const int nvalid = min(NBE, a.B - b0);
    const uint32_t red_tx_bytes = (uint32_t)(ncta * BLK * 2);
    const int rounds = a.S - 1;

    // ---- one-time: zero the B operand (rows >= NBE stay zero); without pre-transposed weights, stage this
    // CTA's 128 gate rows of W_hh in shared memory (row-major, coalesced 16-byte loads) for the transposition
    for (int idx = tid; idx < (2 * NB * 128) / 16; idx += kThreads)
        reinterpret_cast<uint4*>(sB)[idx] = make_uint4(0, 0, 0, 0);
    const __nv_bfloat16* W = dir ? a.Whh[1] : a.Whh[0];
    const __nv_bfloat16* WT = dir ? a.WhhT[1] : a.WhhT[0];
    if (WT == nullptr) dnb::pdl_wait();      // this path reads global memory before the allocation below
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
    dnb::pdl_wait();               // on-chip prologue done; global memory from here on
    // A operand in TENSOR MEMORY: tile mt, lane m = hidden unit mt*128 + m, 32-bit column j2 =
    // W_hh[own gate row 2*j2, unit] | W_hh[own gate row 2*j2+1, unit] << 16 — i.e. W_hh^T restricted to
    // this CTA's gate rows, K = 128.
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
        // in-kernel transposition: thread m reads column `unit` of the staged slice (consecutive lanes ->
        // consecutive 2-byte elements: conflict-free)
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

    // thread role: (batch row b2, UPT consecutive units from u0)
    const int b2 = tid / TPR, u0 = (tid % TPR) * UPT;
    const bool b_ok = b2 < nvalid;
    const int ucol = cta * U + u0;                      // first of this thread's hidden units
    float dc[UPT], dh_out[UPT];
#pragma unroll
    for (int j = 0; j < UPT; ++j) dc[j] = dh_out[j] = 0.f;
    if (b_ok) {
        ldv<UPT>(dh_out, a.dHmean + (size_t)(b0 + b2) * ldh + dir * H + ucol);
        const float inv = 1.f / (float)a.S;
#pragma unroll
        for (int j = 0; j < UPT; ++j) dh_out[j] *= inv;
    }
    float dbacc[4][UPT];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < UPT; ++j) dbacc[q][j] = 0.f;

    const uint32_t idesc = dnb::umma_idesc_bf16(128, NB);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t db_base = dnb::umma_desc_k_sw128(__shfl_sync(0xffffffffu, dnb::smem_u32(sB), 0));
    // where this CTA's partial-dh blocks land in an owner's sRed: [parity][source = this CTA]
    const uint32_t red_blk0 = dnb::smem_u32(sRed + (size_t)cta * BLK);      // + parity * red_par_bytes
    const uint32_t red_par_bytes = (uint32_t)(ncta * BLK * 2);
    const uint32_t rbar_u = dnb::smem_u32(&rbar[0]);                        // + parity * 8
    const uint32_t sB_off[4] = {sw128_off(NB, b2, 0 * U + u0), sw128_off(NB, b2, 1 * U + u0),
                                sw128_off(NB, b2, 2 * U + u0), sw128_off(NB, b2, 3 * U + u0)};

    // pointers walk the sequence backwards in processing order
    const long long row_step = dir ? (long long)a.B : -(long long)a.B;      // towards step t-1
    const long long row_last = (long long)(dir ? 0 : a.S - 1) * a.B + b0 + b2;   // time row of step S-1
    const float* p_ptr = a.P + (size_t)row_last * ldp + (size_t)dir * 4 * H + ucol;       // + q*H
    const long long p_step = row_step * (long long)ldp;
    __nv_bfloat16* dp_ptr = a.dP + (size_t)row_last * ldp + (size_t)dir * 4 * H + ucol;
    const float* c_ptr = a.C + (((size_t)dir * a.S + a.S - 1) * a.B + b0 + b2) * H + ucol;
    const size_t c_step = (size_t)a.B * H;

    // Software pipeline (see forward): the saved state of step t-1 is fetched while step t is
    // processed; dP_t goes to global memory one step late; the activation recompute (independent of dh)
    // overlaps the MMA and the DSMEM reduction.
    float pq_n[4][UPT], cc_n[UPT], cp_n[UPT];
#pragma unroll
    for (int j = 0; j < UPT; ++j) cc_n[j] = cp_n[j] = pq_n[0][j] = pq_n[1][j] = pq_n[2][j] = pq_n[3][j] = 0.f;
    if (b_ok) {
#pragma unroll
        for (int q = 0; q < 4; ++q) ldv<UPT>(pq_n[q], p_ptr + q * H);
        ldv<UPT>(cc_n, c_ptr);
        if (a.S > 1) ldv<UPT>(cp_n, c_ptr - c_step);
    }
    BfPack<UPT> dp_store[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) dp_store[q] = zero_bf<UPT>();

    DNB_STAMP_DECL;
    if (mark) a.dbg[13] = clock64();
    for (int t = a.S - 1; t >= 0; --t) {
        const bool stamp = a.dbg != nullptr && blockIdx.x == 0 && tid == 0 && t >= a.S / 4 && t < a.S - a.S / 4;
        stamp_n += stamp ? 1 : 0;
        DNB_STAMP(0);
        if (t == 0 && tid == 0) dnb::pdl_launch();         // last step: the next kernel may start its prologue
        float pq[4][UPT], cv[UPT], cpv[UPT];
#pragma unroll
        for (int j = 0; j < UPT; ++j) {
            cv[j] = cc_n[j]; cpv[j] = cp_n[j];
#pragma unroll
            for (int q = 0; q < 4; ++q) pq[q][j] = pq_n[q][j];
        }

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
            for (int q = 0; q < 4; ++q) st_bf<UPT>(dprev + q * H, dp_store[q]);
        }
        if (t > 0 && b_ok) {
            const float* pn = p_ptr + p_step;
#pragma unroll
            for (int q = 0; q < 4; ++q) ldv<UPT>(pq_n[q], pn + q * H);
            ldv<UPT>(cc_n, c_ptr - c_step);
            if (t > 1) ldv<UPT>(cp_n, c_ptr - 2 * c_step);
            else {
#pragma unroll
                for (int j = 0; j < UPT; ++j) cp_n[j] = 0.f;
            }
        }

        // local Jacobian factors of the cell (recomputed from the saved pre-activations)
        float k_i[UPT], k_f[UPT], k_g[UPT], k_h[UPT], k_tc[UPT], k_fg[UPT];
#pragma unroll
        for (int j = 0; j < UPT; ++j) {
            const float si = dnb::sigmoid_fast(pq[0][j]), sf = dnb::sigmoid_fast(pq[1][j]), so = dnb::sigmoid_fast(pq[2][j]);
            const float i = dnb::sigmoid01_poly(si), f = dnb::sigmoid01_poly(sf), o = dnb::sigmoid01_poly(so);
            const float g = dnb::tanh_fast(pq[3][j]);
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
        for (int j = 0; j < UPT; ++j)
            asm volatile("" ::"f"(k_i[j]), "f"(k_f[j]), "f"(k_g[j]), "f"(k_h[j]), "f"(k_tc[j]), "f"(k_fg[j]));
        float dh[UPT];
#pragma unroll
        for (int j = 0; j < UPT; ++j) dh[j] = dh_out[j];
        DNB_STAMP(2);
        if (t < a.S - 1) {
            dnb::mbar_wait(mma_bar, n & 1);
            dnb::tc_fence_after();
            DNB_STAMP(3);
            // Each warp owns a 32-unit x NBE-batch block of the partial dh per tile; tile mt of warp w
            // belongs to CTA mt*4 + w.  Blocks travel as bf16 (the SM-to-SM network moves ~17-21 B/clk per
            // SM — the exchange, not the MMA, bounds this kernel), laid out [batch row][unit] so that the
            // owner reads its UPT units with one load per source; ONE bulk copy per block = one
            // complete_tx on the owner's barrier.
            uint32_t v[2][NBE];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
                if (mt < MT) tmem_ld_cols<NBE>(tmem_base + ((uint32_t)(warp * 32) << 16) + mt * NB, v[mt]);
            dnb::tmem_ld_wait();
            DNB_STAMP(4);
            __nv_bfloat16* stg = sStg + (((n & 1) * 4 + warp) * 2) * BLK;     // this warp's two blocks of round n
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
                if (mt < MT && mt * 4 + warp < ncta) {
#pragma unroll
                    for (int b = 0; b < NBE; ++b)
                        stg[mt * BLK + b * U + lane] = __float2bfloat16_rn(__uint_as_float(v[mt][b]));
                }
            dnb::fence_proxy_async_smem();                 // one fence for both blocks
            __syncwarp();
            DNB_STAMP(5);
            if (dnb::elect_one()) {
#pragma unroll
                for (int mt = 0; mt < 2; ++mt)
                    if (mt < MT && mt * 4 + warp < ncta)
                        dnb::bulk_copy_smem_to_cluster(dnb::mapa(red_blk0 + (n & 1) * red_par_bytes, mt * 4 + warp),
                                                       dnb::smem_u32(stg + mt * BLK), BLK * 2,
                                                       dnb::mapa(rbar_u + (n & 1) * 8, mt * 4 + warp));
            }
            DNB_STAMP(6);
            dnb::tc_fence_before();
            dnb::mbar_wait(&rbar[n & 1], (n >> 1) & 1);      // all partial blocks for my units arrived
            DNB_STAMP(7);
            const __nv_bfloat16* red_par = sRed + (size_t)(n & 1) * ncta * BLK + b2 * U + u0;
            BfPack<UPT> part[kMaxCtas];
#pragma unroll
            for (int src = 0; src < kMaxCtas; ++src)       // constant offsets; all loads issued before the adds
                part[src] = (src < ncta) ? ld_bf<UPT>(red_par + src * BLK) : zero_bf<UPT>();
#pragma unroll
            for (int src = 0; src < kMaxCtas; ++src) add_bf<UPT>(dh, part[src]);
        }
        DNB_STAMP(8);
        float dp[4][UPT];
#pragma unroll
        for (int j = 0; j < UPT; ++j) {
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
            const BfPack<UPT> pk = pack_bf<UPT>(dp[q]);
            dp_store[q] = pk;
            st_bf<UPT>(sB + sB_off[q], pk);               // local B operand of the next (earlier) step
#pragma unroll
            for (int j = 0; j < UPT; ++j) dbacc[q][j] += dp[q][j];
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
        for (int q = 0; q < 4; ++q) st_bf<UPT>(dprev + q * H, dp_store[q]);
    }

    // bias gradients: sum over the batch rows — lanes that share u0 inside a warp by shuffles, then the
    // four warps through shared memory.  (16-way contended shared atomics here cost ~10 us per launch.)
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < UPT; ++j) {
            float v = dbacc[q][j];
#pragma unroll
            for (int o = TPR; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane < TPR) sDb[warp * 128 + q * U + u0 + j] = v;
        }
    __syncthreads();
    {
        const int gr = (tid >> 5) * H + cta * U + (tid & 31);
        const float v = sDb[tid] + sDb[128 + tid] + sDb[256 + tid] + sDb[384 + tid];
        float* dba = dir ? a.db_a[1] : a.db_a[0];
        float* dbb = dir ? a.db_b[1] : a.db_b[0];
        // batch tiles of the same direction add into the same vector
        if (a.accumulate || gridDim.x > (unsigned)(ncta * a.D)) {
            atomicAdd(dba + gr, v);
            if (dbb) atomicAdd(dbb + gr, v);
        } else {
            dba[gr] = v;
            if (dbb) dbb[gr] = v;
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
    DNB_CUDA_CHECK(dnb_launch(kernel, dim3(grid), dim3(kThreads), smem, st, ncta, args));
    dnb_count_launch(1);
    return 0;
}

// How many clusters of `ncta` CTAs (1 CTA per SM, kThreads threads, `smem` bytes) can be resident at once.
template <typename Args>
int max_active_clusters(void (*kernel)(const Args), int ncta, size_t smem) {
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(ncta, 1, 1);
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ncta;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int check_shape(int B, int S, int H, int D) {
    DNB_REQUIRE(B >= 1 && S >= 1, "lstm: empty batch/sequence");
    DNB_REQUIRE(D == 1 || D == 2, "lstm: directions must be 1 or 2");
    DNB_REQUIRE(H % 32 == 0 && H >= 32 && H <= 32 * kMaxCtas,
                "lstm: hidden size per direction must be a multiple of 32 in [32, %d] (got %d)", 32 * kMaxCtas, H);
    return 0;
}

size_t fwd_smem(int H, int nbe) {
    return 2 * (size_t)NB * H * 2 + 4 * (size_t)nbe * U * 4 + 2 * (size_t)((nbe + 7) / 8) * 512 + 64 + 1024;
}
size_t bwd_smem(int H, int nbe) {
    const int ncta = H / U, blk = nbe * U;
    return (size_t)128 * (H * 2 + 16) + 2 * NB * 128 + (size_t)2 * ncta * blk * 2 + (size_t)2 * 4 * 2 * blk * 2 +
           5 * 128 * 4 + 64 + 1024;
}

// Batch sub-tile: 8 rows per cluster when all the clusters that makes can be resident together (they
// cannot run in waves: a cluster lives for the whole sequence), else 16.  Cached per (direction count is
// part of the cluster count, so only the residency limit is cached).
int pick_nbe(int B, int D, int H, bool backward) {
    static int cached_limit[2][kMaxCtas + 1];      // [backward][ncta] -> max active clusters (+1; 0 = unknown)
    const int ncta = H / U;
    int& c = cached_limit[backward ? 1 : 0][ncta];
    if (c == 0) {
        // the NBE = 8 instantiation has the larger footprint of the two small ones
        const int n = backward ? max_active_clusters(lstm_bwd_kernel<8>, ncta, bwd_smem(H, 8))
                               : max_active_clusters(lstm_fwd_kernel<8, 0>, ncta, fwd_smem(H, 8));
        c = n + 1;
    }
    // DNB_LSTM_NBE=4|8|16 forces the choice (tests exercise every instantiation; a forced value that does
    // not fit dead-locks nothing — clusters are independent — but runs in waves)
    if (const char* e = getenv("DNB_LSTM_NBE")) {
        if (e[0] == '4') return 4;
        if (e[0] == '8') return 8;
        if (e[0] == '1') return 16;
    }
    if (((B + 3) / 4) * D <= c - 1) return 4;
    if (((B + 7) / 8) * D <= c - 1) return 8;
    return 16;
}

template <typename Args>
int launch_nbe(int nbe, void (*k4)(const Args), void (*k8)(const Args), void (*k16)(const Args), const Args& a, int B, int D,
               int ncta, size_t (*smem_of)(int, int), int H, cudaStream_t st) {
    const int tiles = (B + nbe - 1) / nbe;
    void (*k)(const Args) = nbe == 4 ? k4 : (nbe == 8 ? k8 : k16);
    return launch_cluster(k, a, ncta * D * tiles, ncta, smem_of(H, nbe), st);
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
    const int nbe = pick_nbe(d->B, d->D, d->H, false);
    switch (d->H / 16) {      // specialised recurrent-GEMM length for the common hidden sizes
        case 4: return launch_nbe(nbe, lstm_fwd_kernel<4, 4>, lstm_fwd_kernel<8, 4>, lstm_fwd_kernel<16, 4>, a, d->B, d->D, d->H / U, fwd_smem, d->H, st);
        case 8: return launch_nbe(nbe, lstm_fwd_kernel<4, 8>, lstm_fwd_kernel<8, 8>, lstm_fwd_kernel<16, 8>, a, d->B, d->D, d->H / U, fwd_smem, d->H, st);
        case 12: return launch_nbe(nbe, lstm_fwd_kernel<4, 12>, lstm_fwd_kernel<8, 12>, lstm_fwd_kernel<16, 12>, a, d->B, d->D, d->H / U, fwd_smem, d->H, st);
        case 16: return launch_nbe(nbe, lstm_fwd_kernel<4, 16>, lstm_fwd_kernel<8, 16>, lstm_fwd_kernel<16, 16>, a, d->B, d->D, d->H / U, fwd_smem, d->H, st);
        default: break;
    }
    return launch_nbe(nbe, lstm_fwd_kernel<4, 0>, lstm_fwd_kernel<8, 0>, lstm_fwd_kernel<16, 0>, a, d->B, d->D, d->H / U, fwd_smem, d->H, st);
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
    for (int i = 0; i < d->D; ++i) {
        DNB_REQUIRE(a.Whh[i] && a.db_a[i], "lstm bwd: W_hh/db[%d] missing", i);
        DNB_REQUIRE(a.WhhT[i] == nullptr || ((uintptr_t)a.WhhT[i] & 15) == 0, "lstm bwd: W_hh^T[%d] not 16B aligned", i);
    }
    DNB_REQUIRE(a.P && a.C && a.dHmean && a.dP, "lstm bwd: null buffer");
    const int ncta = d->H / U;
    const int nbe = pick_nbe(d->B, d->D, d->H, true);
    const int tiles = (d->B + nbe - 1) / nbe;
    if (tiles > 1 && !d->accumulate) {
        // several batch tiles add into the same bias-gradient vectors
        for (int i = 0; i < d->D; ++i) {
            DNB_CUDA_CHECK(cudaMemsetAsync(a.db_a[i], 0, sizeof(float) * 4 * d->H, st));
            if (a.db_b[i]) DNB_CUDA_CHECK(cudaMemsetAsync(a.db_b[i], 0, sizeof(float) * 4 * d->H, st));
        }
    }
    return launch_nbe(nbe, lstm_bwd_kernel<4>, lstm_bwd_kernel<8>, lstm_bwd_kernel<16>, a, d->B, d->D, ncta, bwd_smem, d->H, st);
}
