/* C ABI of libdinunet_b200.so — hand-written sm_100a kernels + the small host runtime
 * around them.  No PyTorch dependency: callers pass raw device pointers and a CUDA stream.
 * Python binds these through ctypes (dinunet_implementations_b200/ops/native.py).
 *
 * Conventions
 *   - every function returns 0 on success; on failure dnb_last_error() describes why;
 *   - `stream` is a cudaStream_t passed as void*;
 *   - matrices are row-major; "ld" is the leading dimension in elements;
 *   - kernels that wait on other GPUs or on mbarriers bound their spins and trap instead of
 *     hanging.
 *
 * Reference citations give the math each entry point reproduces (paths relative to the
 * reference tree).
 */
#ifndef DINUNET_B200_H
#define DINUNET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DNB_MAX_RANKS 8
#define DNB_SIGNAL_BLOCKS 160
#define DNB_ORTH_MAX_BATCH 32
#define DNB_ORTH_MAX_COLS 64
/* uint32 slots a signal pad needs: 2 phases x blocks x ranks */
#define DNB_SIGNAL_WORDS (2 * DNB_SIGNAL_BLOCKS * DNB_MAX_RANKS)

/* ---- runtime ------------------------------------------------------------------------- */
const char* dnb_last_error(void);
int dnb_version(void);
long long dnb_launch_count(void);      /* kernels launched by this library so far */
void dnb_reset_launch_count(void);

/* symmetric arena: cudaMalloc'ed, zeroed, exportable through CUDA IPC */
int dnb_arena_alloc(size_t bytes, void** out);
int dnb_arena_free(void* p);
int dnb_ipc_handle_size(void);
int dnb_ipc_get_handle(void* dptr, void* handle_out);
int dnb_ipc_open_handle(const void* handle, void** out);
int dnb_ipc_close_handle(void* p);
int dnb_enable_peer_access(int peer_device);
/* Plain asynchronous copy on `stream` (kind: 1 host->device, 2 device->host, 3 device->device).  The pipelined
 * training step uses it to stage batches without per-step framework overhead (stream-context switches and
 * tensor allocations cost more host time than the whole GPU step takes). */
int dnb_copy_async(void* dst, const void* src, size_t bytes, int kind, void* stream);

/* COINSTAC file mode: mean over n_src flat arrays (fp32, or fp16 when is_half) -> fp32.
 * Replaces the aggregator's multiprocessing pool (reference remote.py:20-21). */
int dnb_cpu_mean_f32(const void* const* srcs, int n_src, long long n, int is_half, float* out, int n_threads);

/* Native parser for one FreeSurfer aseg-stats file (reference comps/fs/__init__.py:33-39): returns the
 * number of values written to out (<0 on error); normalise != 0 divides by the subject's maximum. */
int dnb_parse_aseg(const char* path, double* out, int max_n, int normalise);
/* Host-side modified Gram-Schmidt of the r columns of P [n, r] (row-major fp32, in place; fp64 accumulation) with the
 * same dead-column rule as the CUDA kernel and distrib/lowrank.orthogonalize: a column whose residual norm is
 * <= max(eps * ||P||_F, rel * its original norm) is zeroed.  The file-mode rankDAD / powerSGD reducers call it
 * thousands of times per run on 10-column factors; the Python loop was 40 % of a simulated rankDAD run. */
int dnb_cpu_orthogonalize(float* P, long long n, int r, float eps, float rel);

/* ---- K1: dSGD all-reduce fused with Adam/SGD (reference engine "dSGD", comps/__init__.py:14;
 *      precision_bits compspec.json:161-176) ----------------------------------------------- */
typedef struct DnbDsgdDesc {
    const void* grads[DNB_MAX_RANKS];   /* every rank's flat gradient buffer (wire dtype), peer-mapped */
    float* params[DNB_MAX_RANKS];       /* every rank's flat fp32 parameter buffer, peer-mapped */
    void* shadow_bf16[DNB_MAX_RANKS];   /* optional bf16 copy of the parameters (NULL to skip) */
    uint32_t* flags[DNB_MAX_RANKS];     /* every rank's signal pad, DNB_SIGNAL_WORDS uint32, zero-initialised */
    float* m;                           /* local optimiser state, length n (Adam m / SGD momentum) */
    float* v;                           /* local Adam v, length n */
    float* out;                         /* opt == 2: reduced mean goes here */
    long long n;                        /* flat length, multiple of 8 */
    int nranks, rank;
    int wire;                           /* 0 fp32, 1 bf16, 2 fp16 */
    int mode;                           /* 0 two-shot (shard + push params), 1 one-shot (pull all), 2 two-shot over NVLS multicast */
    int opt;                            /* 0 Adam, 1 SGD(+momentum), 2 none (reduce only) */
    int use_flags;                      /* 0: no cross-GPU barrier (single-GPU emulation of S sites) */
    int step;                           /* optimiser step count t >= 1 (Adam bias correction) */
    float lr, beta1, beta2, eps, weight_decay, momentum;
    float grad_scale;                   /* 1 / (nranks * loss_scale) */
    uint32_t epoch;                     /* strictly increasing per launch; barrier generation */
    uint32_t* dev_state;                /* optional device words {steps done, scratch counter, lr as float bits, 0}:
                                           when set, step / epoch (= epoch + steps done + 1) / lr are read on the
                                           device and the step count is advanced by the kernel, so the identical
                                           launch can be replayed from a CUDA graph */
    float* stage;                       /* optional (16-bit wire only): this rank's fp32 gradients.  The kernel casts them
                                           into grads[rank] itself before the barrier (each block packs exactly the elements
                                           the same-index block of every peer will read) — no separate cast launch */
    int zero_stage;                     /* also clear `stage` while packing: the next step needs no zero-fill launch */
    /* mode 2 (NVLS): multicast (NVSwitch) addresses of the SAME symmetric gradient / parameter / shadow buffers.
     * Rank r reduces its shard with multimem.ld_reduce (the switch adds the sites' values; 16-bit wires accumulate in
     * fp32 and round once), applies the optimiser and publishes the updated values with multimem.st (one store
     * reaches every site).  Per GPU 1/N of the buffer crosses NVLink each way instead of (N-1)/N. */
    const void* grads_mc;
    float* params_mc;
    void* shadow_mc;
} DnbDsgdDesc;
int dnb_dsgd_fused_step(const DnbDsgdDesc* desc, void* stream);

/* ---- tcgen05 + TMA GEMM with fused epilogue --------------------------------------------
 * D[M,N] = alpha * A[M,K] * B[N,K]^T ; A and B bf16 row-major (lda, ldb multiples of 8),
 * fp32 accumulation in TMEM.  Used for: encoder Linear+ReLU (reference comps/icalstm/models.py:87,107),
 * LSTM input projection i2h for all timesteps (models.py:31), the wide FreeSurfer layers
 * (comps/fs/models.py:13-25), their weight/data-gradient GEMMs, and the powerSGD products. */
typedef struct DnbGemmDesc {
    const void* A; long long lda;       /* [M,K] bf16 */
    const void* B; long long ldb;       /* [N,K] bf16 */
    int M, N, K;
    void* out; long long ldo;           /* [M,N] fp32 or bf16 (may be NULL) */
    int out_bf16;
    void* outT; long long ldt;          /* optional transposed bf16 copy [N,M] */
    const float* bias_n;                /* optional [N], added per column */
    const float* bias_m;                /* optional [M], added per row */
    const void* mask; long long ldmask; /* optional bf16 [M,N]: result zeroed where mask <= 0 (ReLU backward) */
    float* colsum;                      /* optional [N]: += column sums of the result (bias gradient) */
    float alpha;
    int relu;
    int accumulate;                     /* fp32 out += result (atomic) */
    int split_k;                        /* >1: split the K loop over grid.z (plain fp32 out, atomically accumulated);
                                           0: choose automatically when `accumulate` is set */
    int block_n;                        /* 0 = auto, else 16/32/64/128/256 */
    int split_k_auto;                   /* reserved (0) */
    /* BatchNorm-over-columns epilogue (reference comps/fs/models.py:15,19: Linear -> BatchNorm1d with batch
     * statistics -> ReLU): rows = features (weights are the A operand), columns = the whole batch (N <= 256).
     * result = [relu](gamma * (x - mean_row) * invstd_row + beta).  Enabled when bn_gamma != NULL. */
    const float* bn_gamma; const float* bn_beta;   /* [M] */
    float bn_eps;
    float* bn_xhat; long long ldxh;     /* optional [M,N] normalised values */
    float* bn_invstd;                   /* optional [M] */
} DnbGemmDesc;
int dnb_gemm_bf16(const DnbGemmDesc* desc, void* stream);

/* ---- K5b/K5c: persistent bidirectional LSTM recurrence, forward and BPTT -----------------
 * Reference cell: comps/icalstm/models.py:20-45 (double sigmoid on i/f/o, gate order i|f|o|g),
 * bidirectional wrapper :59-66 (reverse direction on flipped time, outputs not flipped back).
 * Rows of all [S*B, .] tensors are time-major: row = time * B + b.  H = per-direction hidden
 * size, multiple of 32, <= 256.  D = 1 or 2 directions. */
typedef struct DnbLstmFwdDesc {
    float* PX;               /* [S*B, D*4H] in: i2h(x)+b_ih for every timestep; out: full pre-activations */
    const void* Whh[2];      /* per direction [4H, H] bf16, 16-byte aligned */
    const float* bhh[2];     /* per direction [4H] */
    float* C;                /* [D, S, B, H] cell states per processing step (saved for backward) */
    void* Hprev;             /* [S*B, D*H] bf16: h_{t-1} stored at the time row of step t */
    float* Hmean;            /* [B, D*H] mean over time of h (reference models.py:109) */
    float* hT; float* cT;    /* [B, D*H] final states (may be NULL) */
    long long* dbg;          /* optional: 16 clock64 stamps of one mid-sequence step (NULL normally) */
    int B, S, H, D;
} DnbLstmFwdDesc;
int dnb_lstm_forward(const DnbLstmFwdDesc* desc, void* stream);

typedef struct DnbLstmBwdDesc {
    const float* P;          /* [S*B, D*4H] pre-activations from the forward */
    const float* C;          /* [D, S, B, H] */
    const void* Whh[2];
    const void* WhhT[2];     /* optional bf16 [H, 4H] = Whh transposed: lets the kernel fill its tensor-memory operand
                              * straight from global memory instead of transposing the slice in shared memory */
    const float* dHmean;     /* [B, D*H] gradient w.r.t. Hmean */
    void* dP;                /* [S*B, D*4H] bf16 out: gradient w.r.t. the pre-activations */
    float* db_ih[2];         /* per direction [4H] bias gradient (required) */
    float* db_hh[2];         /* same values, second bias (may be NULL) */
    long long* dbg;          /* optional profiling stamps, as in the forward */
    int accumulate;          /* add to the bias gradients instead of overwriting */
    int B, S, H, D;
} DnbLstmBwdDesc;
int dnb_lstm_backward(const DnbLstmBwdDesc* desc, void* stream);

/* ---- layout / cast helpers ---------------------------------------------------------------- */
/* x fp32 [B,S,F] -> rows bf16 [S*B, ldr] (row = t*B + b) and, if colsT != NULL, its transpose
 * colsT bf16 [F, ldt].  Replaces the per-sample Python encoder loop (reference
 * comps/icalstm/models.py:107) with one time-major matrix for the encoder GEMM. */
int dnb_cast_time_major(const float* x, void* rows, long long ldr, void* colsT, long long ldt, int B, int S, int F,
                        void* stream);
int dnb_transpose_bf16(const void* in, long long ldi, void* out, long long ldo, int R, int C, void* stream);
int dnb_cast_f32_bf16(const float* in, void* out, long long n, void* stream);
/* K10: [N,C,T] -> [N,S,C,W] windows (reference comps/icalstm/__init__.py:28-32) */
int dnb_window_gather(const float* in, float* out, int N, int C, int T, int S, int W, int stride, void* stream);
/* Device-resident batches (replaces DataLoader + collate + per-step H2D; reference comps/icalstm/__init__.py:36-38,
 * comps/fs/__init__.py:33-39): out[b] = table[perm[(cursor + b) % n_perm]] for b < B, labels / row indices likewise;
 * state = {cursor, block counter} (two uint64, zero-initialised) is advanced by B on the device, so one captured launch
 * walks through the epoch when replayed.  labels, labels_out and index_out may be NULL. */
int dnb_batch_gather(const float* table, long long row_floats, const long long* labels, const long long* perm,
                     long long n_perm, unsigned long long* state, int B, float* out, long long* labels_out,
                     long long* index_out, void* stream);
/* K6: softmax + mean cross-entropy forward and gradient in one pass (reference
 * comps/icalstm/__init__.py:60-61, comps/fs/__init__.py:54-55).  prob / dlogits / loss may be NULL;
 * log_softmax_out != 0 writes log-probabilities into prob. */
int dnb_softmax_ce(const float* logits, const long long* labels, float* prob, float* dlogits, float* loss, int B,
                   int ncls, int log_softmax_out, void* stream);

/* ---- K8: ICA-LSTM classifier head (reference comps/icalstm/models.py:95-103) --------------- */
typedef struct DnbHeadFwdDesc {
    const float* x;            /* [B,N0] time-mean of the LSTM outputs */
    const float* drop;         /* [B,N0] dropout scale (0 or 1/(1-p)), NULL when not training */
    const void *W1, *W2, *W3;  /* bf16 [N1,N0], [N2,N1], [N3,N2] */
    const float *b1, *b2, *b3, *bn_w, *bn_b;
    float *run_mean, *run_var; /* BatchNorm running statistics (updated in training, used in eval) */
    float momentum, eps;
    int training;
    float* a0;                 /* [B,N0] dropped input (NULL: not saved) */
    float *xhat, *invstd, *a1, *a2, *logits;   /* [B,N1], [N1], [B,N1], [B,N2], [B,N3] */
    int B, N0, N1, N2, N3;
    /* optional, all may be 0/NULL.  In-kernel dropout: when drop == NULL, drop_p > 0 and drop_state != NULL the
     * kernel draws the mask itself (counter-based hash of drop_seed, *drop_state and the element index), writes
     * the scale values (0 or 1/(1-p)) to drop_out [B,N0] for the backward pass and advances *drop_state by one.
     * num_batches_tracked (int64 scalar) is incremented in training mode (BatchNorm1d bookkeeping). */
    float* drop_out;
    float drop_p;
    unsigned long long drop_seed;
    unsigned long long* drop_state;
    long long* num_batches_tracked;
} DnbHeadFwdDesc;
int dnb_head_forward(const DnbHeadFwdDesc* desc, void* stream);

typedef struct DnbHeadBwdDesc {
    const float* dlogits;      /* [B,N3] */
    const float *a0, *drop, *xhat, *invstd, *a1, *a2;
    const void *W1, *W2, *W3;
    const float* bn_w;
    float *dW1, *db1, *dbn_w, *dbn_b, *dW2, *db2, *dW3, *db3;  /* parameter gradients (fp32) */
    float* dx;                 /* [B,N0] gradient w.r.t. x */
    float* dz1;                /* [B,N1] scratch handed from the first to the second backward launch */
    int training, accumulate;
    int B, N0, N1, N2, N3;
    float* dz2;                /* optional [B,N2] out: gradient at the second layer's pre-activation (rankDAD pairs) */
} DnbHeadBwdDesc;
int dnb_head_backward(const DnbHeadBwdDesc* desc, void* stream);

/* Allocation-free variant of dnb_head_backward (no counterpart in the reference: it exists for
 * host programs that may not allocate device memory on this path).  Same inputs and outputs;
 * desc->dz1 is ignored and the [B,N1] scratch is a 128 KiB __device__ array inside the library,
 * held by every GPU that loads it.  Limits: B <= 64, N1 <= 512; anything larger is refused.  No
 * cudaMalloc/cudaFree.  The first call on a device looks up the array's device address — make it
 * before any stream capture on that device.  Calls share the array: overlapping calls on one GPU race.
 * The device owning `stream` must be the current device; the cached address does not survive cudaDeviceReset. */
int dnb_head_backward_noalloc(const DnbHeadBwdDesc* desc, void* stream);

/* ---- K4: FreeSurfer MLP hidden layer backward through ReLU + BatchNorm(batch statistics)
 * (reference comps/fs/models.py:13-23).  Activations are feature-major: daT / xhatT are [N, B]
 * (leading dims ldd / ldx).  Writes dz as bf16 feature-major (dzT [N, ldzt]) and/or batch-major
 * (dz [B, ldz]) for the weight- and data-gradient GEMMs, and the gamma/beta gradients. */
int dnb_bn_relu_backward(const float* daT, long long ldd, const float* xhatT, long long ldx, const float* gamma,
                         const float* beta, const float* invstd, int N, int B, void* dzT, long long ldzt, void* dz,
                         long long ldz, float* dgamma, float* dbeta, int accumulate, void* stream);

/* ---- K4, default size: the whole FreeSurfer MLP (reference comps/fs/models.py:13-31; default 66-256-128-64-32-2,
 * compspec.json:228-235) in one launch forward and one launch backward, one CTA, activations in shared memory.
 * Hidden layer l: z = a W_l^T (W_l bf16 [width_l, width_{l-1}], no bias) -> BatchNorm with batch statistics (biased
 * variance, also in eval) -> ReLU; then logits = a W_fc^T + b_fc.  The forward saves act / xhat / invstd per layer
 * for the backward, which writes or adds to every parameter gradient (accumulate bit 0: hidden layers add, bit 1:
 * fc_out adds) and optionally writes dx.
 * Limits: 2 <= B <= 64, widths and input <= 256, <= 8 hidden layers, <= 16 classes, and the working set (three
 * activation planes + the largest layer's weights) must fit 227 KB of shared memory: batch <= 48 for the default net. */
#define DNB_MLP_SMALL_MAX_WIDTH 256
#define DNB_MLP_SMALL_MAX_BATCH 64
#define DNB_MLP_SMALL_MAX_LAYERS 8
#define DNB_MLP_SMALL_MAX_CLASSES 16
typedef struct DnbMlpSmallDesc {
    const float* x;                                   /* [B, in0] fp32 */
    const void* W[DNB_MLP_SMALL_MAX_LAYERS];          /* bf16 [width_l, width_{l-1}] */
    const float* gamma[DNB_MLP_SMALL_MAX_LAYERS];
    const float* beta[DNB_MLP_SMALL_MAX_LAYERS];
    const void* Wfc; const float* bfc;                /* bf16 [C, width_last], fp32 [C] */
    float* act[DNB_MLP_SMALL_MAX_LAYERS];             /* [B, width_l] post-ReLU (saved) */
    float* xhat[DNB_MLP_SMALL_MAX_LAYERS];            /* [B, width_l] normalised pre-activations (saved) */
    float* invstd[DNB_MLP_SMALL_MAX_LAYERS];          /* [width_l] */
    float* logits;                                    /* [B, C] forward output */
    const float* dlogits;                             /* [B, C] backward input */
    float* dW[DNB_MLP_SMALL_MAX_LAYERS];
    float* dgamma[DNB_MLP_SMALL_MAX_LAYERS];
    float* dbeta[DNB_MLP_SMALL_MAX_LAYERS];
    float* dWfc; float* dbfc;
    float* dx;                                        /* optional [B, in0] */
    int width[DNB_MLP_SMALL_MAX_LAYERS];
    int B, in0, L, C, accumulate;
    float eps;
} DnbMlpSmallDesc;
int dnb_mlp_small_forward(const DnbMlpSmallDesc* desc, void* stream);
int dnb_mlp_small_backward(const DnbMlpSmallDesc* desc, void* stream);

/* ---- K3b: in-place column orthonormalisation (modified Gram-Schmidt, fp32) of `count` tall-skinny
 * matrices P[i] of shape [n[i], r[i]] with leading dimension ld[i]; one CTA per matrix, one launch per
 * 32 matrices.  powerSGD's orthogonalise(P) and the rankDAD subspace iteration (engine names
 * comps/__init__.py:15-16).  Pointer / shape arrays are HOST arrays. */
int dnb_orthogonalize_batch(float* const* P, const int* n, const int* r, const int* ld, int count, float eps,
                            void* stream);

/* ---- K2a / K2b: device-mode rankDAD (engine "rankDAD", reference comps/__init__.py:15; tunables
 * dad_reduction_rank / dad_num_pow_iters, compspec.json:236-238).  One launch per call, one thread-block
 * cluster per tapped Linear layer.
 *   mode 0: rank-r subspace iteration on (delta [N,out], act [N,in]) -> U [out,16], V [in,16] with
 *           delta^T act ~= U V^T (columns >= rank_cols are zero); the tail stores the factors and the dense
 *           gradients into slot `rank` of every site's factor buffer (P2P stores) and raises a flag per peer.
 *   mode 1: waits for all sites' flags, stacks the received factors ([nsplit*nranks*16, .] bf16 operands built
 *           by the kernel in the workspace passed as A/AT/D/DT), re-compresses the stack to rank r, and
 *           writes G = grad_scale * U V^T [out,in] (fp32) to layer.G; dense gradients are averaged into flat_grad.
 * Every big operand is needed in both orientations (bf16, leading dimensions multiples of 8).  The iterates
 * ZN/ZI/ZO/V0 are [32, ld] bf16 (16 hi rows + 16 lo rows); V0 is the seeded start matrix (identical on all sites).
 * Limits: out <= DNB_DAD_MAX_OUT per layer, <= DNB_DAD_MAX_LAYERS layers per call, rank <= 16.
 * dad_tol is not evaluated on the device: all `iters` iterations always run. */
#define DNB_DAD_R 16
#define DNB_DAD_MAX_LAYERS 8
#define DNB_DAD_MAX_OUT 1152
#define DNB_DAD_MAX_SPANS 32
#define DNB_DAD_MAX_PACK 8
typedef struct DnbDadLayerDesc {
    const void *A, *AT, *D, *DT;          /* bf16: act [N,in], act^T [in,N], delta [N,out], delta^T [out,N] */
    long long lda, ldat, ldd, lddt;
    int N, in, out, pad_;
    void *ZN, *ZI, *ZO;                   /* iterate workspaces [32, ldzn|ldzi|ldzo] bf16 */
    const void* V0;                       /* [32, ldzi] bf16 seeded start (hi rows 0-15, lo rows 16-31) */
    long long ldzn, ldzi, ldzo;
    float *W, *U, *V;                     /* fp32 [out,16] scratch, [out,16] left factor, [in,16] right factor */
    long long fact_off;                   /* float offset of this layer's (U,V) inside a site's slot */
    void *Uc, *Vc;                        /* mode 1: bf16 [out,64] / [in,64] hi|lo stacks (workspace) */
    float* G; long long ldg;              /* mode 1: gradient destination [out,in] */
} DnbDadLayerDesc;
typedef struct DnbDadDesc {
    DnbDadLayerDesc layer[DNB_DAD_MAX_LAYERS];
    float* fact[DNB_MAX_RANKS];           /* every site's factor buffer [2][nranks][slot_floats] fp32, peer-mapped */
    uint32_t* flags[DNB_MAX_RANKS];       /* every site's flag words [nranks], zero-initialised, peer-mapped */
    long long slot_floats, dense_off;     /* floats per site slot; offset of the dense gradients inside it */
    long long span_off[DNB_DAD_MAX_SPANS];/* dense gradients = spans of flat_grad (float offsets / lengths) */
    int span_len[DNB_DAD_MAX_SPANS];
    float* flat_grad;
    uint32_t* dev_state;                  /* device words {generation, CTA counter}; required with use_flags */
    float grad_scale;                     /* mode 1: 1 / nranks */
    int n_layers, mode, iters, cluster, nsplit, nranks, rank, use_flags, n_spans, rank_cols;
} DnbDadDesc;
int dnb_dad_step(const DnbDadDesc* desc, void* stream);

/* fp32 [N,d] (leading dimension lds) -> bf16 row-major [N,ldr] and/or its transpose [d,ldt]; up to
 * DNB_DAD_MAX_PACK matrices per launch (the classifier-head (delta, activation) pairs of the rankDAD step). */
typedef struct DnbDadPackDesc {
    const float* src[DNB_DAD_MAX_PACK];
    void* rows[DNB_DAD_MAX_PACK];
    void* colsT[DNB_DAD_MAX_PACK];
    long long lds[DNB_DAD_MAX_PACK], ldr[DNB_DAD_MAX_PACK], ldt[DNB_DAD_MAX_PACK];
    int N[DNB_DAD_MAX_PACK], d[DNB_DAD_MAX_PACK];
    int count;
} DnbDadPackDesc;
int dnb_dad_pack(const DnbDadPackDesc* desc, void* stream);

/* ---- K3a / K3c: the three streaming passes of a device-mode powerSGD step (engine "powerSGD", reference
 * comps/__init__.py:16; results nnlogs.ipynb:48-50) on one gradient matrix [n, m] (fp32, row-major, contiguous):
 *   pass 0: E += grad (E then holds M = grad + error feedback); P[n,r] = M * Q[m,r]
 *   pass 1: Q[m,r] += M^T * P[n,r]      (P orthonormalised; Q must be zero on entry: fp32 atomics over row blocks)
 *   pass 2: grad = P * Q^T (the rank-r approximation, formed on the fly); E = M - grad
 * r <= DNB_PSGD_MAX_RANK.  ~1 flop per byte: HBM-bound streaming kernels, exact fp32. */
#define DNB_PSGD_MAX_RANK 8
#define DNB_PSGD_MAX_SEGS 32
typedef struct DnbPsgdDesc {
    float* grad;          /* [n,m] in (pass 0) / out (pass 2) */
    float* E;             /* [n,m] error feedback, updated in place */
    float* Q;             /* [m,r] */
    float* P;             /* [n,r] */
    int n, m, r, pass;
} DnbPsgdDesc;
int dnb_psgd_pass(const DnbPsgdDesc* desc, void* stream);
/* dst[i][0..len[i]) = src[i][0..len[i]) for `count` fp32 segments (host arrays of device pointers), one launch per 32 */
int dnb_seg_copy(const float* const* src, float* const* dst, const int* len, int count, void* stream);
/* cudaMemsetAsync on `stream` (a memset node inside a captured graph, no kernel launch) */
int dnb_memset_async(void* p, int value, size_t bytes, void* stream);

/* K9: device-side metric accumulators (reference hooks: new_metrics().add / new_averages().add,
 * comps/fs/__init__.py:58-61, comps/icalstm/__init__.py:64-68).  Appends column `col` of scores_in [B, ld]
 * and the labels to scores/labs (capacity cap), updates state = {cursor, tp, fp, tn, fn} (uint64) at
 * threshold thr and lossacc = {sum(loss*B), sum(B)}.  loss may be NULL. */
int dnb_metrics_append(const float* scores_in, long long ld, int col, const long long* labels, const float* loss,
                       int B, float* scores, int* labs, int cap, unsigned long long* state, float* lossacc, float thr,
                       void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DINUNET_B200_H */
