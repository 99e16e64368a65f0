/* libb200iq — C ABI of the B200-native IQ / BLT-VQG training-step kernels.
 *
 * Every entry point replaces a stock-PyTorch op sequence of the reference
 * (nihirv/blt-vqg); the `replaces:` note on each declaration cites it as
 * file:line relative to the reference root.  Conventions:
 *   - extern "C", plain pointers and sizes, no torch / C++ types;
 *   - all pointers are DEVICE pointers owned by the caller unless stated;
 *   - every launch is asynchronous on the `stream` argument (a cudaStream_t
 *     passed as void*); nothing allocates, frees or retains pointers;
 *   - return 0 = ok, <0 = argument error (nothing launched), >0 = cudaError_t;
 *     b200iq_last_error() gives the thread-local message;
 *   - there is NO CPU fallback and no other architecture: the library only
 *     contains sm_100a code.
 * Floating-point tensors are fp32 row-major; token ids are int32.
 */
#ifndef B200IQ_H_
#define B200IQ_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200IQ_ABI_VERSION 3

int b200iq_abi_version(void);
/* 1 when the current device is compute capability 10.0, else 0 */
int b200iq_device_supported(void);
const char* b200iq_last_error(void);
/* Numerics mode of the whole library (default 1).  1: the tensor-core path — every activation / gradient tensor a
 * kernel writes is rounded to the nearest TF32 value, because tcgen05.mma kind::tf32 truncates its fp32 operands and
 * truncation biases every dot product towards zero.  0: nothing is rounded (use with B200IQ_GEMM_SIMT for an exact
 * fp32 run, e.g. the bit-exact greedy decode).  Returns the previous mode. */
int b200iq_set_tf32_rounding(int on);
/* Programmatic dependent launch for the hot kernels (GEMM, LayerNorm, column sums, attention, cross entropy, axpby):
 * 1 lets each of them become resident under the tail of the previous kernel of its stream and block until that kernel
 * has completed (griddepcontrol.wait) — same results, no launch gap.  Default 0.  Returns the previous setting. */
int b200iq_set_pdl(int on);

/* ------------------------------------------------------------------ GEMM */
/* flags for B200iqGemm.flags */
#define B200IQ_GEMM_A_MN      1   /* A stored [K,M] row-major (transposed operand) */
#define B200IQ_GEMM_B_MN      2   /* B stored [K,N] row-major; default is [N,K] (nn.Linear weight) */
#define B200IQ_GEMM_BIAS      4   /* + bias[N] */
#define B200IQ_GEMM_RELU      8   /* relu after bias/residual */
#define B200IQ_GEMM_RESIDUAL  16  /* + residual[M,N] (ld = ldr) */
#define B200IQ_GEMM_ACCUM     32  /* C += result (atomic; required when split_k > 1) */
#define B200IQ_GEMM_SIMT      64  /* exact-fp32 CUDA-core path instead of tcgen05 TF32 */
#define B200IQ_GEMM_RELU_MASK 128 /* multiply result by (gate[M,N] > 0): relu backward fused */
#define B200IQ_GEMM_FP32_OUT  256 /* keep the full fp32 result (logits, returned tensors): no TF32 rounding of C */
#define B200IQ_GEMM_DROPOUT   512 /* after relu: C *= keep/(1-p), keep = 16 bits of Philox4x32-7(seed; counter = offset +
                                     (row*N+col)/8) >= round(65536 p) — the mask b200iq_dropout makes on the contiguous [M,N]
                                     tensor (needs ldc == N, N % 4 == 0);
                                     replaces the nn.Dropout after each FFN linear, transformer_layers.py:405-406 */

typedef struct B200iqGemm {
  const float* A;        /* [M,K] ld=lda, or [K,M] when A_MN */
  const float* B;        /* [N,K] ld=ldb, or [K,N] when B_MN */
  float* C;              /* [M,N] ld=ldc */
  const float* bias;     /* [N] or NULL */
  const float* residual; /* [M,N] ld=ldr, or NULL; with RELU_MASK: the gate tensor */
  int M, N, K;
  int lda, ldb, ldc, ldr;
  int flags;
  int split_k;           /* >=1; >1 needs ACCUM and forbids RELU */
  float alpha;           /* scales A·B before the epilogue */
  float drop_p;          /* DROPOUT: drop probability in [0,1) */
  unsigned long long drop_seed, drop_offset; /* DROPOUT: Philox key / counter base */
} B200iqGemm;

/* C = epilogue(alpha * op(A) · op(B)^T).
 * replaces: every nn.Linear / torch.matmul on the path, e.g.
 *   models/transformer_layers.py:489-491,530 (Q/K/V/O), :403 (FFN),
 *   models/iq.py:76 (emb projection), models/decoder_transformer.py:35,40,
 *   models/mlp.py:37, models/encoder_cnn.py:20, models/decoder_rnn.py:107,
 *   and their autograd backward (dX = dY·W, dW = dYᵀ·X).
 * Default path: TMA -> shared memory (128B swizzle) -> tcgen05.mma kind::tf32 ->
 * TMEM accumulator -> tcgen05.ld epilogue.  Needs 16-byte aligned bases and
 * leading dimensions that are multiples of 4 floats. */
int b200iq_gemm(const B200iqGemm* p, void* stream);
/* tile policy of the tcgen05 path: 0 = automatic (default), 128 / 256 = one CTA per 128 x tile output tile,
 * 512 = CTA pairs (tcgen05.mma.cta_group::2, 256 x 256 per pair) wherever M, N > 128.  Returns the previous setting.
 * Environment default: B200IQ_GEMM_TILE. */
int b200iq_set_gemm_tile(int tile);

/* ------------------------------------------------------------------ fp16-plane GEMM (the transformer path's precision scheme)
 * A GEMM operand x is stored as fp16 PLANES written by the kernel that produces x: hi = fp16(x) and, for operands of forward
 * products, lo = fp16(x - hi) (together ~22 significand bits).  With both lo planes given the product is the 3-term
 * A_hi.B_hi + A_lo.B_hi + A_hi.B_lo (K-major operands only); without them it is the single pass A_hi.B_hi.  Why: the forward
 * products feed ReLU gates, and 11-bit operands there flip enough gates to put ~1e-2 of error on the gradients, while the
 * backward products (dX = dY.W, dW = dY^T.X) tolerate 11 bits (profiles/r02a_cpu_precision_study.txt).  Backward operands carry
 * the loss scale of the caller (fp16 range).  Output: fp32 C (TMA store; ACCUM / split-K as in b200iq_gemm) and / or fp16 planes
 * C_hi (+ C_lo) for the next GEMM.  flags: A_MN, B_MN, BIAS, RELU, RESIDUAL (fp32 tile, needs C), ACCUM, RELU_MASK (gate = fp16
 * plane, > 0 passes), DROPOUT.  Planes: 16-byte aligned, leading dimensions multiples of 8 halves.
 * replaces: the same nn.Linear call sites as b200iq_gemm (models/transformer_layers.py:489-491,530,403; models/iq.py:76;
 * models/decoder_transformer.py:35,40; models/mlp.py:37; models/encoder_cnn.py:20) and their autograd backward. */
typedef struct B200iqGemmF16 {
  const void* A_hi; const void* A_lo;   /* [M,K] ld=lda halves ([K,M] when A_MN); A_lo NULL = single pass */
  const void* B_hi; const void* B_lo;   /* [N,K] ld=ldb halves ([K,N] when B_MN) */
  float* C;                             /* [M,N] ld=ldc floats, or NULL */
  void* C_hi; void* C_lo;               /* fp16 planes [M,N] ld=ldh halves, or NULL */
  const float* bias;                    /* [N] or NULL */
  const float* residual;                /* [M,N] ld=ldr floats (RESIDUAL) */
  const void* gate;                     /* fp16 plane [M,N] ld=ldg halves (RELU_MASK) */
  int M, N, K;
  int lda, ldb, ldc, ldh, ldr, ldg;
  int flags;
  int split_k;
  float alpha;
  float drop_p;
  unsigned long long drop_seed, drop_offset;
  const unsigned long long* drop_seed_dev; /* NULL, or a DEVICE word added to drop_seed when the kernel runs: a step captured in a
                                              CUDA graph keeps its launch arguments, so the per-step part of the Philox key lives
                                              in device memory and is bumped by the graph itself */
} B200iqGemmF16;
int b200iq_gemm_f16(const B200iqGemmF16* p, void* stream);
/* hi[r,c] = fp16(scale * src[r,c]), lo[r,c] = fp16(scale * src[r,c] - hi[r,c]) (lo may be NULL): the plane copy of the fp32
 * master weights (rows = 1, cols = arena size) and of small fp32 tensors that feed a GEMM.  lds / ldd: row pitches (floats / halves). */
int b200iq_split_f16(const float* src, void* hi, void* lo, int rows, long long cols, long long lds, long long ldd, float scale, void* stream);
/* b200iq_embed_gather writing planes (ldo halves).   replaces: nn.Embedding forward, models/iq.py:71-78 */
int b200iq_embed_gather_f16(const int* ids, const float* table, void* out_hi, void* out_lo, int M, int E, int V, int ldo, void* stream);
/* LayerNorm forward with plane outputs: sum = x + add when add != NULL (sum may be NULL otherwise); y (fp32), y_hi, y_lo are
 * each optional outputs of the same normalised row (ldy halves for the planes). */
int b200iq_layernorm_fwd_f16(const float* x, const float* add, float* sum, const float* gamma, const float* beta, float* y,
                             void* y_hi, void* y_lo, int ldy, float* mean, float* rstd, int M, int H, float eps, void* stream);
/* b200iq_layernorm_bwd that also writes fp16(dx) (the accumulated result) to dx_hi (ld H halves) when dx_hi != NULL */
int b200iq_layernorm_bwd_f16(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                             float* dx, void* dx_hi, float* dgamma, float* dbeta, int M, int H, int accumulate_dx, void* stream);
/* dpre_hi = fp16(scale * dy * (y > 0)); dbias[c] += colsum (dbias may be NULL) */
int b200iq_relu_bwd_colsum_f16(const float* dy, const float* y, void* dpre_hi, float* dbias, int M, int N, float scale, void* stream);
/* out[c] += sum_m x_hi[m,c] over an fp16 plane (bias gradients of tensors that only exist as planes) */
int b200iq_colsum_add_f16(const void* x_hi, float* out, int M, int N, int ld, void* stream);
/* b200iq_ce_rows that leaves the logits untouched and writes the gradient rows as an fp16 plane (ldg halves) */
int b200iq_ce_rows_f16(const float* logits, int ld, const int* targets, int rows, int V, int pad_id, const float* count,
                       float coef, float* loss_sum, void* grad_hi, int ldg, void* stream);

/* ---------------------------------------------------------- row / elementwise ops */
/* out[m,:] = table[ids[m],:]   replaces: nn.Embedding fwd, models/iq.py:71-78 (also decoder_rnn.py:102, encoder_rnn.py:78) */
int b200iq_embed_gather(const int* ids, const float* table, float* out, int M, int E, int V, void* stream);
/* dtable[ids[m],:] += dout[m,:] unless ids[m]==pad_id (pad_id<0: no padding row)   replaces: Embedding backward */
int b200iq_embed_scatter_add(const int* ids, const float* dout, float* dtable, int M, int E, int V, int pad_id, void* stream);
/* x[b,t,:] += timing[t,:]; token 0 also gets add0[b,:] + add1[b,:] (either may be NULL)
 * replaces: models/transformer_layers.py:146,214; models/decoder_transformer.py:31-34,46 */
int b200iq_add_timing(float* x, const float* timing, const float* add0, const float* add1, int B, int n, int H, void* stream);
/* dst[b*ldd + c] (+)= src[(b*n)*H + c]  /  dst[(b*n)*H + c] += src[b*lds + c]  for c < H: moves token 0 of every sequence
 * (x[:,0] reads of models/encoder_transformer.py:32,35, iq.py:110 and the matching gradient plumbing) */
int b200iq_tok0_gather(const float* src, float* dst, int B, int n, int H, int ldd, int accumulate, void* stream);
int b200iq_tok0_scatter_add(const float* src, float* dst, int B, int n, int H, int lds, void* stream);
/* torch.nn.LayerNorm(H) forward (saves mean/rstd per row) and backward (dgamma/dbeta accumulate)
 * replaces: models/transformer_layers.py:134,151,202,220,256-257,320-322 */
int b200iq_layernorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* mean, float* rstd,
                         int M, int H, float eps, void* stream);
/* sum = x + add (rounded like b200iq_axpby), y = LayerNorm(sum): the residual add in front of a LayerNorm and the LayerNorm in
 * one pass (H in {128, 256, 512, 1024}, 16-byte aligned pointers; other shapes: b200iq_axpby + b200iq_layernorm_fwd).
 * replaces: `x + y` at models/transformer_layers.py:282,364 followed by the next layer's layer_norm (:271,338) / the final one */
int b200iq_add_layernorm_fwd(const float* x, const float* add, float* sum, const float* gamma, const float* beta, float* y,
                             float* mean, float* rstd, int M, int H, float eps, void* stream);
int b200iq_layernorm_bwd(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                         float* dx, float* dgamma, float* dbeta, int M, int H, int accumulate_dx, void* stream);
/* out[c] += sum_m x[m,c]  (bias gradients) */
int b200iq_colsum_add(const float* x, float* out, int M, int N, int ld, void* stream);
/* dpre = dy * (y > 0); dbias[c] += colsum(dpre) (dbias may be NULL)   replaces: ReLU backward, transformer_layers.py:405 */
int b200iq_relu_bwd_colsum(const float* dy, const float* y, float* dpre, float* dbias, int M, int N, float scale, void* stream);
/* inverted dropout in place; element i is kept when 16 bits of Philox4x32-7(seed; counter = offset + i/8) — word i & 3,
 * upper half when (i >> 2) & 1 — are >= round(65536 p).  Re-applying it to the gradient with the same key is the backward.   replaces: nn.Dropout at transformer_layers.py:406,520; base_rnn.py:41 */
int b200iq_dropout(float* x, size_t n, float p, uint64_t seed, uint64_t offset, void* stream);
/* same with the device word *seed_dev (may be NULL) added to the seed at run time (steps replayed from a CUDA graph) */
int b200iq_dropout_dev(float* x, size_t n, float p, uint64_t seed, uint64_t offset, const unsigned long long* seed_dev, void* stream);
/* out ~ N(0,1), Philox4x32-10(seed; counter = offset + i/4) + Box-Muller (moments checked in tests/test_ops_gpu.py)
 * replaces: torch.randn on the CPU + H2D copy, models/transformer_layers.py:45 */
int b200iq_randn(float* out, size_t n, uint64_t seed, uint64_t offset, void* stream);
/* same with the device word *seed_dev (may be NULL) added to the seed at run time (CUDA-graph replays draw fresh numbers) */
int b200iq_randn_dev(float* out, size_t n, uint64_t seed, uint64_t offset, const unsigned long long* seed_dev, void* stream);
/* out = a*x + b*y (out may alias x or y): residual adds that cannot ride a GEMM epilogue */
int b200iq_axpby(const float* x, const float* y, float* out, size_t n, float a, float b, void* stream);
/* dst = round_to_nearest_tf32(src): the GEMM-operand copy of the fp32 master weights / of user inputs */
int b200iq_round_tf32(const float* src, float* dst, size_t n, void* stream);

/* ---------------------------------------------------------- fused attention (n <= 128) */
typedef struct B200iqAttention {
  const float* q;        /* [B*nq, ldq], head h at columns [h*head_dim, (h+1)*head_dim) */
  const float* k;        /* [B*nk, ldk] */
  const float* v;        /* [B*nk, ldv] */
  float* out;            /* [B*nq, ldo] merged heads (fwd only) */
  int ldq, ldk, ldv, ldo;
  const int* key_ids;    /* [B, nk] token ids; keys equal to pad_id are masked; NULL = no padding mask */
  int pad_id;
  int causal;            /* 1: key j > query i is masked (transformer_layers.py:207) */
  int B, num_heads, nq, nk, head_dim;
  float scale;           /* head_dim^-0.5 (transformer_layers.py:449,499) */
  float drop_p;          /* attention dropout (transformer_layers.py:520); key j of score row r = (b*heads + h)*nq + i is kept when 16 bits
                            of Philox4x32-7(seed; counter = offset + r*ceil(nk/8) + j/8, 1) — word (j >> 1) & 3, upper half when
                            j & 1 — are >= round(65536 p) */
  uint64_t seed, offset;
  long long kv_batch_stride; /* elements between consecutive samples' K (and V) blocks; 0 = nk*ldk / nk*ldv (dense [B*nk, ld]).
                                A KV cache [B, max_len, H] read at length nk passes max_len*H here (forward only). */
  int precise;           /* 1 (head_dim 64, n <= 80): q / k / v are unrounded fp32; the forward's QK^T and PV run as 3-term
                            products (hi.hi + lo.hi + hi.lo, ~2^-21: fp16 mma.sync with ldmatrix fragments for n <= 32, TF32 terms
                            split in registers above that), the backward recomputes the scores the same way and rounds (instead of
                            truncating) the operands of its single-pass products.  0: legacy TF32 single pass. */
  void* out_hi;          /* precise forward: fp16 planes of the merged-head output (ldo halves) written instead of `out` when */
  void* out_lo;          /*   out_hi != NULL (out_lo optional) */
  const unsigned long long* seed_dev; /* NULL, or a device word added to `seed` at run time (see B200iqGemmF16.drop_seed_dev) */
  /* planes_in = 1 (precise mode, head_dim 64, nq, nk <= 32): q / k / v are the fp16 HI planes of the projections (what
   * b200iq_gemm_f16 writes through C_hi / C_lo), q_lo / k_lo / v_lo their lo planes, ldq / ldk / ldv row pitches in halves
   * (multiples of 8, 16-byte aligned bases).  The kernels then stage the operands without a conversion pass; the backward reads
   * v as hi only (v_lo may be NULL there) and takes `dout` of b200iq_attention_bwd_f16 as an fp16 hi plane with pitch ldo halves. */
  int planes_in;
  const void* q_lo; const void* k_lo; const void* v_lo;
} B200iqAttention;
/* replaces: models/transformer_layers.py:494-526; mask fill is the finite -1e18 of :506 */
int b200iq_attention_fwd(const B200iqAttention* p, void* stream);
/* recomputes the probabilities from q/k (nothing is stashed by fwd); writes dq/dk/dv (not accumulated) */
int b200iq_attention_bwd(const B200iqAttention* p, const float* dout, float* dq, float* dk, float* dv,
                         int lddq, int lddk, int lddv, void* stream);
/* same, gradients written as fp16 planes (ld in halves): the operands of the projections' backward GEMMs */
int b200iq_attention_bwd_f16(const B200iqAttention* p, const float* dout, void* dq_hi, void* dk_hi, void* dv_hi,
                             int lddq, int lddk, int lddv, void* stream);

/* ---------------------------------------------------------- losses */
/* count[0] = #(targets != pad_id)  (the CE "mean over non-pad" denominator, train_iq.py:82-83) */
int b200iq_count_nonpad(const int* targets, int n, int pad_id, float* count, void* stream);
/* Softmax cross-entropy over `rows` logits rows, forward AND gradient in one pass: loss_sum += sum_r (lse_r - logit_r[t_r]) / count;
 * with write_grad the rows are overwritten by coef * (softmax - onehot) / count (zero rows where t_r == pad_id).  The caller feeds
 * row CHUNKS of the vocabulary projection, so the (B*T, V) logits tensor never exists.
 * replaces: models/decoder_transformer.py:40 + train_iq.py:82-83; models/decoder_rnn.py:107 + old_train_iq.py:122-126 */
int b200iq_ce_rows(float* logits, int ld, const int* targets, int rows, int V, int pad_id, const float* count,
                   float coef, float* loss_sum, int write_grad, void* stream);
/* CE(repeat(z_logit, T), targets) without materialising the repeat.   replaces: train_iq.py:92-94 */
int b200iq_ce_aux(const float* z_logit, float* dz_logit, const int* targets, int B, int T, int V, int pad_id,
                  const float* count, float coef, float* loss_sum, void* stream);
/* prior/post are [B,2L] = [mu | logvar]; z = eps*exp(.5*logvar_q)+mu_q (prior when post==NULL); kld_sum += mean_b KL(q||p)
 * replaces: models/transformer_layers.py:43-58,536-540 */
int b200iq_latent_fwd(const float* prior, const float* post, const float* eps, float* z, float* kld_sum, int B, int L, void* stream);
int b200iq_latent_bwd(const float* prior, const float* post, const float* eps, const float* dz, float kld_coef,
                      float* dprior, float* dpost, int B, int L, void* stream);
/* same with the coefficient read from device memory (kld_coef_dev[0]): the KL weight anneals every step (train_iq.py:96), a
 * captured step cannot carry it as a launch argument */
int b200iq_latent_bwd_dev(const float* prior, const float* post, const float* eps, const float* dz, const float* kld_coef_dev,
                          float* dprior, float* dpost, int B, int L, void* stream);
/* nn.BatchNorm1d(H, momentum) on [B,H]; training: batch statistics + running-stat update.   replaces: models/encoder_cnn.py:21,34 */
int b200iq_bn_fwd(const float* x, const float* gamma, const float* beta, float* y, float* save_mean, float* save_rstd,
                  float* running_mean, float* running_var, int B, int H, float eps, float momentum, int training, void* stream);
int b200iq_bn_bwd(const float* dy, const float* x, const float* gamma, const float* save_mean, const float* save_rstd,
                  float* dx, float* dgamma, float* dbeta, int B, int H, void* stream);
/* loss_sum += mean((a-b)^2); da += coef*2(a-b)/n; db (+)= -coef*2(a-b)/n (da/db/loss_sum may be NULL)   replaces: train_iq.py:84 */
int b200iq_mse(const float* a, const float* b, size_t n, float coef, float* loss_sum, float* da, float* db, int accumulate_db, void* stream);

/* ---------------------------------------------------------- decode-time helpers */
/* x[r,:] -= logsumexp(x[r,:])   replaces: F.log_softmax at models/decoder_rnn.py:107 */
int b200iq_log_softmax_rows(float* x, int ld, int rows, int V, void* stream);
/* per row the k (1 or 6) largest logits, ties to the lowest index; val_out (or NULL) = their softmax probabilities
 * replaces: torch.max + torch.topk(softmax, 6) at models/iq.py:136-137; topk(1) at models/decoder_rnn.py:124 */
int b200iq_topk_softmax_rows(const float* x, int ld, int rows, int V, int k, int* idx_out, float* val_out, void* stream);

/* ---------------------------------------------------------- LSTM recurrence (time-major, gate order i|f|g|o) */
/* One step of nn.LSTM: `gates` [B,4H] holds the pre-activations h_{t-1}.W_hh^T + x_t.W_ih^T + b (written by b200iq_gemm with
 * the residual epilogue) and is overwritten with the activated gates (kept for the backward pass); c_next = f*c_prev + i*g,
 * h_next = o*tanh(c_next).  lengths (or NULL): rows with t >= lengths[b] copy their state (packed-sequence semantics).
 * replaces: models/decoder_rnn.py:105 / models/encoder_rnn.py:80-86 (cuDNN LSTM forward) */
int b200iq_lstm_cell_fwd(float* gates, const float* c_prev, const float* h_prev, float* c_next, float* h_next,
                         const int* lengths, int t, int B, int H, void* stream);
/* Backward of one step: dgates [B,4H] = pre-activation gradients; dc updated in place to d/dc_{t-1}; dh_run (d/dh_t from the
 * future) is left holding only the pass-through of masked rows so that  dh_run = dgates.W_hh + dh_run  (b200iq_gemm) is
 * d/dh_{t-1}.  dh_out: gradient from the consumers of h_t, or NULL. */
int b200iq_lstm_cell_bwd(const float* gates, const float* c_prev, const float* c_t, float* dh_run, const float* dh_out,
                         float* dc, float* dgates, const int* lengths, int t, int B, int H, void* stream);
/* Persistent recurrence: all T steps of nn.LSTM(H, H, 1) in one launch (8-CTA clusters, W_hh slice resident in shared
 * memory, mma.sync TF32, c in registers, one hardware cluster barrier per step).  Same buffers as the per-step pair
 * b200iq_gemm + b200iq_lstm_cell_fwd: gates [T,B,4H] holds the input projection (+ both biases) on entry and the ACTIVATED
 * gates on return; hbuf / cbuf [T+1,B,H] with slot 0 = the initial state; lengths [B] or NULL (packed-sequence masking,
 * models/encoder_rnn.py:80-86).   replaces: models/decoder_rnn.py:105, models/encoder_rnn.py:83 (cuDNN recurrence). */
int b200iq_lstm_seq_supported(int H);
int b200iq_lstm_seq_fwd(float* gates, const float* whh, float* hbuf, float* cbuf, const int* lengths, int T, int B, int H, void* stream);
/* BPTT of the same recurrence in one launch: gates (activated, from the forward) and cbuf in; dh_last [B,H] (gradient of the
 * final hidden state) and d_out [T,B,H] (gradient of every output) may each be NULL; writes dgates [T,B,4H] (pre-activation
 * gate gradients: the operand of the dW_hh / dW_ih / dx GEMMs), dh0 and dc0 [B,H].  workspace: caller-owned scratch of
 * b200iq_lstm_seq_bwd_workspace_bytes(B, H) bytes (partial products exchanged between the CTAs of a cluster). */
size_t b200iq_lstm_seq_bwd_workspace_bytes(int B, int H);
int b200iq_lstm_seq_bwd(const float* gates, const float* cbuf, const float* whh, const float* dh_last, const float* d_out,
                        float* dgates, float* dh0, float* dc0, float* workspace, const int* lengths, int T, int B, int H, void* stream);
/* b200iq_lstm_seq_bwd for callers that cannot allocate device memory (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc): same kernel, same arguments minus `workspace` — the scratch is a file-scope device array inside the library
 * (B200IQ_LSTM_SEQ_BWD_STATIC_MAX_BYTES = 5 MiB per GPU that loads the library, held until exit; covers B <= 256 at any supported
 * H, e.g. the B = 256, H = 300 full-size LSTM config).  Calls no cudaMalloc / cudaFree / allocating routine.  Returns < 0 with a
 * message when b200iq_lstm_seq_bwd_workspace_bytes(B, H) exceeds the limit.  Every call on a device shares the array: calls that
 * OVERLAP on the GPU (different streams) race; back-to-back calls are fine (nothing in the array has to start at zero).
 * No counterpart in the reference. */
#define B200IQ_LSTM_SEQ_BWD_STATIC_MAX_BYTES (5u << 20)
size_t b200iq_lstm_seq_bwd_static_max_bytes(void);
int b200iq_lstm_seq_bwd_static(const float* gates, const float* cbuf, const float* whh, const float* dh_last, const float* d_out,
                               float* dgates, float* dh0, float* dc0, const int* lengths, int T, int B, int H, void* stream);
/* out[t*B+b] = in[b*T_in + col0 + t], t < T_out: batch-major token ids -> time-major (inputs[:, :-1] of decoder_rnn.py:141) */
int b200iq_transpose_ids(const int* in, int* out, int B, int T_in, int T_out, int col0, void* stream);

/* ---------------------------------------------------------- input side: feature cache + token block */
/* The trunk is frozen (models/encoder_cnn.py:18-19) and an image repeats across its questions (utils/store_dataset.py:107-121), so
 * its pooled features are computed once and kept in HBM: `cache` is [num_images, F], bf16 (cache_bf16 = 1, F % 8 == 0) or fp32
 * (cache_bf16 = 0, F % 4 == 0) rows, 16-byte aligned.  A batch then names its images by row index (int32 idx[B]).
 * gather: out[b,:] = cache[idx[b],:] widened to fp32 (fp32 rows are rounded to TF32 like every GEMM operand unless
 * b200iq_set_tf32_rounding(0)); store: cache[idx[b],:] = feats[b,:] (round-to-nearest-even to bf16).  Rows whose index is outside
 * [0, num_images) read as zeros / are not written, and are counted into *bad_rows (device int, may be NULL).
 * replaces: the per-step `self.encoder_cnn.cnn(images)` trunk pass + avgpool of models/encoder_cnn.py:32-33 and the (B,3,224,224)
 * fp32 host->device copy of the batch dict (utils/data_loader.py:160,168) */
int b200iq_feature_gather(const void* cache, int cache_bf16, const int* idx, float* out, int B, int F, int num_images,
                          int* bad_rows, void* stream);
int b200iq_feature_store(const float* feats, const int* idx, void* cache, int cache_bf16, int B, int F, int num_images,
                         int* bad_rows, void* stream);
/* dst[i] = (int32) src[i]: the reference's collate_fn builds int64 token tensors (`.long()`, utils/data_loader.py:162-166);
 * the kernels read int32 */
int b200iq_ids_narrow(const int64_t* src, int* dst, size_t n, void* stream);

/* ---------------------------------------------------------- optimizer */
/* hyper[0..3] = {lr, 1-beta1^t, 1-beta2^t, grad_scale}, written by a kernel whose ARGUMENTS carry the host scalars (no pinned
 * staging buffer to race with).   replaces: the param_groups[0]["lr"] write of train_iq.py:257 */
int b200iq_set_hyper(float* hyper, float lr, float bias_corr1, float bias_corr2, float grad_scale, void* stream);
/* out[0] = sum(grad^2) over the flat gradient arena, bit-reproducible (fixed reduction order; no float atomics), so the clip
 * factor is identical on every data-parallel rank.  scratch: >= 1186 floats of caller-owned device memory, ZERO-initialised once by the caller and then
 * reused across calls (the kernel leaves its ticket word at 0) */
int b200iq_grad_sumsq(const float* grad, size_t n, float* out, float* scratch, int scratch_floats, void* stream);
/* clip_grad_norm_(max_norm) + Adam on flat arenas; hyper (device) = {lr, 1-beta1^t, 1-beta2^t, grad_scale}.
 * replaces: train_iq.py:259-261 (Adam), :372 (gradient_clip_val=5), :252-257 (LR written by the host into hyper[0]) */
int b200iq_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, size_t n, const float* hyper,
                     const float* normsq, float beta1, float beta2, float eps, float max_norm, int zero_grad, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200IQ_H_ */
