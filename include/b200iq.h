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

#define B200IQ_ABI_VERSION 1

int b200iq_abi_version(void);
/* 1 when the current device is compute capability 10.0, else 0 */
int b200iq_device_supported(void);
const char* b200iq_last_error(void);

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

#ifdef __cplusplus
}
#endif
#endif /* B200IQ_H_ */
