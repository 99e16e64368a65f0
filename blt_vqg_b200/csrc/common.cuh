// Shared device/host helpers for libb200iq (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/b200iq.h"

namespace b200iq {

// thread-local error string returned by b200iq_last_error()
void set_error(const char* fmt, ...);

#define B200IQ_CHECK_ARG(cond, ...)                     \
  do {                                                  \
    if (!(cond)) {                                      \
      ::b200iq::set_error(__VA_ARGS__);                 \
      return -1;                                        \
    }                                                   \
  } while (0)

// returns positive cudaError_t on launch failure
static inline int launch_status(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
int sm_count();
// 1: every activation / gradient tensor a kernel writes is rounded to the nearest TF32 value (b200iq_set_tf32_rounding)
int round_mode();

// Programmatic dependent launch (b200iq_set_pdl): a kernel launched through launch_dep() may become resident while the
// previous kernel of its stream is still draining, runs its prologue, and blocks in pdl_wait() until that kernel has
// completed and its writes are visible.  Every kernel launched this way calls pdl_wait() before its first global access
// (reads AND writes) and pdl_trigger() at its top; both are no-ops for a plain launch.
int pdl_mode();
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <class... KA, class... A>
static inline cudaError_t launch_dep(void (*kernel)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_mode() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KA(args)...);
}

// ---------------------------------------------------------------- device utils
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Round-to-nearest (ties away) fp32 -> tf32, kept in an fp32 container.  tcgen05.mma kind::tf32 TRUNCATES its fp32
// operands; truncation shrinks every product towards zero (a systematic ~7e-4 relative bias per GEMM), so every
// tensor that can become a GEMM operand is rounded once, by the kernel that produces it.
__device__ __forceinline__ float tf32_rn(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float rnd(float x, int on) { return on ? tf32_rn(x) : x; }

// block-wide sum; `red` must hold >= 32 floats of shared memory. All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* red) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  r = warp_sum(r);
  return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : -3.0e38f;
  r = warp_max(r);
  return r;
}

// Philox4x32-10 counter-based RNG (Salmon et al. 2011); same constants as curand/torch.
struct Philox {
  uint32_t k0, k1;
  __host__ __device__ Philox(uint64_t seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
  __host__ __device__ static inline void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
    uint64_t p = (uint64_t)a * b;
    hi = (uint32_t)(p >> 32);
    lo = (uint32_t)p;
  }
  // counter = (c0,c1,c2,c3) -> 4 random words
  __host__ __device__ inline void operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                             uint32_t out[4]) const {
    rounds<10>(c0, c1, c2, c3, out);
  }
  template <int ROUNDS>
  __host__ __device__ inline void rounds(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t out[4]) const {
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
      uint32_t h0, l0, h1, l1;
      mulhilo(0xD2511F53u, c0, h0, l0);
      mulhilo(0xCD9E8D57u, c2, h1, l1);
      uint32_t n0 = h1 ^ c1 ^ a, n1 = l1, n2 = h0 ^ c3 ^ b, n3 = l0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      a += 0x9E3779B9u; b += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
  }
};
__host__ __device__ inline float u32_to_unit(uint32_t x) {  // (0,1]
  return ((float)(x >> 8) + 1.0f) * (1.0f / 16777216.0f);
}

// The feed-forward dropout mask (b200iq_dropout and the GEMM epilogue make the SAME one): element i of the contiguous tensor
// takes 16 bits of Philox4x32-7(seed; counter = offset + i/8) — word i & 3, upper half when (i >> 2) & 1 — and is kept when
// they are >= round(p * 65536).  One call covers 8 elements and 7 rounds is the depth Salmon et al. report as passing
// BigCrush: 2.9x less integer work per element than a 10-round word per element, which the GEMM epilogue could not hide
// (profiles/r01f_gemm_dropout_epilogue_icache.txt).
__host__ __device__ inline uint32_t dropout_threshold(float p) { return (uint32_t)(p * 65536.0f + 0.5f); }
__device__ __forceinline__ void dropout_words(uint64_t seed, uint64_t offset, size_t i, uint32_t w[4]) {
  const uint64_t ctr = offset + (i >> 3);
  Philox(seed).rounds<7>((uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u, w);
}
__device__ __forceinline__ bool dropout_keep(const uint32_t w[4], size_t i, uint32_t thr) {
  return ((w[i & 3] >> (((i >> 2) & 1) << 4)) & 0xFFFFu) >= thr;
}

}  // namespace b200iq
