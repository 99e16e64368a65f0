// Fused multi-head attention for the short sequences of the IQ path (n <= 128; 5/20/21 tokens at cfg3).
// One CTA per (batch, head): Q/K/V tiles live in shared memory, scores never touch HBM.
// replaces: models/transformer_layers.py:494-526 (split heads, q*scale, QK^T, masked_fill(-1e18), softmax,
//           dropout, PV, merge heads) and its autograd backward.
// Semantics kept from the reference: the mask fill is the FINITE value -1e18, so a fully masked row
// softmaxes to a uniform distribution, and masked positions receive no gradient.
#include "common.cuh"

namespace b200iq {

static constexpr int ATT_THREADS = 128;
static constexpr int ATT_MAXJ = 4;  // keys per lane -> nk <= 128

struct AttnArgs {
  const float *q, *k, *v;
  int ldq, ldk, ldv;
  const int* key_ids;  // [B, nk] or NULL
  int pad_id, causal;
  int B, nh, nq, nk, d;
  float scale, drop_p;
  uint64_t seed, offset;
  int rn;  // round outputs to TF32
};

__device__ __forceinline__ float keep_scale(const AttnArgs& a, size_t elem) {
  if (a.drop_p <= 0.f) return 1.f;
  Philox ph(a.seed);
  uint32_t r[4];
  uint64_t ctr = a.offset + (elem >> 2);
  ph((uint32_t)ctr, (uint32_t)(ctr >> 32), 1u, 0u, r);
  return (u32_to_unit(r[elem & 3]) > a.drop_p) ? 1.f / (1.f - a.drop_p) : 0.f;
}

// softmax row for query i; lane owns keys j = lane + 32*jj.  Returns p[] (pre-dropout) and the mask bits.
__device__ __forceinline__ void softmax_row(const AttnArgs& a, const float* Qs, const float* Ks, int dp, int b, int i,
                                            int lane, float (&p)[ATT_MAXJ], bool (&msk)[ATT_MAXJ]) {
  float mx = -3.0e38f;
#pragma unroll
  for (int jj = 0; jj < ATT_MAXJ; ++jj) {
    int j = lane + 32 * jj;
    p[jj] = -3.0e38f;
    msk[jj] = true;
    if (j < a.nk) {
      float s = 0.f;
      const float* qr = Qs + i * dp;
      const float* kr = Ks + j * dp;
      for (int c = 0; c < a.d; ++c) s = fmaf(qr[c], kr[c], s);
      s *= a.scale;
      bool m = (a.causal && j > i) || (a.key_ids && a.key_ids[b * a.nk + j] == a.pad_id);
      msk[jj] = m;
      p[jj] = m ? -1e18f : s;
      mx = fmaxf(mx, p[jj]);
    }
  }
  mx = warp_max(mx);
  float sum = 0.f;
#pragma unroll
  for (int jj = 0; jj < ATT_MAXJ; ++jj) {
    int j = lane + 32 * jj;
    p[jj] = (j < a.nk) ? __expf(p[jj] - mx) : 0.f;
    sum += p[jj];
  }
  sum = warp_sum(sum);
  float inv = 1.f / sum;
#pragma unroll
  for (int jj = 0; jj < ATT_MAXJ; ++jj) p[jj] *= inv;
}

__device__ __forceinline__ void load_tile(float* dst, const float* src, int rows, int d, int dp, int ld) {
  for (int idx = threadIdx.x; idx < rows * d; idx += blockDim.x) {
    int r = idx / d, c = idx - r * d;
    dst[r * dp + c] = src[(size_t)r * ld + c];
  }
}

__global__ void __launch_bounds__(ATT_THREADS)
attention_fwd_kernel(AttnArgs a, float* __restrict__ out, int ldo) {
  extern __shared__ float sm[];
  const int b = blockIdx.x / a.nh, h = blockIdx.x % a.nh, dp = a.d + 1;
  float* Qs = sm;
  float* Ks = Qs + a.nq * dp;
  float* Vs = Ks + a.nk * dp;
  float* Ps = Vs + a.nk * dp;  // [4 warps][nk]
  load_tile(Qs, a.q + (size_t)b * a.nq * a.ldq + h * a.d, a.nq, a.d, dp, a.ldq);
  load_tile(Ks, a.k + (size_t)b * a.nk * a.ldk + h * a.d, a.nk, a.d, dp, a.ldk);
  load_tile(Vs, a.v + (size_t)b * a.nk * a.ldv + h * a.d, a.nk, a.d, dp, a.ldv);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* Pw = Ps + warp * a.nk;
  for (int i = warp; i < a.nq; i += ATT_THREADS / 32) {
    float p[ATT_MAXJ]; bool msk[ATT_MAXJ];
    softmax_row(a, Qs, Ks, dp, b, i, lane, p, msk);
#pragma unroll
    for (int jj = 0; jj < ATT_MAXJ; ++jj) {
      int j = lane + 32 * jj;
      if (j < a.nk) Pw[j] = p[jj] * keep_scale(a, ((size_t)blockIdx.x * a.nq + i) * a.nk + j);
    }
    __syncwarp();
    float* orow = out + ((size_t)b * a.nq + i) * ldo + h * a.d;
    for (int c = lane; c < a.d; c += 32) {
      float o = 0.f;
      for (int j = 0; j < a.nk; ++j) o = fmaf(Pw[j], Vs[j * dp + c], o);
      orow[c] = rnd(o, a.rn);
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(ATT_THREADS)
attention_bwd_kernel(AttnArgs a, const float* __restrict__ dout, int ldo, float* __restrict__ dq, float* __restrict__ dk,
                     float* __restrict__ dv, int lddq, int lddk, int lddv) {
  extern __shared__ float sm[];
  const int b = blockIdx.x / a.nh, h = blockIdx.x % a.nh, dp = a.d + 1, np = a.nk + 1;
  float* Qs = sm;
  float* Ks = Qs + a.nq * dp;
  float* Vs = Ks + a.nk * dp;
  float* dOs = Vs + a.nk * dp;
  float* Ws = dOs + a.nq * dp;   // post-dropout weights [nq][np]
  float* dSs = Ws + a.nq * np;   // d(logits) * scale     [nq][np]
  load_tile(Qs, a.q + (size_t)b * a.nq * a.ldq + h * a.d, a.nq, a.d, dp, a.ldq);
  load_tile(Ks, a.k + (size_t)b * a.nk * a.ldk + h * a.d, a.nk, a.d, dp, a.ldk);
  load_tile(Vs, a.v + (size_t)b * a.nk * a.ldv + h * a.d, a.nk, a.d, dp, a.ldv);
  load_tile(dOs, dout + (size_t)b * a.nq * ldo + h * a.d, a.nq, a.d, dp, ldo);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = warp; i < a.nq; i += ATT_THREADS / 32) {
    float p[ATT_MAXJ]; bool msk[ATT_MAXJ];
    softmax_row(a, Qs, Ks, dp, b, i, lane, p, msk);
    float dP[ATT_MAXJ], dot = 0.f;
#pragma unroll
    for (int jj = 0; jj < ATT_MAXJ; ++jj) {
      int j = lane + 32 * jj;
      dP[jj] = 0.f;
      if (j < a.nk) {
        float ks = keep_scale(a, ((size_t)blockIdx.x * a.nq + i) * a.nk + j);
        float g = 0.f;
        const float* dor = dOs + i * dp;
        const float* vr = Vs + j * dp;
        for (int c = 0; c < a.d; ++c) g = fmaf(dor[c], vr[c], g);
        Ws[i * np + j] = p[jj] * ks;
        dP[jj] = g * ks;
        dot += dP[jj] * p[jj];
      }
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int jj = 0; jj < ATT_MAXJ; ++jj) {
      int j = lane + 32 * jj;
      if (j < a.nk) dSs[i * np + j] = msk[jj] ? 0.f : p[jj] * (dP[jj] - dot) * a.scale;
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < a.nq * a.d; idx += blockDim.x) {
    int i = idx / a.d, c = idx - i * a.d;
    float s = 0.f;
    for (int j = 0; j < a.nk; ++j) s = fmaf(dSs[i * np + j], Ks[j * dp + c], s);
    dq[((size_t)b * a.nq + i) * lddq + h * a.d + c] = rnd(s, a.rn);
  }
  for (int idx = threadIdx.x; idx < a.nk * a.d; idx += blockDim.x) {
    int j = idx / a.d, c = idx - j * a.d;
    float sk = 0.f, sv = 0.f;
    for (int i = 0; i < a.nq; ++i) {
      sk = fmaf(dSs[i * np + j], Qs[i * dp + c], sk);
      sv = fmaf(Ws[i * np + j], dOs[i * dp + c], sv);
    }
    dk[((size_t)b * a.nk + j) * lddk + h * a.d + c] = rnd(sk, a.rn);
    dv[((size_t)b * a.nk + j) * lddv + h * a.d + c] = rnd(sv, a.rn);
  }
}

static int check_attn(const B200iqAttention* p) {
  B200IQ_CHECK_ARG(p && p->q && p->k && p->v, "attention: null pointer");
  B200IQ_CHECK_ARG(p->B > 0 && p->num_heads > 0 && p->nq > 0 && p->nk > 0 && p->head_dim > 0, "attention: bad shape");
  B200IQ_CHECK_ARG(p->nk <= 32 * ATT_MAXJ, "attention: nk=%d > %d unsupported", p->nk, 32 * ATT_MAXJ);
  B200IQ_CHECK_ARG(p->drop_p >= 0.f && p->drop_p < 1.f, "attention: bad dropout p");
  return 0;
}
static AttnArgs to_args(const B200iqAttention* p) {
  AttnArgs a;
  a.q = p->q; a.k = p->k; a.v = p->v; a.ldq = p->ldq; a.ldk = p->ldk; a.ldv = p->ldv;
  a.key_ids = p->key_ids; a.pad_id = p->pad_id; a.causal = p->causal;
  a.B = p->B; a.nh = p->num_heads; a.nq = p->nq; a.nk = p->nk; a.d = p->head_dim;
  a.scale = p->scale; a.drop_p = p->drop_p; a.seed = p->seed; a.offset = p->offset;
  a.rn = round_mode();
  return a;
}

}  // namespace b200iq
using namespace b200iq;

extern "C" int b200iq_attention_fwd(const B200iqAttention* p, void* stream) {
  if (int rc = check_attn(p)) return rc;
  B200IQ_CHECK_ARG(p->out, "attention_fwd: null out");
  AttnArgs a = to_args(p);
  size_t smem = sizeof(float) * ((size_t)(a.nq + 2 * a.nk) * (a.d + 1) + (ATT_THREADS / 32) * a.nk);
  B200IQ_CHECK_ARG(smem <= 227 * 1024, "attention_fwd: tile does not fit shared memory (%zu B)", smem);
  if (smem > 48 * 1024) cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  attention_fwd_kernel<<<a.B * a.nh, ATT_THREADS, smem, (cudaStream_t)stream>>>(a, p->out, p->ldo);
  return launch_status("attention_fwd");
}

extern "C" int b200iq_attention_bwd(const B200iqAttention* p, const float* dout, float* dq, float* dk, float* dv,
                                    int lddq, int lddk, int lddv, void* stream) {
  if (int rc = check_attn(p)) return rc;
  B200IQ_CHECK_ARG(dout && dq && dk && dv, "attention_bwd: null pointer");
  AttnArgs a = to_args(p);
  size_t smem = sizeof(float) * ((size_t)(2 * a.nq + 2 * a.nk) * (a.d + 1) + 2 * (size_t)a.nq * (a.nk + 1));
  B200IQ_CHECK_ARG(smem <= 227 * 1024, "attention_bwd: tile does not fit shared memory (%zu B)", smem);
  if (smem > 48 * 1024) cudaFuncSetAttribute(attention_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  attention_bwd_kernel<<<a.B * a.nh, ATT_THREADS, smem, (cudaStream_t)stream>>>(a, dout, p->ldo, dq, dk, dv, lddq, lddk, lddv);
  return launch_status("attention_bwd");
}
