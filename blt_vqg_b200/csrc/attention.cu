// Fused multi-head attention for the short sequences of the IQ path (n <= 128; 5/20/21 tokens at cfg3).
// One CTA per (batch, head): Q/K/V tiles live in shared memory, scores never touch HBM.
// replaces: models/transformer_layers.py:494-526 (split heads, q*scale, QK^T, masked_fill(-1e18), softmax,
//           dropout, PV, merge heads) and its autograd backward.
// Semantics kept from the reference: the mask fill is the FINITE value -1e18, so a fully masked row
// softmaxes to a uniform distribution, and masked positions receive no gradient.
#include <stdlib.h>
#include <type_traits>
#include <cuda_fp16.h>
#include "common.cuh"

namespace b200iq {

// where a kernel's result rows go: fp32 (f32 != NULL) or fp16 planes hi (+ lo) for the next GEMM (gemm_f16.cu); ld in elements
struct AttOut { float* f32; __half* hi; __half* lo; int ld; };
__device__ __forceinline__ AttOut att_out_at(const AttOut& o, size_t off) {
  AttOut r = o;
  if (r.f32) r.f32 += off;
  if (r.hi) r.hi += off;
  if (r.lo) r.lo += off;
  return r;
}
__device__ __forceinline__ void att_store1(const AttOut& o, size_t idx, float v, int rn) {
  if (o.hi) {
    const __half h = __float2half_rn(v);
    o.hi[idx] = h;
    if (o.lo) o.lo[idx] = __float2half_rn(v - __half2float(h));
  } else {
    o.f32[idx] = rnd(v, rn);
  }
}
// 8 consecutive elements of one row (idx a multiple of 8 for planes / of 4 for fp32)
__device__ __forceinline__ void att_store8(const AttOut& o, size_t idx, const float (&v)[8], int rn) {
  if (o.hi) {
    uint32_t hb[4], lb[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __half2 h = __floats2half2_rn(v[2 * j], v[2 * j + 1]);
      hb[j] = *reinterpret_cast<const uint32_t*>(&h);
      const float2 f = __half22float2(h);
      const __half2 l = __floats2half2_rn(v[2 * j] - f.x, v[2 * j + 1] - f.y);
      lb[j] = *reinterpret_cast<const uint32_t*>(&l);
    }
    *reinterpret_cast<uint4*>(o.hi + idx) = make_uint4(hb[0], hb[1], hb[2], hb[3]);
    if (o.lo) *reinterpret_cast<uint4*>(o.lo + idx) = make_uint4(lb[0], lb[1], lb[2], lb[3]);
  } else {
    float4* dst = reinterpret_cast<float4*>(o.f32 + idx);
    dst[0] = make_float4(rnd(v[0], rn), rnd(v[1], rn), rnd(v[2], rn), rnd(v[3], rn));
    dst[1] = make_float4(rnd(v[4], rn), rnd(v[5], rn), rnd(v[6], rn), rnd(v[7], rn));
  }
}

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
  const unsigned long long* seed_dev;  // device word added to seed at run time (CUDA-graph replays), or NULL
  int rn;  // round outputs to TF32
  int dq2; // backward: dQ = (dS + dS_lo) . K, dS as two TF32 terms (B200IQ_ATT_DQ2, default 1)
  size_t bsk, bsv;  // elements between consecutive samples' K / V blocks (nk*ldk, nk*ldv unless reading a KV cache)
  const __half *ql, *kl, *vl;  // planes_in: q / k / v above are fp16 hi planes (cast), these their lo planes; ld* in halves
};

// The attention-dropout mask: the 8 keys j = 8c .. 8c+7 of score row `row` (= (sample*heads + head)*nq + i) share ONE call of
// Philox4x32-7(seed; counter = offset + row*ceil(nk/8) + c, 1): key j takes 16 bits — word (j >> 1) & 3, upper half when j & 1 —
// and is kept when they are >= round(65536 p) (the feed-forward mask's recipe, common.cuh).  A lane of the tensor-core kernels
// owns keys 2t, 2t+1 of every 8-key tile of its rows: one call per (row, tile) instead of a 10-round call per element, which
// was most of the instruction stream of these kernels (5.5 k instructions per warp, ncu profiles/r01k_attention_fwd_rows.txt).
__device__ __forceinline__ void att_drop_words(const AttnArgs& a, size_t row, int tile, uint32_t (&w)[4]) {
  const uint64_t ctr = a.offset + row * (uint64_t)((a.nk + 7) >> 3) + (uint64_t)tile;
  Philox(a.seed + ((a.drop_p > 0.f && a.seed_dev) ? __ldg(a.seed_dev) : 0ull)).rounds<7>((uint32_t)ctr, (uint32_t)(ctr >> 32), 1u, 0u, w);
}
__device__ __forceinline__ float att_keep(const uint32_t (&w)[4], int j, uint32_t thr, float inv) {
  return (((w[(j >> 1) & 3] >> ((j & 1) << 4)) & 0xFFFFu) >= thr) ? inv : 0.f;
}
__device__ __forceinline__ float keep_scale(const AttnArgs& a, size_t row, int j) {
  if (a.drop_p <= 0.f) return 1.f;
  uint32_t w[4];
  att_drop_words(a, row, j >> 3, w);
  return att_keep(w, j, dropout_threshold(a.drop_p), 1.f / (1.f - a.drop_p));
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
attention_fwd_kernel(AttnArgs a, AttOut out) {
  const int ldo = out.ld;
  extern __shared__ float sm[];
  const int b = blockIdx.x / a.nh, h = blockIdx.x % a.nh, dp = a.d + 1;
  float* Qs = sm;
  float* Ks = Qs + a.nq * dp;
  float* Vs = Ks + a.nk * dp;
  float* Ps = Vs + a.nk * dp;  // [4 warps][nk]
  load_tile(Qs, a.q + (size_t)b * a.nq * a.ldq + h * a.d, a.nq, a.d, dp, a.ldq);
  load_tile(Ks, a.k + (size_t)b * a.bsk + h * a.d, a.nk, a.d, dp, a.ldk);
  load_tile(Vs, a.v + (size_t)b * a.bsv + h * a.d, a.nk, a.d, dp, a.ldv);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* Pw = Ps + warp * a.nk;
  for (int i = warp; i < a.nq; i += ATT_THREADS / 32) {
    float p[ATT_MAXJ]; bool msk[ATT_MAXJ];
    softmax_row(a, Qs, Ks, dp, b, i, lane, p, msk);
#pragma unroll
    for (int jj = 0; jj < ATT_MAXJ; ++jj) {
      int j = lane + 32 * jj;
      if (j < a.nk) Pw[j] = p[jj] * keep_scale(a, (size_t)blockIdx.x * a.nq + i, j);
    }
    __syncwarp();
    const size_t orow = ((size_t)b * a.nq + i) * ldo + h * a.d;
    for (int c = lane; c < a.d; c += 32) {
      float o = 0.f;
      for (int j = 0; j < a.nk; ++j) o = fmaf(Pw[j], Vs[j * dp + c], o);
      att_store1(out, orow + c, o, a.rn);
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(ATT_THREADS)
attention_bwd_kernel(AttnArgs a, const float* __restrict__ dout, int ldo, AttOut dq, AttOut dk, AttOut dv) {
  const int lddq = dq.ld, lddk = dk.ld, lddv = dv.ld;
  extern __shared__ float sm[];
  const int b = blockIdx.x / a.nh, h = blockIdx.x % a.nh, dp = a.d + 1, np = a.nk + 1;
  float* Qs = sm;
  float* Ks = Qs + a.nq * dp;
  float* Vs = Ks + a.nk * dp;
  float* dOs = Vs + a.nk * dp;
  float* Ws = dOs + a.nq * dp;   // post-dropout weights [nq][np]
  float* dSs = Ws + a.nq * np;   // d(logits) * scale     [nq][np]
  load_tile(Qs, a.q + (size_t)b * a.nq * a.ldq + h * a.d, a.nq, a.d, dp, a.ldq);
  load_tile(Ks, a.k + (size_t)b * a.bsk + h * a.d, a.nk, a.d, dp, a.ldk);
  load_tile(Vs, a.v + (size_t)b * a.bsv + h * a.d, a.nk, a.d, dp, a.ldv);
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
        float ks = keep_scale(a, (size_t)blockIdx.x * a.nq + i, j);
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
    att_store1(dq, ((size_t)b * a.nq + i) * lddq + h * a.d + c, s, a.rn);
  }
  for (int idx = threadIdx.x; idx < a.nk * a.d; idx += blockDim.x) {
    int j = idx / a.d, c = idx - j * a.d;
    float sk = 0.f, sv = 0.f;
    for (int i = 0; i < a.nq; ++i) {
      sk = fmaf(dSs[i * np + j], Qs[i * dp + c], sk);
      sv = fmaf(Ws[i * np + j], dOs[i * dp + c], sv);
    }
    att_store1(dk, ((size_t)b * a.nk + j) * lddk + h * a.d + c, sk, a.rn);
    att_store1(dv, ((size_t)b * a.nk + j) * lddv + h * a.d + c, sv, a.rn);
  }
}

// ------------------------------------------------------------------ tensor-core path (head_dim 64, nq,nk <= 32)
// The SIMT kernels above are issue-bound (ncu profiles/r01c_attention_*: 69 % issue-active at 7 % of HBM peak: one LDS
// pair per FMA).  Here ONE WARP owns one (sample, head) and keeps everything in registers: QK^T and PV are
// mma.sync.m16n8k8 TF32 with operands loaded straight from global memory as float4 — the reduction index of an MMA may
// be permuted freely as long as A and B agree, and so may the n index as long as the store agrees, which is what makes
// 16-byte loads and 32-byte-per-lane stores line up with the fragment layouts.  No shared memory in the forward.
//   A frag: a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4);  B frag: b0 (k=t, n=g) b1 (k=t+4, n=g);
//   C frag: c0 (g, 2t) c1 (g, 2t+1) c2 (g+8, 2t) c3 (g+8, 2t+1)          with g = lane>>2, t = lane&3
__device__ __forceinline__ void mma_tf32(float (&c)[4], float a0, float a1, float a2, float a3, float b0, float b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(__float_as_uint(a0)), "r"(__float_as_uint(a1)), "r"(__float_as_uint(a2)), "r"(__float_as_uint(a3)),
                 "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}
__device__ __forceinline__ float4 ld4(const float* base, int row, int nrows, int ld, int col) {
  if (row < nrows) return *reinterpret_cast<const float4*>(base + (size_t)row * ld + col);   // global or shared (generic load)
  return make_float4(0.f, 0.f, 0.f, 0.f);
}

// The operands arrive as unrounded fp32 and mma.sync TRUNCATES to TF32, so every operand is split in registers:
// hi = round-to-nearest TF32, lo = x - hi (exact; its own truncation is a 2^-22 effect).
__device__ __forceinline__ float4 hi4(float4 v) { return make_float4(tf32_rn(v.x), tf32_rn(v.y), tf32_rn(v.z), tf32_rn(v.w)); }
__device__ __forceinline__ float4 sub4(float4 a, float4 b) { return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }

// acc[mt][nt] += X[16mt.., :64] . Y[8nt.., :64]^T   (both K-major in global memory; rows beyond nx / ny read as zero)
// TERMS = 1: hi.hi (operands rounded to nearest);  TERMS = 3: hi.hi + lo.hi + hi.lo (~2^-21: the forward products)
template <int MT, int NT, int TERMS>
__device__ __forceinline__ void mma_xyT(float (&acc)[MT][NT][4], const float* X, int nx, int ldx, const float* Y, int ny, int ldy,
                                        int g, int t) {
  // rolled: the body (all tiles, three terms) is large and these kernels were instruction-fetch bound when every k-step was
  // unrolled (ncu r02i: stall_no_instruction the top stall of the backward kernel)
#pragma unroll 1
  for (int kk = 0; kk < 4; ++kk) {
    float4 xa[MT], xb[MT], yf[NT];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) { xa[mt] = ld4(X, 16 * mt + g, nx, ldx, 16 * kk + 4 * t); xb[mt] = ld4(X, 16 * mt + g + 8, nx, ldx, 16 * kk + 4 * t); }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) yf[nt] = ld4(Y, 8 * nt + g, ny, ldy, 16 * kk + 4 * t);
    float4 xal[MT], xbl[MT], yl[NT];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      const float4 ah = hi4(xa[mt]), bh = hi4(xb[mt]);
      if (TERMS == 3) { xal[mt] = sub4(xa[mt], ah); xbl[mt] = sub4(xb[mt], bh); }
      xa[mt] = ah; xb[mt] = bh;
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const float4 yh = hi4(yf[nt]);
      if (TERMS == 3) yl[nt] = sub4(yf[nt], yh);
      yf[nt] = yh;
    }
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        if (TERMS == 3) {
          mma_tf32(acc[mt][nt], xal[mt].x, xbl[mt].x, xal[mt].y, xbl[mt].y, yf[nt].x, yf[nt].y);
          mma_tf32(acc[mt][nt], xal[mt].z, xbl[mt].z, xal[mt].w, xbl[mt].w, yf[nt].z, yf[nt].w);
          mma_tf32(acc[mt][nt], xa[mt].x, xb[mt].x, xa[mt].y, xb[mt].y, yl[nt].x, yl[nt].y);
          mma_tf32(acc[mt][nt], xa[mt].z, xb[mt].z, xa[mt].w, xb[mt].w, yl[nt].z, yl[nt].w);
        }
        mma_tf32(acc[mt][nt], xa[mt].x, xb[mt].x, xa[mt].y, xb[mt].y, yf[nt].x, yf[nt].y);
        mma_tf32(acc[mt][nt], xa[mt].z, xb[mt].z, xa[mt].w, xb[mt].w, yf[nt].z, yf[nt].w);
      }
  }
}

// out[16mt.., :64] = P . Y with P[mt][nt] in C-fragment layout (rows i, cols j) and Y[j, :64] in global memory (rows >= ny zero).
// Lane (g,t) ends up with 8 contiguous output floats per row and 32-column block: cols 32cq + 8t .. +7.
// With PLO != nullptr the left operand is P + PLO (two TF32 terms: see the dQ product of the backward kernels).
// Y is rounded to nearest here; P (and PLO) must hold TF32 values.  TERMS = 3 adds the P.Y_lo term (with P = hi, PLO = lo of the
// attention weights this is the forward's 3-term P.V).
template <int MT, int NT, int TERMS>
__device__ __forceinline__ void mma_pY_store(const float (&P)[MT][NT][4], const float* Y, int ny, int ldy, const AttOut& out, int nrows,
                                             int g, int t, int rn, const float (*PLO)[NT][4] = nullptr) {
  const int ldo = out.ld;
#pragma unroll 1
  for (int cq = 0; cq < 2; ++cq) {
    float acc[MT][4][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mt][u][e] = 0.f;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const float4 y0f = ld4(Y, 8 * nt + 2 * t, ny, ldy, 32 * cq + 4 * g), y1f = ld4(Y, 8 * nt + 2 * t + 1, ny, ldy, 32 * cq + 4 * g);
      const float4 y0 = hi4(y0f), y1 = hi4(y1f);
      const float4 y0l = sub4(y0f, y0), y1l = sub4(y1f, y1);
      const float* y0a = &y0.x; const float* y1a = &y1.x;
      const float* y0b = &y0l.x; const float* y1b = &y1l.x;
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (TERMS == 3) mma_tf32(acc[mt][u], P[mt][nt][0], P[mt][nt][2], P[mt][nt][1], P[mt][nt][3], y0b[u], y1b[u]);
          if (PLO != nullptr) mma_tf32(acc[mt][u], PLO[mt][nt][0], PLO[mt][nt][2], PLO[mt][nt][1], PLO[mt][nt][3], y0a[u], y1a[u]);
          mma_tf32(acc[mt][u], P[mt][nt][0], P[mt][nt][2], P[mt][nt][1], P[mt][nt][3], y0a[u], y1a[u]);
        }
    }
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int row = 16 * mt + g + 8 * hh;
        if (row < nrows) {
          const float v[8] = {acc[mt][0][2 * hh], acc[mt][1][2 * hh], acc[mt][2][2 * hh], acc[mt][3][2 * hh],
                              acc[mt][0][2 * hh + 1], acc[mt][1][2 * hh + 1], acc[mt][2][2 * hh + 1], acc[mt][3][2 * hh + 1]};
          att_store8(out, (size_t)row * ldo + 32 * cq + 8 * t, v, rn);
        }
      }
  }
}

// softmax over j of the masked, scaled scores in C-fragment layout; S becomes P (pre-dropout). msk bit (nt*4+e) per mt.
// i0 = absolute index of the fragment's first query row (a warp of the row-block kernels owns rows i0 .. i0+15).
template <int MT, int NT, class MaskT = uint32_t>
__device__ __forceinline__ void softmax_frag(const AttnArgs& a, float (&S)[MT][NT][4], MaskT (&msk)[MT], int b, int g, int t, int i0 = 0) {
  bool padj[NT][2];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int j = 8 * nt + 2 * t + e;
      padj[nt][e] = (a.key_ids != nullptr) && j < a.nk && __ldg(a.key_ids + b * a.nk + j) == a.pad_id;
    }
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
    msk[mt] = 0;
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int i = i0 + 16 * mt + g + 8 * hh;
      float mx = -3.0e38f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = 8 * nt + 2 * t + e;
          const bool m = (a.causal && j > i) || padj[nt][e];
          if (m) msk[mt] |= (MaskT)1 << (nt * 4 + 2 * hh + e);
          float s = m ? -1e18f : S[mt][nt][2 * hh + e] * a.scale;
          if (j >= a.nk) s = -3.0e38f;
          S[mt][nt][2 * hh + e] = s;
          mx = fmaxf(mx, s);
        }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      float sum = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = 8 * nt + 2 * t + e;
          const float p = (j < a.nk) ? __expf(S[mt][nt][2 * hh + e] - mx) : 0.f;
          S[mt][nt][2 * hh + e] = p;
          sum += p;
        }
      sum += __shfl_xor_sync(0xffffffffu, sum, 1);
      sum += __shfl_xor_sync(0xffffffffu, sum, 2);
      const float inv = 1.f / sum;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) S[mt][nt][2 * hh + e] *= inv;
    }
  }
}

static constexpr int ATT_TC_WARPS = 2;
// The register-resident kernels first pull the (sample, head) tiles they multiply — Q, K, V (and dO) — into shared memory with
// cp.async (16-byte chunks, every chunk of a warp's tiles in flight at once) and read the MMA fragments from there: with the
// fragments loaded straight from global memory every k-step waited for its own L2 round trip (ncu: long-scoreboard stalls, 8 warps
// per SM, 49 % / 36 % of the HBM roof on the cfg3 shapes).  Row pitch 76 floats: the P.V-style fragment reads (rows 2t, columns 4g)
// are bank-conflict-free, the Q.K^T-style reads (rows g, columns 4t) 2-way on 4 of 16 banks.
static constexpr int ATT_SP = 76;
__device__ __forceinline__ void att_stage(float* dst, const float* __restrict__ src, int rows, int ld, int lane) {
  for (int idx = lane; idx < rows * 16; idx += 32) {
    const int r = idx >> 4, c4 = idx & 15;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                 ::"r"((uint32_t)__cvta_generic_to_shared(dst + r * ATT_SP + 4 * c4)), "l"(src + (size_t)r * ld + 4 * c4) : "memory");
  }
}
__device__ __forceinline__ void att_stage_wait() {
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncwarp();
}

template <int MT, int NT>
__global__ void __launch_bounds__(ATT_TC_WARPS * 32)
attention_fwd_tc_kernel(AttnArgs a, AttOut out) {
  const int ldo = out.ld;
  pdl_trigger();
  pdl_wait();
  const int bh = blockIdx.x * ATT_TC_WARPS + (threadIdx.x >> 5);
  if (bh >= a.B * a.nh) return;
  const int b = bh / a.nh, h = bh - b * a.nh, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  extern __shared__ __align__(16) float att_dyn[];
  float* Q = att_dyn + (size_t)(threadIdx.x >> 5) * (a.nq + 2 * a.nk) * ATT_SP;
  float* K = Q + a.nq * ATT_SP;
  float* V = K + a.nk * ATT_SP;
  att_stage(Q, a.q + (size_t)b * a.nq * a.ldq + h * 64, a.nq, a.ldq, lane);
  att_stage(K, a.k + (size_t)b * a.bsk + h * 64, a.nk, a.ldk, lane);
  att_stage(V, a.v + (size_t)b * a.bsv + h * 64, a.nk, a.ldv, lane);
  float S[MT][NT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) S[mt][nt][e] = 0.f;
  att_stage_wait();
  mma_xyT<MT, NT, 3>(S, Q, a.nq, ATT_SP, K, a.nk, ATT_SP, g, t);
  uint32_t msk[MT];
  softmax_frag<MT, NT>(a, S, msk, b, g, t);
  float SL[MT][NT][4];   // low-order TF32 term of the attention weights (3-term P.V)
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int i = 16 * mt + g + 8 * hh;
        uint32_t dw[4];
        if (a.drop_p > 0.f) att_drop_words(a, (size_t)bh * a.nq + i, nt, dw);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = 8 * nt + 2 * t + e;
          float w = S[mt][nt][2 * hh + e];
          if (a.drop_p > 0.f && i < a.nq && j < a.nk) w *= att_keep(dw, j, dropout_threshold(a.drop_p), 1.f / (1.f - a.drop_p));
          const float wh = tf32_rn(w);
          S[mt][nt][2 * hh + e] = wh;
          SL[mt][nt][2 * hh + e] = w - wh;
        }
      }
  mma_pY_store<MT, NT, 3>(S, V, a.nk, ATT_SP, att_out_at(out, (size_t)b * a.nq * ldo + h * 64), a.nq, g, t, a.rn, SL);
}

// out[j, :64] = sum_i T[i][j] * X[i, :64] for j < nj, with T ([32][33] floats, rows i, written by this warp in C layout)
// used as the TRANSPOSED A operand.  Reduction slot t of k-step s is i = 4*MT*t + 2s (slot t+4: i+1), which makes the
// shared-memory reads of a fragment hit 32 distinct banks for MT = 2.
template <int MT, int NT>
__device__ __forceinline__ void mma_TtX_store(const float* T, const float* X, int nx, int ldx, const AttOut& out, int nj,
                                              int g, int t, int rn) {
  const int ldo = out.ld;
  constexpr int MTJ = (NT + 1) / 2;
#pragma unroll 1
  for (int cq = 0; cq < 2; ++cq) {
    float acc[MTJ][4][4];
#pragma unroll
    for (int mj = 0; mj < MTJ; ++mj)
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mj][u][e] = 0.f;
#pragma unroll
    for (int s = 0; s < 2 * MT; ++s) {
      const int i = 4 * MT * t + 2 * s;
      const float4 x0 = hi4(ld4(X, i, nx, ldx, 32 * cq + 4 * g)), x1 = hi4(ld4(X, i + 1, nx, ldx, 32 * cq + 4 * g));
      const float* x0a = &x0.x; const float* x1a = &x1.x;
#pragma unroll
      for (int mj = 0; mj < MTJ; ++mj) {
        const int j0 = 16 * mj + g, j1 = j0 + 8;
        const float a0 = T[i * 33 + j0], a2 = T[(i + 1) * 33 + j0];
        const float a1 = (j1 < 8 * NT) ? T[i * 33 + j1] : 0.f, a3 = (j1 < 8 * NT) ? T[(i + 1) * 33 + j1] : 0.f;
#pragma unroll
        for (int u = 0; u < 4; ++u) mma_tf32(acc[mj][u], a0, a1, a2, a3, x0a[u], x1a[u]);
      }
    }
#pragma unroll
    for (int mj = 0; mj < MTJ; ++mj)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int row = 16 * mj + g + 8 * hh;
        if (row < nj) {
          const float v[8] = {acc[mj][0][2 * hh], acc[mj][1][2 * hh], acc[mj][2][2 * hh], acc[mj][3][2 * hh],
                              acc[mj][0][2 * hh + 1], acc[mj][1][2 * hh + 1], acc[mj][2][2 * hh + 1], acc[mj][3][2 * hh + 1]};
          att_store8(out, (size_t)row * ldo + 32 * cq + 8 * t, v, rn);
        }
      }
  }
}

template <int MT, int NT>
__global__ void __launch_bounds__(ATT_TC_WARPS * 32)
attention_bwd_tc_kernel(AttnArgs a, const float* __restrict__ dout, int ldo, AttOut dq, AttOut dk, AttOut dv) {
  const int lddq = dq.ld, lddk = dk.ld, lddv = dv.ld;
  __shared__ float tsm[ATT_TC_WARPS][2][32 * 33];
  pdl_trigger();
  pdl_wait();
  const int wid = threadIdx.x >> 5;
  const int bh = blockIdx.x * ATT_TC_WARPS + wid;
  if (bh >= a.B * a.nh) return;
  const int b = bh / a.nh, h = bh - b * a.nh, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  extern __shared__ __align__(16) float att_dyn[];
  float* Q = att_dyn + (size_t)wid * (2 * a.nq + 2 * a.nk) * ATT_SP;
  float* K = Q + a.nq * ATT_SP;
  float* V = K + a.nk * ATT_SP;
  float* dO = V + a.nk * ATT_SP;
  att_stage(Q, a.q + (size_t)b * a.nq * a.ldq + h * 64, a.nq, a.ldq, lane);
  att_stage(K, a.k + (size_t)b * a.bsk + h * 64, a.nk, a.ldk, lane);
  att_stage(V, a.v + (size_t)b * a.bsv + h * 64, a.nk, a.ldv, lane);
  att_stage(dO, dout + (size_t)b * a.nq * ldo + h * 64, a.nq, ldo, lane);
  float S[MT][NT][4], dW[MT][NT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) { S[mt][nt][e] = 0.f; dW[mt][nt][e] = 0.f; }
  att_stage_wait();
  mma_xyT<MT, NT, 3>(S, Q, a.nq, ATT_SP, K, a.nk, ATT_SP, g, t);   // scores: the forward's 3-term product, so P is the forward's P
  mma_xyT<MT, NT, 1>(dW, dO, a.nq, ATT_SP, V, a.nk, ATT_SP, g, t);   // d(post-dropout weights) = dO . V^T
  uint32_t msk[MT];
  softmax_frag<MT, NT>(a, S, msk, b, g, t);
  float* Ws = tsm[wid][0];
  float* Ds = tsm[wid][1];
  // S: P -> dS (tf32) ; dW -> W (tf32); both also go to shared memory as the transposed operands of dK / dV
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int i = 16 * mt + g + 8 * hh;
      float ks[NT][2], dot = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        uint32_t dw[4];
        if (a.drop_p > 0.f) att_drop_words(a, (size_t)bh * a.nq + i, nt, dw);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = 8 * nt + 2 * t + e;
          ks[nt][e] = (a.drop_p > 0.f && i < a.nq && j < a.nk) ? att_keep(dw, j, dropout_threshold(a.drop_p), 1.f / (1.f - a.drop_p)) : 1.f;
          const float dP = dW[mt][nt][2 * hh + e] * ks[nt][e];
          dW[mt][nt][2 * hh + e] = dP;
          dot += dP * S[mt][nt][2 * hh + e];
        }
      }
      dot += __shfl_xor_sync(0xffffffffu, dot, 1);
      dot += __shfl_xor_sync(0xffffffffu, dot, 2);
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = 8 * nt + 2 * t + e, idx = 2 * hh + e;
          const float p = S[mt][nt][idx];
          const bool m = (msk[mt] >> (nt * 4 + idx)) & 1u;
          const float dsf = (m || i >= a.nq) ? 0.f : p * (dW[mt][nt][idx] - dot) * a.scale;
          const float ds = tf32_rn(dsf);
          const float w = (i >= a.nq) ? 0.f : tf32_rn(p * ks[nt][e]);
          S[mt][nt][idx] = ds;
          dW[mt][nt][idx] = tf32_rn(dsf - ds);   // low-order term of dS, used by the dQ product only
          Ws[i * 33 + j] = w;
          Ds[i * 33 + j] = ds;
        }
    }
  __syncwarp();
  // dQ = dS . K with dS as TWO TF32 terms: every row of dS sums to zero, so when the keys share a large common component (LayerNorm
  // + a bias-free projection at initialisation) the product cancels it and the 2^-11 rounding of a single-term dS is what is left
  // (measured 7e-3 of |dQ|, scripts/gpu_attention_err.py; 3e-4 with the second term)
  mma_pY_store<MT, NT, 1>(S, K, a.nk, ATT_SP, att_out_at(dq, (size_t)b * a.nq * lddq + h * 64), a.nq, g, t, a.rn, a.dq2 ? dW : nullptr);
  mma_TtX_store<MT, NT>(Ds, Q, a.nq, ATT_SP, att_out_at(dk, (size_t)b * a.nk * lddk + h * 64), a.nk, g, t, a.rn);  // dK = dS^T . Q
  mma_TtX_store<MT, NT>(Ws, dO, a.nq, ATT_SP, att_out_at(dv, (size_t)b * a.nk * lddv + h * 64), a.nk, g, t, a.rn);   // dV = W^T . dO
}

// ---------------------------------------------------------------- fp16 mma.sync backward (precise mode, fp16-plane gradients)
// attention_bwd_tc_kernel above executes ~5 k SASS instructions per (sample, head) — m16n8k8 TF32 MMAs whose fragments are
// assembled from scalar / 16-byte shared loads and split into TF32 terms at every use, 12 Philox calls per lane — at 6 warps per
// SM, and sits at 12-20 % of the HBM roof although its DRAM traffic is the algorithmic minimum (ncu profiles/r02i_attention_bwd.txt).
// This kernel does the same five products with mma.sync.m16n8k16 on fp16 tiles: Q, K (hi + lo planes, for the 3-term score
// product that reproduces the forward's probabilities), V and dO (hi) are converted ONCE while they are staged, fragments come
// from ldmatrix (.trans for the operands that are read transposed: dS^T, W^T, and K / Q / dO as row-major B operands), dS goes
// from the accumulator layout of the score product straight into the A operand of dQ = dS . K (two fp16 terms: rows of dS sum
// to zero, see attention_bwd_tc_kernel), and the 4 lanes that share a score row compute each Philox call once between them.
// Every operand of a backward product carries 11 significant bits either way (fp16 here, TF32 there); gradients are under the
// caller's loss scale, as the fp16 planes they are written to require anyway.  Tiles hold exactly nq / nk rows: fragment rows
// beyond them read one shared row of zeros.
// Staging: a row arrives as 64 fp32 (256 B, cp.async: every 16-byte piece of all four tensors in flight at once, no registers
// held) and is converted IN PLACE into its fp16 hi half-row (bytes 0..127) and lo half-row (bytes 128..255); row pitch 272 B.
static constexpr int ATT_HP = 136;  // halves per row of a 64-column tile (272 B: the 8 rows of an ldmatrix hit 8 distinct 16-byte slots)
static constexpr int ATT_HL = 64;   // the lo half-row starts 64 halves into the row
static constexpr int ATT_HS = 40;   // halves per row of a score tile (<= 32 keys; 80 B: likewise conflict-free)
static constexpr int ATT_HZ = 72;   // halves of the shared row of zeros
static size_t att_h_bytes(int nq, int nk, bool planes = false) {
  return (size_t)2 * ((size_t)ATT_HP * (nq + nk) + (size_t)(planes ? ATT_HZ : ATT_HP) * (nq + nk) + ATT_HZ + (size_t)ATT_HS * 2 * nq);
}

__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], const __half* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], const __half* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ void mma_f16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  const __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}
// rows x 64 fp32 (global, 16-byte aligned rows) -> the tile's rows, still fp32: one cp.async per 16-byte piece.  Lane l owns
// piece l & 15 of rows (l >> 4), (l >> 4) + 2, ...: both addresses advance by constants
__device__ __forceinline__ void att_stage_h(__half* tile, const float* __restrict__ src, int rows, int ld, int lane) {
  const int r0 = lane >> 4, c4 = lane & 15;
  uint32_t d = (uint32_t)__cvta_generic_to_shared(tile + r0 * ATT_HP + 8 * c4);
  const float* sp = src + (size_t)r0 * ld + 4 * c4;
  const size_t step = (size_t)2 * ld;
  for (int r = r0; r < rows; r += 2, d += 4 * ATT_HP, sp += step)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(sp) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
}
// planes_in: rows of an fp16 hi plane (and lo plane) straight into the tile's half-rows — 8 pieces of 16 bytes per half-row
template <bool LO>
__device__ __forceinline__ void att_stage_p(__half* tile, int pitch, const __half* __restrict__ hi, const __half* __restrict__ lo,
                                            int rows, int ld, int lane) {
  constexpr int PPR = LO ? 16 : 8;   // pieces per row
  const int r0 = lane / PPR, pc = lane % PPR;
  const __half* sp = (LO && pc >= 8 ? lo : hi) + (size_t)r0 * ld + 8 * (pc & 7);
  uint32_t d = (uint32_t)__cvta_generic_to_shared(tile + r0 * pitch + (LO && pc >= 8 ? ATT_HL : 0) + 8 * (pc & 7));
  constexpr int RS = 32 / PPR;       // rows per warp pass
  const size_t step = (size_t)RS * ld;
  for (int r = r0; r < rows; r += RS, d += 2 * RS * pitch, sp += step)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(sp) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
}
// fp32 rows -> fp16 hi (+ lo = fp16(x - hi)) in place.  The 16 lanes that own the pieces of one row read them in the same
// instruction and write that row's half-rows after a warp sync; other rows are untouched, so rows go through in batches of 8
template <bool LO>
__device__ __forceinline__ void att_convert_h(__half* tile, int rows, int lane) {
  const int r0 = lane >> 4, c4 = lane & 15;
  for (int rb = 0; rb < rows; rb += 8) {
    float4 v[4];
    __half* base = tile + (rb + r0) * ATT_HP;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (rb + r0 + 2 * u < rows) v[u] = *reinterpret_cast<const float4*>(base + 2 * u * ATT_HP + 8 * c4);
    __syncwarp();
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (rb + r0 + 2 * u < rows) {
        __half* o = base + 2 * u * ATT_HP + 4 * c4;
        const __half2 h0 = __floats2half2_rn(v[u].x, v[u].y), h1 = __floats2half2_rn(v[u].z, v[u].w);
        *reinterpret_cast<uint2*>(o) = make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
        if (LO) {
          const float2 f0 = __half22float2(h0), f1 = __half22float2(h1);
          *reinterpret_cast<uint2*>(o + ATT_HL) = make_uint2(pack_h2(v[u].x - f0.x, v[u].y - f0.y), pack_h2(v[u].z - f1.x, v[u].w - f1.y));
        }
      }
  }
}

// out[16mj + .., :64] (rows < nj) = T^T . X: T = score tile [ni][ATT_HS] (rows = queries, read transposed through ldmatrix.trans),
// X = [ni][xp] tile (queries x 64, row pitch xp halves); Z = the row of zeros.  MT = 16-query steps, NP = 16-key output blocks.
template <int MT, int NP>
__device__ __forceinline__ void att_h_TtX_store(const __half* T, const __half* X, int xp, const __half* Z, int ni, const AttOut& out, int nj, int lane) {
  const int g = lane >> 2, t = lane & 3, r8 = lane & 7, mi = lane >> 3;
  // per-lane fragment rows (query index = the reduction index of both operands)
  const __half* xrow[MT];
  const __half* trow[MT];
#pragma unroll
  for (int ks = 0; ks < MT; ++ks) {
    const int xr = 16 * ks + r8 + 8 * (mi & 1), tr = 16 * ks + r8 + 8 * (mi >> 1);
    xrow[ks] = (xr < ni ? X + xr * xp : Z) + 8 * (mi >> 1);
    trow[ks] = (tr < ni ? T + tr * ATT_HS : Z) + 8 * (mi & 1);
  }
  __half* orow[NP][2];
#pragma unroll
  for (int mj = 0; mj < NP; ++mj)
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) orow[mj][hh] = out.hi + (size_t)(16 * mj + g + 8 * hh) * out.ld + 2 * t;
#pragma unroll 1
  for (int cq = 0; cq < 2; ++cq) {
    float acc[NP][4][4];
#pragma unroll
    for (int mj = 0; mj < NP; ++mj)
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mj][u][e] = 0.f;
#pragma unroll
    for (int ks = 0; ks < MT; ++ks) {
      // B = X rows (k = query) as a row-major operand: .trans; two loads cover the four 8-column tiles of this 32-column half
      uint32_t xb[2][4];
#pragma unroll
      for (int x = 0; x < 2; ++x) ldsm4t(xb[x], xrow[ks] + 32 * cq + 16 * x);
#pragma unroll
      for (int mj = 0; mj < NP; ++mj) {
        uint32_t af[4];
        ldsm4t(af, trow[ks] + 16 * mj);
#pragma unroll
        for (int x = 0; x < 2; ++x) {
          mma_f16(acc[mj][2 * x], af[0], af[1], af[2], af[3], xb[x][0], xb[x][1]);
          mma_f16(acc[mj][2 * x + 1], af[0], af[1], af[2], af[3], xb[x][2], xb[x][3]);
        }
      }
    }
#pragma unroll
    for (int mj = 0; mj < NP; ++mj)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh)
        if (16 * mj + g + 8 * hh < nj) {
#pragma unroll
          for (int u = 0; u < 4; ++u)
            *reinterpret_cast<uint32_t*>(orow[mj][hh] + 32 * cq + 8 * u) = pack_h2(acc[mj][u][2 * hh], acc[mj][u][2 * hh + 1]);
        }
  }
}

// Precise forward in the same form: scores = Q . K^T and out = W . V both as 3-term fp16 products (hi.hi + lo.hi + hi.lo) with
// ldmatrix fragments; the post-dropout weights W go from the accumulator layout of the score product straight into the A operand of
// W . V as two fp16 terms.  (attention_fwd_tc_kernel: the TF32 version that splits every operand in registers at every use.)
static size_t att_hf_bytes(int nq, int nk, int nt) { return (size_t)2 * ((size_t)ATT_HP * (nq + 2 * nk) + ATT_HZ) + (size_t)16 * nq * nt; }

template <int MT, int NT, bool PLANES>
__global__ void __launch_bounds__(ATT_TC_WARPS * 32)
attention_fwd_h_kernel(AttnArgs a, AttOut out, int warp_bytes) {
  constexpr int NP = (NT + 1) / 2;   // 16-key steps
  pdl_trigger();
  pdl_wait();
  const int wid = threadIdx.x >> 5;
  const int bh = blockIdx.x * ATT_TC_WARPS + wid;
  if (bh >= a.B * a.nh) return;
  const int b = bh / a.nh, h = bh - b * a.nh, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3, r8 = lane & 7, mi = lane >> 3;
  const int nq = a.nq, nk = a.nk;
  extern __shared__ __align__(16) float att_dyn[];
  __half* Qh = reinterpret_cast<__half*>(reinterpret_cast<uint8_t*>(att_dyn) + (size_t)wid * warp_bytes);
  __half* Kh = Qh + nq * ATT_HP;
  __half* Vh = Kh + nk * ATT_HP;
  __half* Z = Vh + nk * ATT_HP;      // one row of zeros
  uint4* dwords = reinterpret_cast<uint4*>(Z + ATT_HZ);   // [nq * NT] dropout words
  if (PLANES) {   // the projections' fp16 planes: half-rows land where the conversion would have put them
    const size_t oq = (size_t)b * nq * a.ldq + h * 64, ok = (size_t)b * a.bsk + h * 64, ov = (size_t)b * a.bsv + h * 64;
    att_stage_p<true>(Qh, ATT_HP, reinterpret_cast<const __half*>(a.q) + oq, a.ql + oq, nq, a.ldq, lane);
    att_stage_p<true>(Kh, ATT_HP, reinterpret_cast<const __half*>(a.k) + ok, a.kl + ok, nk, a.ldk, lane);
    att_stage_p<true>(Vh, ATT_HP, reinterpret_cast<const __half*>(a.v) + ov, a.vl + ov, nk, a.ldv, lane);
  } else {
    att_stage_h(Qh, a.q + (size_t)b * nq * a.ldq + h * 64, nq, a.ldq, lane);
    att_stage_h(Kh, a.k + (size_t)b * a.bsk + h * 64, nk, a.ldk, lane);
    att_stage_h(Vh, a.v + (size_t)b * a.bsv + h * 64, nk, a.ldv, lane);
  }
  for (int i = lane; i < ATT_HZ / 2; i += 32) reinterpret_cast<uint32_t*>(Z)[i] = 0u;
  const bool drop = a.drop_p > 0.f;
  if (drop) {
    for (int c = lane; c < nq * NT; c += 32) {
      uint32_t w[4];
      att_drop_words(a, (size_t)bh * nq + c / NT, c % NT, w);
      dwords[c] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  uint32_t jvalid = 0u, jpad = 0u;
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int j = 8 * nt + 2 * t + e;
      if (j < nk) {
        jvalid |= 1u << (2 * nt + e);
        if (a.key_ids != nullptr && __ldg(a.key_ids + b * nk + j) == a.pad_id) jpad |= 1u << (2 * nt + e);
      }
    }
  if (PLANES) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  } else {
    asm volatile("cp.async.wait_group 2;" ::: "memory");
    __syncwarp();
    att_convert_h<true>(Qh, nq, lane);
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    att_convert_h<true>(Kh, nk, lane);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    att_convert_h<true>(Vh, nk, lane);
  }
  __syncwarp();

  float S[MT][NT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) S[mt][nt][e] = 0.f;
  {
    const __half* kp[NT]; const __half* klp[NT];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int r = 8 * nt + r8;
      const bool ok = r < nk;
      kp[nt] = (ok ? Kh + r * ATT_HP : Z) + 8 * mi;
      klp[nt] = (ok ? Kh + r * ATT_HP + ATT_HL : Z) + 8 * mi;
    }
    const __half* qp[MT]; const __half* qlp[MT];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      const int r = 16 * mt + r8 + 8 * (mi & 1);
      const bool ok = r < nq;
      qp[mt] = (ok ? Qh + r * ATT_HP : Z) + 8 * (mi >> 1);
      qlp[mt] = (ok ? Qh + r * ATT_HP + ATT_HL : Z) + 8 * (mi >> 1);
    }
#pragma unroll 1
    for (int k2 = 0; k2 < 2; ++k2) {
      uint32_t kh[NT][4], kl[NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) { ldsm4(kh[nt], kp[nt] + 32 * k2); ldsm4(kl[nt], klp[nt] + 32 * k2); }
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int kq = 0; kq < 2; ++kq) {
          uint32_t qh[4], ql[4];
          ldsm4(qh, qp[mt] + 32 * k2 + 16 * kq);
          ldsm4(ql, qlp[mt] + 32 * k2 + 16 * kq);
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) mma_f16(S[mt][nt], ql[0], ql[1], ql[2], ql[3], kh[nt][2 * kq], kh[nt][2 * kq + 1]);
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) mma_f16(S[mt][nt], qh[0], qh[1], qh[2], qh[3], kl[nt][2 * kq], kl[nt][2 * kq + 1]);
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) mma_f16(S[mt][nt], qh[0], qh[1], qh[2], qh[3], kh[nt][2 * kq], kh[nt][2 * kq + 1]);
        }
    }
  }
  // softmax, dropout; the weights as two fp16 terms in A-fragment order
  const uint32_t thr = dropout_threshold(a.drop_p);
  const float inv_keep = 1.f / (1.f - a.drop_p);
  const int gd = g - 2 * t;
  uint32_t wh[MT][2 * NP][2], wl[MT][2 * NP][2];
  __syncwarp();
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int i = 16 * mt + g + 8 * hh;
      const bool row_ok = i < nq;
      float p[NT][2];
      float mx = -3.0e38f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const bool m = (a.causal && (8 * nt + e - 16 * mt - 8 * hh) > gd) || ((jpad >> (2 * nt + e)) & 1u);
          float sv = m ? -1e18f : S[mt][nt][2 * hh + e] * a.scale;
          sv = ((jvalid >> (2 * nt + e)) & 1u) ? sv : -3.0e38f;
          p[nt][e] = sv;
          mx = fmaxf(mx, sv);
        }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      float sum = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float pe = ((jvalid >> (2 * nt + e)) & 1u) ? __expf(p[nt][e] - mx) : 0.f;
          p[nt][e] = pe;
          sum += pe;
        }
      sum += __shfl_xor_sync(0xffffffffu, sum, 1);
      sum += __shfl_xor_sync(0xffffffffu, sum, 2);
      const float inv = 1.f / sum;
#pragma unroll
      for (int nt = 0; nt < 2 * NP; ++nt) {
        float w0 = 0.f, w1 = 0.f;
        if (nt < NT) {
          uint32_t w = 0xFFFFFFFFu;
          if (drop && row_ok) w = reinterpret_cast<const uint32_t*>(dwords + (i * NT + nt))[t];
          const float k0 = (!drop || ((w & 0xFFFFu) >= thr)) ? (drop ? inv_keep : 1.f) : 0.f;
          const float k1 = (!drop || ((w >> 16) >= thr)) ? (drop ? inv_keep : 1.f) : 0.f;
          w0 = p[nt][0] * inv * k0;
          w1 = p[nt][1] * inv * k1;
        }
        const __half2 hv = __floats2half2_rn(w0, w1);
        const float2 hf = __half22float2(hv);
        wh[mt][nt][hh] = *reinterpret_cast<const uint32_t*>(&hv);
        wl[mt][nt][hh] = pack_h2(w0 - hf.x, w1 - hf.y);
      }
    }
  // out = W . V (3-term); V rows as a row-major B operand (.trans), hi and lo half-rows
  {
    const __half* vr[NP]; const __half* vlr[NP];
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const int r = 16 * j + r8 + 8 * (mi & 1);
      const bool ok = r < nk;
      vr[j] = (ok ? Vh + r * ATT_HP : Z) + 8 * (mi >> 1);
      vlr[j] = (ok ? Vh + r * ATT_HP + ATT_HL : Z) + 8 * (mi >> 1);
    }
    const size_t obase = (size_t)b * nq * out.ld + h * 64;
#pragma unroll 1
    for (int cq = 0; cq < 2; ++cq) {
      float acc[MT][4][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[mt][u][e] = 0.f;
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        uint32_t vb[2][4], vlb[2][4];
#pragma unroll
        for (int x = 0; x < 2; ++x) { ldsm4t(vb[x], vr[j] + 32 * cq + 16 * x); ldsm4t(vlb[x], vlr[j] + 32 * cq + 16 * x); }
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
#pragma unroll
          for (int x = 0; x < 2; ++x) {
            mma_f16(acc[mt][2 * x], wl[mt][2 * j][0], wl[mt][2 * j][1], wl[mt][2 * j + 1][0], wl[mt][2 * j + 1][1], vb[x][0], vb[x][1]);
            mma_f16(acc[mt][2 * x + 1], wl[mt][2 * j][0], wl[mt][2 * j][1], wl[mt][2 * j + 1][0], wl[mt][2 * j + 1][1], vb[x][2], vb[x][3]);
          }
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
#pragma unroll
          for (int x = 0; x < 2; ++x) {
            mma_f16(acc[mt][2 * x], wh[mt][2 * j][0], wh[mt][2 * j][1], wh[mt][2 * j + 1][0], wh[mt][2 * j + 1][1], vlb[x][0], vlb[x][1]);
            mma_f16(acc[mt][2 * x + 1], wh[mt][2 * j][0], wh[mt][2 * j][1], wh[mt][2 * j + 1][0], wh[mt][2 * j + 1][1], vlb[x][2], vlb[x][3]);
          }
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
#pragma unroll
          for (int x = 0; x < 2; ++x) {
            mma_f16(acc[mt][2 * x], wh[mt][2 * j][0], wh[mt][2 * j][1], wh[mt][2 * j + 1][0], wh[mt][2 * j + 1][1], vb[x][0], vb[x][1]);
            mma_f16(acc[mt][2 * x + 1], wh[mt][2 * j][0], wh[mt][2 * j][1], wh[mt][2 * j + 1][0], wh[mt][2 * j + 1][1], vb[x][2], vb[x][3]);
          }
      }
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const int row = 16 * mt + g + 8 * hh;
          if (row < nq) {
            const size_t o = obase + (size_t)row * out.ld + 32 * cq + 2 * t;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const float v0 = acc[mt][u][2 * hh], v1 = acc[mt][u][2 * hh + 1];
              if (out.hi) {
                const __half2 hv = __floats2half2_rn(v0, v1);
                *reinterpret_cast<uint32_t*>(out.hi + o + 8 * u) = *reinterpret_cast<const uint32_t*>(&hv);
                if (out.lo) {
                  const float2 hf = __half22float2(hv);
                  *reinterpret_cast<uint32_t*>(out.lo + o + 8 * u) = pack_h2(v0 - hf.x, v1 - hf.y);
                }
              } else {
                *reinterpret_cast<float2*>(out.f32 + o + 8 * u) = make_float2(rnd(v0, a.rn), rnd(v1, a.rn));
              }
            }
          }
        }
    }
  }
}

template <int MT, int NT, bool PLANES>
__global__ void __launch_bounds__(ATT_TC_WARPS * 32)
attention_bwd_h_kernel(AttnArgs a, const float* __restrict__ dout, int ldo, AttOut dq, AttOut dk, AttOut dv, int warp_bytes) {
  constexpr int NP = (NT + 1) / 2;   // 16-key steps
  constexpr int PV = PLANES ? ATT_HZ : ATT_HP;   // row pitch of the V and dO tiles: hi half-rows only when they arrive as planes
  pdl_trigger();
  pdl_wait();
  const int wid = threadIdx.x >> 5;
  const int bh = blockIdx.x * ATT_TC_WARPS + wid;
  if (bh >= a.B * a.nh) return;
  const int b = bh / a.nh, h = bh - b * a.nh, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3, r8 = lane & 7, mi = lane >> 3;
  const int nq = a.nq, nk = a.nk;
  extern __shared__ __align__(16) float att_dyn[];
  __half* Qh = reinterpret_cast<__half*>(reinterpret_cast<uint8_t*>(att_dyn) + (size_t)wid * warp_bytes);
  __half* Kh = Qh + nq * ATT_HP;
  __half* Vh = Kh + nk * ATT_HP;
  __half* Oh = Vh + nk * PV;
  __half* Z = Oh + nq * PV;          // one row of zeros
  __half* Ds = Z + ATT_HZ;           // [nq][ATT_HS] dS (hi)
  __half* Ws = Ds + nq * ATT_HS;     // [nq][ATT_HS] post-dropout weights
  if (PLANES) {
    const size_t oq = (size_t)b * nq * a.ldq + h * 64, ok = (size_t)b * a.bsk + h * 64, ov = (size_t)b * a.bsv + h * 64;
    att_stage_p<true>(Qh, ATT_HP, reinterpret_cast<const __half*>(a.q) + oq, a.ql + oq, nq, a.ldq, lane);
    att_stage_p<true>(Kh, ATT_HP, reinterpret_cast<const __half*>(a.k) + ok, a.kl + ok, nk, a.ldk, lane);
    att_stage_p<false>(Vh, PV, reinterpret_cast<const __half*>(a.v) + ov, nullptr, nk, a.ldv, lane);
    att_stage_p<false>(Oh, PV, reinterpret_cast<const __half*>(dout) + (size_t)b * nq * ldo + h * 64, nullptr, nq, ldo, lane);
  } else {
    att_stage_h(Qh, a.q + (size_t)b * nq * a.ldq + h * 64, nq, a.ldq, lane);
    att_stage_h(Kh, a.k + (size_t)b * a.bsk + h * 64, nk, a.ldk, lane);
    att_stage_h(Vh, a.v + (size_t)b * a.bsv + h * 64, nk, a.ldv, lane);
    att_stage_h(Oh, dout + (size_t)b * nq * ldo + h * 64, nq, ldo, lane);
  }
  for (int i = lane; i < ATT_HZ / 2; i += 32) reinterpret_cast<uint32_t*>(Z)[i] = 0u;
  // the dropout words of this (sample, head): call (row i, key tile c) is made once, by lane (i * NT + c) % 32, and left in the
  // score-tile region (free until dS / W are written) as four words; every lane then picks word t of the calls of its rows
  const bool drop = a.drop_p > 0.f;
  uint4* dwords = reinterpret_cast<uint4*>(Ds);
  if (drop) {
    for (int c = lane; c < nq * NT; c += 32) {
      uint32_t w[4];
      att_drop_words(a, (size_t)bh * nq + c / NT, c % NT, w);
      dwords[c] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  // masks of this lane's keys j = 8 nt + 2 t + e (bit 2 nt + e): inside the sequence, and padding
  uint32_t jvalid = 0u, jpad = 0u;
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int j = 8 * nt + 2 * t + e;
      if (j < nk) {
        jvalid |= 1u << (2 * nt + e);
        if (a.key_ids != nullptr && __ldg(a.key_ids + b * nk + j) == a.pad_id) jpad |= 1u << (2 * nt + e);
      }
    }
  // the tensors land in issue order: convert each as soon as its group is complete, under the loads of the later ones
  if (PLANES) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  } else {
    asm volatile("cp.async.wait_group 3;" ::: "memory");
    __syncwarp();
    att_convert_h<true>(Qh, nq, lane);
    asm volatile("cp.async.wait_group 2;" ::: "memory");
    __syncwarp();
    att_convert_h<true>(Kh, nk, lane);
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    att_convert_h<false>(Vh, nk, lane);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    att_convert_h<false>(Oh, nq, lane);
  }
  __syncwarp();

  float S[MT][NT][4], dW[MT][NT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) { S[mt][nt][e] = 0.f; dW[mt][nt][e] = 0.f; }
  // scores = Q . K^T (3-term) and d(post-dropout weights) = dO . V^T (single pass); K / V rows are the B operand as stored
  {
    const __half* kp[NT]; const __half* klp[NT]; const __half* vp[NT];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int r = 8 * nt + r8;
      const bool ok = r < nk;
      kp[nt] = (ok ? Kh + r * ATT_HP : Z) + 8 * mi;
      klp[nt] = (ok ? Kh + r * ATT_HP + ATT_HL : Z) + 8 * mi;
      vp[nt] = (ok ? Vh + r * PV : Z) + 8 * mi;
    }
    const __half* qp[MT]; const __half* qlp[MT]; const __half* op[MT];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      const int r = 16 * mt + r8 + 8 * (mi & 1);
      const bool ok = r < nq;
      qp[mt] = (ok ? Qh + r * ATT_HP : Z) + 8 * (mi >> 1);
      qlp[mt] = (ok ? Qh + r * ATT_HP + ATT_HL : Z) + 8 * (mi >> 1);
      op[mt] = (ok ? Oh + r * PV : Z) + 8 * (mi >> 1);
    }
#pragma unroll 1
    for (int k2 = 0; k2 < 2; ++k2) {   // 32 head columns per pass
      uint32_t kh[NT][4], kl[NT][4], vh[NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        ldsm4(kh[nt], kp[nt] + 32 * k2);
        ldsm4(kl[nt], klp[nt] + 32 * k2);
        ldsm4(vh[nt], vp[nt] + 32 * k2);
      }
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int kq = 0; kq < 2; ++kq) {
          uint32_t qh[4], ql[4], oh[4];
          ldsm4(qh, qp[mt] + 32 * k2 + 16 * kq);
          ldsm4(ql, qlp[mt] + 32 * k2 + 16 * kq);
          ldsm4(oh, op[mt] + 32 * k2 + 16 * kq);
          // term-major: consecutive MMAs go to different accumulators (a dependent accumulate waits out the whole MMA latency)
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) mma_f16(S[mt][nt], ql[0], ql[1], ql[2], ql[3], kh[nt][2 * kq], kh[nt][2 * kq + 1]);
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) mma_f16(dW[mt][nt], oh[0], oh[1], oh[2], oh[3], vh[nt][2 * kq], vh[nt][2 * kq + 1]);
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) mma_f16(S[mt][nt], qh[0], qh[1], qh[2], qh[3], kl[nt][2 * kq], kl[nt][2 * kq + 1]);
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) mma_f16(S[mt][nt], qh[0], qh[1], qh[2], qh[3], kh[nt][2 * kq], kh[nt][2 * kq + 1]);
        }
    }
  }
  // softmax over the keys, dP = dW * keep, dS = P (dP - sum_j dP P) scale (0 where masked), W = P * keep — per score row
  // (16 mt + g + 8 hh; rows >= nq hold finite junk that no stored result depends on).  dS leaves as two fp16 terms in A-fragment
  // order for dQ and, with W, as the shared tiles whose transposes feed dK / dV.
  const uint32_t thr = dropout_threshold(a.drop_p);
  const float inv_keep = 1.f / (1.f - a.drop_p);
  const int gd = g - 2 * t;   // causal: key j = 8 nt + 2 t + e is masked for row i = 16 mt + 8 hh + g when 8 nt + e - 16 mt - 8 hh > g - 2 t
  uint32_t dsh[MT][2 * NP][2], dsl[MT][2 * NP][2];
  __syncwarp();   // (dropout words are read below, the tiles are written after the next sync)
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int i = 16 * mt + g + 8 * hh;
      const bool row_ok = i < nq;
      float p[NT][2], kp2[NT][2];
      uint32_t mbits = 0u;
      float mx = -3.0e38f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const bool m = (a.causal && (8 * nt + e - 16 * mt - 8 * hh) > gd) || ((jpad >> (2 * nt + e)) & 1u);
          if (m) mbits |= 1u << (2 * nt + e);
          float sv = m ? -1e18f : S[mt][nt][2 * hh + e] * a.scale;
          sv = ((jvalid >> (2 * nt + e)) & 1u) ? sv : -3.0e38f;
          p[nt][e] = sv;
          mx = fmaxf(mx, sv);
        }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      float sum = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float pe = ((jvalid >> (2 * nt + e)) & 1u) ? __expf(p[nt][e] - mx) : 0.f;
          p[nt][e] = pe;
          sum += pe;
        }
      sum += __shfl_xor_sync(0xffffffffu, sum, 1);
      sum += __shfl_xor_sync(0xffffffffu, sum, 2);
      const float inv = 1.f / sum;
      float dot = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        uint32_t w = 0xFFFFFFFFu;
        if (drop && row_ok) w = reinterpret_cast<const uint32_t*>(dwords + (i * NT + nt))[t];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          p[nt][e] *= inv;
          kp2[nt][e] = (!drop || (((w >> (16 * e)) & 0xFFFFu) >= thr)) ? (drop ? inv_keep : 1.f) : 0.f;
          const float dP = dW[mt][nt][2 * hh + e] * kp2[nt][e];
          dW[mt][nt][2 * hh + e] = dP;
          dot += dP * p[nt][e];
        }
      }
      dot += __shfl_xor_sync(0xffffffffu, dot, 1);
      dot += __shfl_xor_sync(0xffffffffu, dot, 2);
#pragma unroll
      for (int nt = 0; nt < 2 * NP; ++nt) {
        float ds0 = 0.f, ds1 = 0.f;
        if (nt < NT) {
          ds0 = ((mbits >> (2 * nt)) & 1u) ? 0.f : p[nt][0] * (dW[mt][nt][2 * hh] - dot) * a.scale;
          ds1 = ((mbits >> (2 * nt + 1)) & 1u) ? 0.f : p[nt][1] * (dW[mt][nt][2 * hh + 1] - dot) * a.scale;
          S[mt][nt][2 * hh] = p[nt][0] * kp2[nt][0];       // W replaces the scores (written to its tile below)
          S[mt][nt][2 * hh + 1] = p[nt][1] * kp2[nt][1];
        }
        const __half2 dh = __floats2half2_rn(ds0, ds1);
        const float2 df = __half22float2(dh);
        dsh[mt][nt][hh] = *reinterpret_cast<const uint32_t*>(&dh);
        dsl[mt][nt][hh] = pack_h2(ds0 - df.x, ds1 - df.y);
      }
    }
  __syncwarp();   // every lane has read its dropout words: the region becomes the dS / W tiles
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int i = 16 * mt + g + 8 * hh;
      if (i < nq) {
        uint32_t* drow = reinterpret_cast<uint32_t*>(Ds + i * ATT_HS + 2 * t);
        uint32_t* wrow = reinterpret_cast<uint32_t*>(Ws + i * ATT_HS + 2 * t);
#pragma unroll
        for (int nt = 0; nt < 2 * NP; ++nt) {
          drow[4 * nt] = dsh[mt][nt][hh];
          wrow[4 * nt] = (nt < NT) ? pack_h2(S[mt][nt < NT ? nt : 0][2 * hh], S[mt][nt < NT ? nt : 0][2 * hh + 1]) : 0u;   // (index clamp: compile-time)
        }
      }
    }
  __syncwarp();
  // dQ = (dS + dS_lo) . K: dS straight from the accumulator layout (two 8-key tiles = one 16-key A fragment), K rows as a
  // row-major B operand (.trans)
  {
    const __half* kr[NP];
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const int r = 16 * j + r8 + 8 * (mi & 1);
      kr[j] = (r < nk ? Kh + r * ATT_HP : Z) + 8 * (mi >> 1);
    }
    __half* orow[MT][2];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) orow[mt][hh] = dq.hi + ((size_t)b * nq + 16 * mt + g + 8 * hh) * dq.ld + h * 64 + 2 * t;
#pragma unroll 1
    for (int cq = 0; cq < 2; ++cq) {
      float acc[MT][4][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[mt][u][e] = 0.f;
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        uint32_t kb[2][4];
#pragma unroll
        for (int x = 0; x < 2; ++x) ldsm4t(kb[x], kr[j] + 32 * cq + 16 * x);
        if (a.dq2) {
#pragma unroll
          for (int mt = 0; mt < MT; ++mt)
#pragma unroll
            for (int x = 0; x < 2; ++x) {
              mma_f16(acc[mt][2 * x], dsl[mt][2 * j][0], dsl[mt][2 * j][1], dsl[mt][2 * j + 1][0], dsl[mt][2 * j + 1][1], kb[x][0], kb[x][1]);
              mma_f16(acc[mt][2 * x + 1], dsl[mt][2 * j][0], dsl[mt][2 * j][1], dsl[mt][2 * j + 1][0], dsl[mt][2 * j + 1][1], kb[x][2], kb[x][3]);
            }
        }
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
#pragma unroll
          for (int x = 0; x < 2; ++x) {
            mma_f16(acc[mt][2 * x], dsh[mt][2 * j][0], dsh[mt][2 * j][1], dsh[mt][2 * j + 1][0], dsh[mt][2 * j + 1][1], kb[x][0], kb[x][1]);
            mma_f16(acc[mt][2 * x + 1], dsh[mt][2 * j][0], dsh[mt][2 * j][1], dsh[mt][2 * j + 1][0], dsh[mt][2 * j + 1][1], kb[x][2], kb[x][3]);
          }
      }
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh)
          if (16 * mt + g + 8 * hh < nq) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
              *reinterpret_cast<uint32_t*>(orow[mt][hh] + 32 * cq + 8 * u) = pack_h2(acc[mt][u][2 * hh], acc[mt][u][2 * hh + 1]);
          }
    }
  }
  att_h_TtX_store<MT, NP>(Ds, Qh, ATT_HP, Z, nq, att_out_at(dk, (size_t)b * nk * dk.ld + h * 64), nk, lane);   // dK = dS^T . Q
  att_h_TtX_store<MT, NP>(Ws, Oh, PV, Z, nq, att_out_at(dv, (size_t)b * nk * dv.ld + h * 64), nk, lane);       // dV = W^T . dO
}

// ---------------------------------------------------------------- row-block kernels: 32 < n <= 80 (cfg5: T = 64, R = 65)
// One CTA per (sample, head), one WARP per 16 query rows; the warp keeps its 16 x nk slab of scores in registers
// (NT <= 10 key tiles of 8).  Forward: warps are independent, no shared memory.  Backward: every warp writes its rows of
// dS and of the post-dropout weights W into two CTA-wide shared tiles ([16*MT][ATT_LD]); after one barrier warp w owns the
// 16 keys 16w .. 16w+15 and reduces dK = dS^T Q and dV = W^T dO over ALL query rows.  ATT_LD = 84: the transposed-operand
// reads (rows 8s + 2t (+1), columns g / g + 8) then fall into 32 distinct banks (2 * 84 mod 32 = 8).
static constexpr int ATT_LD = 84;
static constexpr int ATT_ROWBLOCK_MAX = 80;

template <int NT>
__global__ void __launch_bounds__(160)
attention_fwd_rows_kernel(AttnArgs a, AttOut out) {
  const int ldo = out.ld;
  using MaskT = typename std::conditional<(NT > 8), unsigned long long, uint32_t>::type;
  pdl_trigger();
  pdl_wait();
  const int bh = blockIdx.x, w = threadIdx.x >> 5, i0 = 16 * w;
  if (i0 >= a.nq) return;
  const int b = bh / a.nh, h = bh - b * a.nh, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const float* Q = a.q + ((size_t)b * a.nq + i0) * a.ldq + h * 64;
  const float* K = a.k + (size_t)b * a.bsk + h * 64;
  const float* V = a.v + (size_t)b * a.bsv + h * 64;
  float S[1][NT][4];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int e = 0; e < 4; ++e) S[0][nt][e] = 0.f;
  mma_xyT<1, NT, 3>(S, Q, a.nq - i0, a.ldq, K, a.nk, a.ldk, g, t);
  MaskT msk[1];
  softmax_frag<1, NT, MaskT>(a, S, msk, b, g, t, i0);
  float SL[1][NT][4];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int i = i0 + g + 8 * hh;
      uint32_t dw[4];
      if (a.drop_p > 0.f) att_drop_words(a, (size_t)bh * a.nq + i, nt, dw);
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int j = 8 * nt + 2 * t + e;
        float wgt = S[0][nt][2 * hh + e];
        if (a.drop_p > 0.f && i < a.nq && j < a.nk) wgt *= att_keep(dw, j, dropout_threshold(a.drop_p), 1.f / (1.f - a.drop_p));
        const float wh = tf32_rn(wgt);
        S[0][nt][2 * hh + e] = wh;
        SL[0][nt][2 * hh + e] = wgt - wh;
      }
    }
  mma_pY_store<1, NT, 3>(S, V, a.nk, a.ldv, att_out_at(out, ((size_t)b * a.nq + i0) * ldo + h * 64), a.nq - i0, g, t, a.rn, SL);
}

// out[16mj + .., :64] = sum_{i < ni} T[i][j] * X[i, :64] for the 16 keys of block mj (rows j < nj are stored); T is a shared
// tile [ni][ATT_LD] (ni a multiple of 8, columns up to 16*ceil(NT/2) readable).  Reduction slot t of k-step s is row 8s + 2t,
// slot t + 4 is row 8s + 2t + 1, on both operands.
__device__ __forceinline__ void mma_TtX_block_store(const float* T, int ni, int ncols, const float* X, int nx, int ldx, const AttOut& out,
                                                    int mj, int nj, int g, int t, int rn) {
  const int ldo = out.ld;
  const int j0 = 16 * mj + g, j1 = j0 + 8;
#pragma unroll 1
  for (int cq = 0; cq < 2; ++cq) {
    float acc[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[u][e] = 0.f;
    // the rows of X come straight from global memory: 5 k-steps of loads are issued before the first MMA needs one (a warp that
    // waits for every pair of loads in turn spends the kernel in L2 latency: 10 trips x 2 products x 2 column halves)
    constexpr int PF = 5;
    const int ns = ni / 8;
    for (int s0 = 0; s0 < ns; s0 += PF) {
      float4 x0[PF], x1[PF];
#pragma unroll
      for (int p = 0; p < PF; ++p) {
        const int i = 8 * (s0 + p) + 2 * t;      // rows >= nx (and every row of a k-step >= ns) read as zero
        x0[p] = hi4(ld4(X, i, nx, ldx, 32 * cq + 4 * g));
        x1[p] = hi4(ld4(X, i + 1, nx, ldx, 32 * cq + 4 * g));
      }
#pragma unroll
      for (int p = 0; p < PF; ++p) {
        if (s0 + p < ns) {
          const int i = 8 * (s0 + p) + 2 * t;
          const float* x0a = &x0[p].x; const float* x1a = &x1[p].x;
          const float a0 = T[i * ATT_LD + j0], a2 = T[(i + 1) * ATT_LD + j0];
          const float a1 = (j1 < ncols) ? T[i * ATT_LD + j1] : 0.f, a3 = (j1 < ncols) ? T[(i + 1) * ATT_LD + j1] : 0.f;
#pragma unroll
          for (int u = 0; u < 4; ++u) mma_tf32(acc[u], a0, a1, a2, a3, x0a[u], x1a[u]);
        }
      }
    }
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int row = 16 * mj + g + 8 * hh;
      if (row < nj) {
        const float v[8] = {acc[0][2 * hh], acc[1][2 * hh], acc[2][2 * hh], acc[3][2 * hh],
                            acc[0][2 * hh + 1], acc[1][2 * hh + 1], acc[2][2 * hh + 1], acc[3][2 * hh + 1]};
        att_store8(out, (size_t)row * ldo + 32 * cq + 8 * t, v, rn);
      }
    }
  }
}

template <int NT>
__global__ void __launch_bounds__(160, 2)
attention_bwd_rows_kernel(AttnArgs a, const float* __restrict__ dout, int ldo, AttOut dq, AttOut dk, AttOut dv) {
  const int lddq = dq.ld, lddk = dk.ld, lddv = dv.ld;
  using MaskT = typename std::conditional<(NT > 8), unsigned long long, uint32_t>::type;
  extern __shared__ float rows_sm[];
  pdl_trigger();
  pdl_wait();
  const int mtiles = (a.nq + 15) / 16, ni = 16 * mtiles;
  float* Ws = rows_sm;
  float* Ds = rows_sm + (size_t)ni * ATT_LD;
  const int bh = blockIdx.x, w = threadIdx.x >> 5, i0 = 16 * w;
  const int b = bh / a.nh, h = bh - b * a.nh, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const float* Qh = a.q + (size_t)b * a.nq * a.ldq + h * 64;
  const float* K = a.k + (size_t)b * a.bsk + h * 64;
  const float* V = a.v + (size_t)b * a.bsv + h * 64;
  const float* dOh = dout + (size_t)b * a.nq * ldo + h * 64;
  if (w < mtiles) {
    float S[1][NT][4], dW[1][NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) { S[0][nt][e] = 0.f; dW[0][nt][e] = 0.f; }
    mma_xyT<1, NT, 3>(S, Qh + (size_t)i0 * a.ldq, a.nq - i0, a.ldq, K, a.nk, a.ldk, g, t);  // scores (the forward's 3-term product)
    mma_xyT<1, NT, 1>(dW, dOh + (size_t)i0 * ldo, a.nq - i0, ldo, V, a.nk, a.ldv, g, t);    // d(post-dropout weights) = dO . V^T
    MaskT msk[1];
    softmax_frag<1, NT, MaskT>(a, S, msk, b, g, t, i0);
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int i = i0 + g + 8 * hh;
      float ks[NT][2], dot = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        uint32_t dw[4];
        if (a.drop_p > 0.f) att_drop_words(a, (size_t)bh * a.nq + i, nt, dw);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = 8 * nt + 2 * t + e;
          ks[nt][e] = (a.drop_p > 0.f && i < a.nq && j < a.nk) ? att_keep(dw, j, dropout_threshold(a.drop_p), 1.f / (1.f - a.drop_p)) : 1.f;
          const float dP = dW[0][nt][2 * hh + e] * ks[nt][e];
          dW[0][nt][2 * hh + e] = dP;
          dot += dP * S[0][nt][2 * hh + e];
        }
      }
      dot += __shfl_xor_sync(0xffffffffu, dot, 1);
      dot += __shfl_xor_sync(0xffffffffu, dot, 2);
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = 8 * nt + 2 * t + e, idx = 2 * hh + e;
          const float p = S[0][nt][idx];
          const bool m = (msk[0] >> (nt * 4 + idx)) & 1u;
          const float dsf = (m || i >= a.nq) ? 0.f : p * (dW[0][nt][idx] - dot) * a.scale;
          const float ds = tf32_rn(dsf);
          const float wgt = (i >= a.nq) ? 0.f : tf32_rn(p * ks[nt][e]);
          S[0][nt][idx] = ds;
          dW[0][nt][idx] = tf32_rn(dsf - ds);    // low-order term of dS for the dQ product (see attention_bwd_tc_kernel)
          Ws[i * ATT_LD + j] = wgt;
          Ds[i * ATT_LD + j] = ds;
        }
    }
    mma_pY_store<1, NT, 1>(S, K, a.nk, a.ldk, att_out_at(dq, ((size_t)b * a.nq + i0) * lddq + h * 64), a.nq - i0, g, t, a.rn, a.dq2 ? dW : nullptr);   // dQ = (dS + dS_lo) . K
  }
  __syncthreads();
  if (16 * w < a.nk) {
    mma_TtX_block_store(Ds, ni, 8 * NT, Qh, a.nq, a.ldq, att_out_at(dk, (size_t)b * a.nk * lddk + h * 64), w, a.nk, g, t, a.rn);  // dK = dS^T . Q
    mma_TtX_block_store(Ws, ni, 8 * NT, dOh, a.nq, ldo, att_out_at(dv, (size_t)b * a.nk * lddv + h * 64), w, a.nk, g, t, a.rn);   // dV = W^T . dO
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
  a.scale = p->scale; a.drop_p = p->drop_p; a.seed = p->seed; a.offset = p->offset; a.seed_dev = p->seed_dev;
  a.rn = round_mode();
  a.ql = (const __half*)p->q_lo; a.kl = (const __half*)p->k_lo; a.vl = (const __half*)p->v_lo;
  static const int dq2 = [] { const char* e = getenv("B200IQ_ATT_DQ2"); return (e && e[0] == '0') ? 0 : 1; }();
  a.dq2 = dq2;
  a.bsk = p->kv_batch_stride > 0 ? (size_t)p->kv_batch_stride : (size_t)p->nk * p->ldk;
  a.bsv = p->kv_batch_stride > 0 ? (size_t)p->kv_batch_stride : (size_t)p->nk * p->ldv;
  return a;
}

// B200IQ_ATT_SIMT=1: every attention call takes the exact SIMT kernels (A/B switch for numerics / timing experiments)
static bool att_force_simt() {
  static const int v = [] { const char* e = getenv("B200IQ_ATT_SIMT"); return (e && e[0] == '1') ? 1 : 0; }();
  return v != 0;
}
// the tensor-core kernels take fp32 rows with 16-byte alignment and plane rows with 16-byte alignment (ld % 8 halves)
static bool outs_ok(const AttOut* outs, int n) {
  for (int i = 0; i < n; ++i) {
    if (outs[i].hi) {
      if ((((uintptr_t)outs[i].hi) | ((uintptr_t)outs[i].lo)) & 15) return false;
      if (outs[i].ld & 7) return false;
    } else {
      if (((uintptr_t)outs[i].f32) & 15) return false;
      if (outs[i].ld & 3) return false;
    }
  }
  return true;
}
static bool tc_common_ok(const B200iqAttention* p, const void* const* ptrs, int nptr, const int* lds, int nld, const AttOut* outs, int nout) {
  if (att_force_simt()) return false;
  // legacy callers (TF32 rounding mode on) or the precise mode of the fp16-plane path; the exact mode (rounding off, precise 0)
  // keeps the CUDA-core kernels
  if (!(round_mode() || p->precise) || p->head_dim != 64) return false;
  for (int i = 0; i < nptr; ++i) if (((uintptr_t)ptrs[i]) & 15) return false;
  for (int i = 0; i < nld; ++i) if (lds[i] & 3) return false;
  return outs_ok(outs, nout);
}
static bool tc_ok(const B200iqAttention* p, const void* const* ptrs, int nptr, const int* lds, int nld, const AttOut* outs, int nout) {
  return p->nq <= 32 && p->nk <= 32 && tc_common_ok(p, ptrs, nptr, lds, nld, outs, nout);
}
// the row-block kernels: anything the register-resident kernels above do not take, up to 80 x 80
static bool rows_ok(const B200iqAttention* p, const void* const* ptrs, int nptr, const int* lds, int nld, const AttOut* outs, int nout) {
  return p->nq <= ATT_ROWBLOCK_MAX && p->nk <= ATT_ROWBLOCK_MAX && tc_common_ok(p, ptrs, nptr, lds, nld, outs, nout);
}
#define B200IQ_ATT_ROWS_DISPATCH(NT_, CALL) \
  do { \
    switch (NT_) { \
      case 1: CALL(1); break; case 2: CALL(2); break; case 3: CALL(3); break; case 4: CALL(4); break; case 5: CALL(5); break; \
      case 6: CALL(6); break; case 7: CALL(7); break; case 8: CALL(8); break; case 9: CALL(9); break; default: CALL(10); break; \
    } \
  } while (0)
#define B200IQ_ATT_DISPATCH(MT_, NT_, CALL) \
  do { \
    if (MT_ == 1) { if (NT_ == 1) { CALL(1, 1); } else if (NT_ == 2) { CALL(1, 2); } else if (NT_ == 3) { CALL(1, 3); } else { CALL(1, 4); } } \
    else          { if (NT_ == 1) { CALL(2, 1); } else if (NT_ == 2) { CALL(2, 2); } else if (NT_ == 3) { CALL(2, 3); } else { CALL(2, 4); } } \
  } while (0)

static int attention_fwd_impl(const B200iqAttention* p, const AttOut& out, void* stream) {
  AttnArgs a = to_args(p);
  {
    const void* ptrs[] = {p->q, p->k, p->v};
    const int lds[] = {p->ldq, p->ldk, p->ldv};
    // B200IQ_ATT_FWD_H=0: the TF32 register kernel for the precise mode too (A/B switch)
    static const bool use_h = [] { const char* e = getenv("B200IQ_ATT_FWD_H"); return !(e && e[0] == '0'); }();
    if (p->planes_in) {
      const int lds8[] = {p->ldq, p->ldk, p->ldv};
      bool okp = p->precise && p->head_dim == 64 && p->nq <= 32 && p->nk <= 32 && p->q_lo && p->k_lo && p->v_lo && outs_ok(&out, 1);
      for (int i = 0; i < 3; ++i) okp = okp && (lds8[i] & 7) == 0;
      const void* pp[] = {p->q, p->k, p->v, p->q_lo, p->k_lo, p->v_lo};
      for (int i = 0; i < 6; ++i) okp = okp && (((uintptr_t)pp[i]) & 15) == 0;
      B200IQ_CHECK_ARG(okp, "attention_fwd: planes_in needs precise mode, head_dim 64, nq, nk <= 32, all three lo planes, 16-byte aligned rows");
      const int mt = a.nq > 16 ? 2 : 1, nt = (a.nk + 7) / 8, grid = (a.B * a.nh + ATT_TC_WARPS - 1) / ATT_TC_WARPS;
      const int wbytes = (int)((att_hf_bytes(a.nq, a.nk, nt) + 15) & ~(size_t)15);
      const size_t tsm = (size_t)ATT_TC_WARPS * wbytes;
#define CALL(M_, N_) \
  do { \
    if (tsm > 48 * 1024) cudaFuncSetAttribute(attention_fwd_h_kernel<M_, N_, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm); \
    launch_dep(attention_fwd_h_kernel<M_, N_, true>, dim3(grid), dim3(ATT_TC_WARPS * 32), tsm, (cudaStream_t)stream, a, out, wbytes); \
  } while (0)
      B200IQ_ATT_DISPATCH(mt, nt, CALL);
#undef CALL
      return launch_status("attention_fwd_h (planes)");
    }
    if (use_h && p->precise && tc_ok(p, ptrs, 3, lds, 3, &out, 1)) {
      const int mt = a.nq > 16 ? 2 : 1, nt = (a.nk + 7) / 8, grid = (a.B * a.nh + ATT_TC_WARPS - 1) / ATT_TC_WARPS;
      const int wbytes = (int)((att_hf_bytes(a.nq, a.nk, nt) + 15) & ~(size_t)15);
      const size_t tsm = (size_t)ATT_TC_WARPS * wbytes;
#define CALL(M_, N_) \
  do { \
    if (tsm > 48 * 1024) cudaFuncSetAttribute(attention_fwd_h_kernel<M_, N_, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm); \
    launch_dep(attention_fwd_h_kernel<M_, N_, false>, dim3(grid), dim3(ATT_TC_WARPS * 32), tsm, (cudaStream_t)stream, a, out, wbytes); \
  } while (0)
      B200IQ_ATT_DISPATCH(mt, nt, CALL);
#undef CALL
      return launch_status("attention_fwd_h");
    }
    if (tc_ok(p, ptrs, 3, lds, 3, &out, 1)) {
      const int mt = a.nq > 16 ? 2 : 1, nt = (a.nk + 7) / 8, grid = (a.B * a.nh + ATT_TC_WARPS - 1) / ATT_TC_WARPS;
      const size_t tsm = sizeof(float) * ATT_TC_WARPS * (size_t)(a.nq + 2 * a.nk) * ATT_SP;
#define CALL(M_, N_) \
  do { \
    if (tsm > 48 * 1024) cudaFuncSetAttribute(attention_fwd_tc_kernel<M_, N_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm); \
    launch_dep(attention_fwd_tc_kernel<M_, N_>, dim3(grid), dim3(ATT_TC_WARPS * 32), tsm, (cudaStream_t)stream, a, out); \
  } while (0)
      B200IQ_ATT_DISPATCH(mt, nt, CALL);
#undef CALL
      return launch_status("attention_fwd_tc");
    }
    if (rows_ok(p, ptrs, 3, lds, 3, &out, 1)) {
      const int nt = (a.nk + 7) / 8, warps = (a.nq + 15) / 16;
#define CALL(N_) launch_dep(attention_fwd_rows_kernel<N_>, dim3(a.B * a.nh), dim3(32 * warps), 0, (cudaStream_t)stream, a, out)
      B200IQ_ATT_ROWS_DISPATCH(nt, CALL);
#undef CALL
      return launch_status("attention_fwd_rows");
    }
  }
  size_t smem = sizeof(float) * ((size_t)(a.nq + 2 * a.nk) * (a.d + 1) + (ATT_THREADS / 32) * a.nk);
  B200IQ_CHECK_ARG(smem <= 227 * 1024, "attention_fwd: tile does not fit shared memory (%zu B)", smem);
  if (smem > 48 * 1024) cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  attention_fwd_kernel<<<a.B * a.nh, ATT_THREADS, smem, (cudaStream_t)stream>>>(a, out);
  return launch_status("attention_fwd");
}

static int attention_bwd_impl(const B200iqAttention* p, const float* dout, const AttOut& dq, const AttOut& dk, const AttOut& dv, void* stream) {
  AttnArgs a = to_args(p);
  {
    const void* ptrs[] = {p->q, p->k, p->v, dout};
    const int lds[] = {p->ldq, p->ldk, p->ldv, p->ldo};
    const AttOut outs[] = {dq, dk, dv};
    // B200IQ_ATT_BWD_H=0: keep the TF32 register kernel for the fp16-plane gradients too (A/B switch)
    static const bool use_h = [] { const char* e = getenv("B200IQ_ATT_BWD_H"); return !(e && e[0] == '0'); }();
    const bool planes_hi = dq.hi && dk.hi && dv.hi && !dq.lo && !dk.lo && !dv.lo;
    if (p->planes_in) {
      const int lds8[] = {p->ldq, p->ldk, p->ldv, p->ldo};
      bool okp = p->precise && planes_hi && p->head_dim == 64 && p->nq <= 32 && p->nk <= 32 && p->q_lo && p->k_lo && outs_ok(outs, 3);
      for (int i = 0; i < 4; ++i) okp = okp && (lds8[i] & 7) == 0;
      const void* pp[] = {p->q, p->k, p->v, p->q_lo, p->k_lo, dout};
      for (int i = 0; i < 6; ++i) okp = okp && (((uintptr_t)pp[i]) & 15) == 0;
      B200IQ_CHECK_ARG(okp, "attention_bwd: planes_in needs precise mode, fp16-plane gradients, head_dim 64, nq, nk <= 32, q_lo / k_lo, 16-byte aligned rows");
      const int mt = a.nq > 16 ? 2 : 1, nt = (a.nk + 7) / 8, grid = (a.B * a.nh + ATT_TC_WARPS - 1) / ATT_TC_WARPS;
      const int wbytes = (int)((att_h_bytes(a.nq, a.nk, true) + 15) & ~(size_t)15);
      const size_t tsm = (size_t)ATT_TC_WARPS * wbytes;
#define CALL(M_, N_) \
  do { \
    if (tsm > 48 * 1024) cudaFuncSetAttribute(attention_bwd_h_kernel<M_, N_, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm); \
    launch_dep(attention_bwd_h_kernel<M_, N_, true>, dim3(grid), dim3(ATT_TC_WARPS * 32), tsm, (cudaStream_t)stream, a, dout, p->ldo, dq, dk, dv, wbytes); \
  } while (0)
      B200IQ_ATT_DISPATCH(mt, nt, CALL);
#undef CALL
      return launch_status("attention_bwd_h (planes)");
    }
    if (use_h && p->precise && planes_hi && tc_ok(p, ptrs, 4, lds, 4, outs, 3)) {
      const int mt = a.nq > 16 ? 2 : 1, nt = (a.nk + 7) / 8, grid = (a.B * a.nh + ATT_TC_WARPS - 1) / ATT_TC_WARPS;
      const int wbytes = (int)((att_h_bytes(a.nq, a.nk) + 15) & ~(size_t)15);
      const size_t tsm = (size_t)ATT_TC_WARPS * wbytes;
#define CALL(M_, N_) \
  do { \
    if (tsm > 48 * 1024) cudaFuncSetAttribute(attention_bwd_h_kernel<M_, N_, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm); \
    launch_dep(attention_bwd_h_kernel<M_, N_, false>, dim3(grid), dim3(ATT_TC_WARPS * 32), tsm, (cudaStream_t)stream, a, dout, p->ldo, dq, dk, dv, wbytes); \
  } while (0)
      B200IQ_ATT_DISPATCH(mt, nt, CALL);
#undef CALL
      return launch_status("attention_bwd_h");
    }
    if (tc_ok(p, ptrs, 4, lds, 4, outs, 3)) {
      const int mt = a.nq > 16 ? 2 : 1, nt = (a.nk + 7) / 8, grid = (a.B * a.nh + ATT_TC_WARPS - 1) / ATT_TC_WARPS;
      const size_t tsm = sizeof(float) * ATT_TC_WARPS * (size_t)(2 * a.nq + 2 * a.nk) * ATT_SP;
#define CALL(M_, N_) \
  do { \
    /* static (the two transposed tiles per warp) + dynamic may pass the 48 KB a kernel gets without opting in */ \
    if (tsm + sizeof(float) * ATT_TC_WARPS * 2 * 32 * 33 > 48 * 1024) \
      cudaFuncSetAttribute(attention_bwd_tc_kernel<M_, N_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm); \
    launch_dep(attention_bwd_tc_kernel<M_, N_>, dim3(grid), dim3(ATT_TC_WARPS * 32), tsm, (cudaStream_t)stream, a, dout, p->ldo, dq, dk, dv); \
  } while (0)
      B200IQ_ATT_DISPATCH(mt, nt, CALL);
#undef CALL
      return launch_status("attention_bwd_tc");
    }
    if (rows_ok(p, ptrs, 4, lds, 4, outs, 3)) {
      const int nt = (a.nk + 7) / 8, mtiles = (a.nq + 15) / 16, jtiles = (a.nk + 15) / 16, warps = mtiles > jtiles ? mtiles : jtiles;
      const size_t smem = sizeof(float) * 2 * (size_t)(16 * mtiles) * ATT_LD;
#define CALL(N_) \
  do { \
    if (smem > 48 * 1024) cudaFuncSetAttribute(attention_bwd_rows_kernel<N_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    launch_dep(attention_bwd_rows_kernel<N_>, dim3(a.B * a.nh), dim3(32 * warps), smem, (cudaStream_t)stream, a, dout, p->ldo, dq, dk, dv); \
  } while (0)
      B200IQ_ATT_ROWS_DISPATCH(nt, CALL);
#undef CALL
      return launch_status("attention_bwd_rows");
    }
  }
  size_t smem = sizeof(float) * ((size_t)(2 * a.nq + 2 * a.nk) * (a.d + 1) + 2 * (size_t)a.nq * (a.nk + 1));
  B200IQ_CHECK_ARG(smem <= 227 * 1024, "attention_bwd: tile does not fit shared memory (%zu B)", smem);
  if (smem > 48 * 1024) cudaFuncSetAttribute(attention_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  attention_bwd_kernel<<<a.B * a.nh, ATT_THREADS, smem, (cudaStream_t)stream>>>(a, dout, p->ldo, dq, dk, dv);
  return launch_status("attention_bwd");
}
}  // namespace b200iq
using namespace b200iq;

extern "C" int b200iq_attention_fwd(const B200iqAttention* p, void* stream) {
  if (int rc = check_attn(p)) return rc;
  B200IQ_CHECK_ARG(p->out || p->out_hi, "attention_fwd: null out");
  B200IQ_CHECK_ARG(!p->out_lo || p->out_hi, "attention_fwd: out_lo without out_hi");
  const AttOut out{p->out_hi ? nullptr : p->out, (__half*)p->out_hi, (__half*)p->out_lo, p->ldo};
  return attention_fwd_impl(p, out, stream);
}

extern "C" int b200iq_attention_bwd(const B200iqAttention* p, const float* dout, float* dq, float* dk, float* dv,
                                    int lddq, int lddk, int lddv, void* stream) {
  if (int rc = check_attn(p)) return rc;
  B200IQ_CHECK_ARG(dout && dq && dk && dv, "attention_bwd: null pointer");
  B200IQ_CHECK_ARG(p->kv_batch_stride == 0, "attention_bwd: kv_batch_stride is a forward-only (KV cache) option");
  return attention_bwd_impl(p, dout, AttOut{dq, nullptr, nullptr, lddq}, AttOut{dk, nullptr, nullptr, lddk}, AttOut{dv, nullptr, nullptr, lddv}, stream);
}

extern "C" int b200iq_attention_bwd_f16(const B200iqAttention* p, const float* dout, void* dq_hi, void* dk_hi, void* dv_hi,
                                        int lddq, int lddk, int lddv, void* stream) {
  if (int rc = check_attn(p)) return rc;
  B200IQ_CHECK_ARG(dout && dq_hi && dk_hi && dv_hi, "attention_bwd_f16: null pointer");
  B200IQ_CHECK_ARG(p->kv_batch_stride == 0, "attention_bwd_f16: kv_batch_stride is a forward-only (KV cache) option");
  return attention_bwd_impl(p, dout, AttOut{nullptr, (__half*)dq_hi, nullptr, lddq}, AttOut{nullptr, (__half*)dk_hi, nullptr, lddk},
                            AttOut{nullptr, (__half*)dv_hi, nullptr, lddv}, stream);
}
