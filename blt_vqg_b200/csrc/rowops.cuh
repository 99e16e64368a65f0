// Vectorised row kernels for the production widths (H % 128 == 0): warp-per-row LayerNorm forward / backward with the
// row held in registers as float4s, and float4 column-sum / relu-backward kernels.  elementwise.cu keeps the generic
// scalar versions for odd widths (the tiny parity configs) and dispatches here when the shape allows.
// Why: ncu (profiles/r01c_*) showed the CTA-per-row-block LayerNorm backward at 15 % of HBM peak (two block-wide
// reductions = 4 __syncthreads per row) and the scalar column sums at 8 %.
#pragma once
#include "common.cuh"
#include <cuda_fp16.h>

namespace b200iq {

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
// fp16 planes of a GEMM operand (gemm_f16.cu): hi = fp16(x), lo = fp16(x - hi); four consecutive elements = one 8-byte store
__device__ __forceinline__ void st4_planes(__half* hi, __half* lo, size_t off, float4 v) {
  const __half2 h0 = __floats2half2_rn(v.x, v.y), h1 = __floats2half2_rn(v.z, v.w);
  uint2 hb;
  hb.x = *reinterpret_cast<const uint32_t*>(&h0); hb.y = *reinterpret_cast<const uint32_t*>(&h1);
  *reinterpret_cast<uint2*>(hi + off) = hb;
  if (lo) {
    const float2 f0 = __half22float2(h0), f1 = __half22float2(h1);
    const __half2 l0 = __floats2half2_rn(v.x - f0.x, v.y - f0.y), l1 = __floats2half2_rn(v.z - f1.x, v.w - f1.y);
    uint2 lb;
    lb.x = *reinterpret_cast<const uint32_t*>(&l0); lb.y = *reinterpret_cast<const uint32_t*>(&l1);
    *reinterpret_cast<uint2*>(lo + off) = lb;
  }
}
__device__ __forceinline__ void st1_planes(__half* hi, __half* lo, size_t off, float v) {
  const __half h = __float2half_rn(v);
  hi[off] = h;
  if (lo) lo[off] = __float2half_rn(v - __half2float(h));
}

// ---------------------------------------------------------------- LayerNorm forward: one warp per row, one HBM read
// ADD: the residual add in front of the LayerNorm rides along — s = rnd(x + add) is written to `sum` (the next block's
// residual input and the LayerNorm backward's x) and normalised in the same pass: one launch and one read of s less than
// axpby + LayerNorm, bit-identical to them.
template <int NV, bool ADD>  // float4s per lane; H = 128 * NV
__global__ void __launch_bounds__(256)
ln_fwd_warp_kernel(const float* __restrict__ x, const float* __restrict__ add, float* __restrict__ sum, const float* __restrict__ g,
                   const float* __restrict__ b, float* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd, int M,
                   float eps, int rn, __half* __restrict__ y_hi, __half* __restrict__ y_lo, int ldy) {
  constexpr int H = 128 * NV;
  const int lane = threadIdx.x & 31, nw = gridDim.x * (blockDim.x >> 5);
  pdl_trigger();
  pdl_wait();
  float4 gv[NV], bv[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) { gv[i] = ldg4(g + (i * 32 + lane) * 4); bv[i] = ldg4(b + (i * 32 + lane) * 4); }
  for (int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < M; row += nw) {
    float4 xv[NV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      xv[i] = ldg4(x + (size_t)row * H + (i * 32 + lane) * 4);
      if (ADD) {
        const float4 av = ldg4(add + (size_t)row * H + (i * 32 + lane) * 4);
        xv[i] = make_float4(rnd(xv[i].x + av.x, rn), rnd(xv[i].y + av.y, rn), rnd(xv[i].z + av.z, rn), rnd(xv[i].w + av.w, rn));
        st4(sum + (size_t)row * H + (i * 32 + lane) * 4, xv[i]);
      }
      s += (xv[i].x + xv[i].y) + (xv[i].z + xv[i].w);
    }
    const float mu = warp_sum(s) * (1.f / H);
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const float a = xv[i].x - mu, c = xv[i].y - mu, d = xv[i].z - mu, e = xv[i].w - mu;
      v += (a * a + c * c) + (d * d + e * e);
    }
    const float rs = rsqrtf(warp_sum(v) * (1.f / H) + eps);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float4 o;
      o.x = rnd((xv[i].x - mu) * rs * gv[i].x + bv[i].x, rn);
      o.y = rnd((xv[i].y - mu) * rs * gv[i].y + bv[i].y, rn);
      o.z = rnd((xv[i].z - mu) * rs * gv[i].z + bv[i].z, rn);
      o.w = rnd((xv[i].w - mu) * rs * gv[i].w + bv[i].w, rn);
      if (y) st4(y + (size_t)row * H + (i * 32 + lane) * 4, o);
      if (y_hi) st4_planes(y_hi, y_lo, (size_t)row * ldy + (i * 32 + lane) * 4, o);
    }
    if (lane == 0) { if (mean) mean[row] = mu; if (rstd) rstd[row] = rs; }
  }
}

// ---------------------------------------------------------------- LayerNorm backward: one warp per row, shuffles only
// dx (+)= rs * (dy*g - mean(dy*g) - xhat * mean(dy*g*xhat));  dg += sum_rows dy*xhat;  db += sum_rows dy.
// Per-lane column partials live in registers across the warp's rows; the 8 warps of a CTA combine through shared
// memory and issue one atomicAdd per column per CTA.
template <int NV>
__global__ void __launch_bounds__(256)
ln_bwd_warp_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ g,
                   const float* __restrict__ mean, const float* __restrict__ rstd, float* __restrict__ dx,
                   float* __restrict__ dg, float* __restrict__ db, int M, int accumulate, int rn, __half* __restrict__ dx_hi) {
  constexpr int H = 128 * NV;
  __shared__ float red[8][H];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = gridDim.x * 8;
  pdl_trigger();
  pdl_wait();
  float4 gv[NV], pg[NV], pb[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    gv[i] = ldg4(g + (i * 32 + lane) * 4);
    pg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    pb[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int row = blockIdx.x * 8 + warp; row < M; row += nw) {
    const float mu = __ldg(mean + row), rs = __ldg(rstd + row);
    float4 dv[NV], xh[NV];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      dv[i] = ldg4(dy + (size_t)row * H + (i * 32 + lane) * 4);
      const float4 xv = ldg4(x + (size_t)row * H + (i * 32 + lane) * 4);
      xh[i] = make_float4((xv.x - mu) * rs, (xv.y - mu) * rs, (xv.z - mu) * rs, (xv.w - mu) * rs);
      const float a = dv[i].x * gv[i].x, c = dv[i].y * gv[i].y, d = dv[i].z * gv[i].z, e = dv[i].w * gv[i].w;
      s1 += (a + c) + (d + e);
      s2 += (a * xh[i].x + c * xh[i].y) + (d * xh[i].z + e * xh[i].w);
      pg[i].x += dv[i].x * xh[i].x; pg[i].y += dv[i].y * xh[i].y; pg[i].z += dv[i].z * xh[i].z; pg[i].w += dv[i].w * xh[i].w;
      pb[i].x += dv[i].x; pb[i].y += dv[i].y; pb[i].z += dv[i].z; pb[i].w += dv[i].w;
    }
    s1 = warp_sum(s1) * (1.f / H);
    s2 = warp_sum(s2) * (1.f / H);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float* dst = dx + (size_t)row * H + (i * 32 + lane) * 4;
      float4 o;
      o.x = rs * (dv[i].x * gv[i].x - s1 - xh[i].x * s2);
      o.y = rs * (dv[i].y * gv[i].y - s1 - xh[i].y * s2);
      o.z = rs * (dv[i].z * gv[i].z - s1 - xh[i].z * s2);
      o.w = rs * (dv[i].w * gv[i].w - s1 - xh[i].w * s2);
      if (accumulate) { const float4 p = *reinterpret_cast<const float4*>(dst); o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w; }
      st4(dst, make_float4(rnd(o.x, rn), rnd(o.y, rn), rnd(o.z, rn), rnd(o.w, rn)));
      if (dx_hi) st4_planes(dx_hi, nullptr, (size_t)row * H + (i * 32 + lane) * 4, o);
    }
  }
#pragma unroll 1
  for (int pass = 0; pass < 2; ++pass) {
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) st4(&red[warp][(i * 32 + lane) * 4], pass == 0 ? pg[i] : pb[i]);
    __syncthreads();
    float* out = pass == 0 ? dg : db;
    for (int c = threadIdx.x; c < H; c += 256) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[w][c];
      atomicAdd(out + c, s);
    }
  }
}

// ---------------------------------------------------------------- column sums (bias gradients), float4 columns
// block (32, 8): thread (tx, ty) owns columns 4*(32*blockIdx.x + tx) .. +3 and rows r0 + ty, +8, ...  A warp reads 512
// contiguous bytes of one row; four rows are in flight per thread.
template <bool RELU_BWD>
__global__ void __launch_bounds__(256)
colsum4_kernel(const float* dy, const float* __restrict__ y, float* dpre, float* __restrict__ out, int M, int N, int ld,
               int rows_per_cta, float scale, int rn, __half* __restrict__ dpre_hi) {
  __shared__ float4 red[8][32];
  const int tx = threadIdx.x, ty = threadIdx.y, c = (blockIdx.x * 32 + tx) * 4;
  const int r0 = blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  pdl_trigger();
  pdl_wait();
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (c < N) {
#pragma unroll 4
    for (int r = r0 + ty; r < r1; r += 8) {
      float4 v = *reinterpret_cast<const float4*>(dy + (size_t)r * ld + c);
      if (RELU_BWD) {
        const float4 yv = ldg4(y + (size_t)r * ld + c);
        const bool sc = rn && scale != 1.f;
        v.x = yv.x > 0.f ? rnd(v.x * scale, sc) : 0.f;
        v.y = yv.y > 0.f ? rnd(v.y * scale, sc) : 0.f;
        v.z = yv.z > 0.f ? rnd(v.z * scale, sc) : 0.f;
        v.w = yv.w > 0.f ? rnd(v.w * scale, sc) : 0.f;
        if (dpre) st4(dpre + (size_t)r * ld + c, v);
        if (dpre_hi) st4_planes(dpre_hi, nullptr, (size_t)r * ld + c, v);
      }
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && c < N && out) {
#pragma unroll
    for (int w = 1; w < 8; ++w) { const float4 o = red[w][tx]; s.x += o.x; s.y += o.y; s.z += o.z; s.w += o.w; }
    atomicAdd(out + c, s.x); atomicAdd(out + c + 1, s.y); atomicAdd(out + c + 2, s.z); atomicAdd(out + c + 3, s.w);
  }
}

}  // namespace b200iq
