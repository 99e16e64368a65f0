// Producers of the fp16 operand planes that gemm_f16.cu reads (include/b200iq.h: "fp16-plane GEMM"): the split of the fp32
// master weights, and the plane-writing variants of the HBM-bound row kernels (embedding gather, LayerNorm forward /
// backward, relu backward, column sums, cross-entropy gradient).  A kernel that produces a GEMM operand writes the planes
// itself, so an operand is never re-read just to be converted.
#include "common.cuh"
#include "rowops.cuh"
#include <stdlib.h>

namespace b200iq {

static inline int planes_grid_for(size_t n, int threads) {
  size_t g = (n + threads - 1) / threads;
  size_t cap = (size_t)sm_count() * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// ---------------------------------------------------------------- fp32 -> (hi, lo) planes
// vec: float4 in, 8-byte stores out (cols, both pitches multiples of 4, aligned bases); four loads in flight per thread
__global__ void __launch_bounds__(256)
split_f16_vec_kernel(const float* __restrict__ src, __half* __restrict__ hi, __half* __restrict__ lo, int rows, long long cols4,
                     long long lds, long long ldd, float scale) {
  const size_t total = (size_t)rows * cols4, stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  auto at = [&](size_t idx, size_t& so, size_t& dof) {
    const size_t r = rows == 1 ? 0 : idx / cols4, c = idx - r * cols4;
    so = r * lds + c * 4; dof = r * ldd + c * 4;
  };
  for (; i + 3 * stride < total; i += 4 * stride) {
    float4 v[4]; size_t so[4], dof[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { at(i + u * stride, so[u], dof[u]); v[u] = *reinterpret_cast<const float4*>(src + so[u]); }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      st4_planes(hi, lo, dof[u], make_float4(v[u].x * scale, v[u].y * scale, v[u].z * scale, v[u].w * scale));
  }
  for (; i < total; i += stride) {
    size_t so, dof;
    at(i, so, dof);
    const float4 v = *reinterpret_cast<const float4*>(src + so);
    st4_planes(hi, lo, dof, make_float4(v.x * scale, v.y * scale, v.z * scale, v.w * scale));
  }
}
__global__ void split_f16_kernel(const float* __restrict__ src, __half* __restrict__ hi, __half* __restrict__ lo, int rows,
                                 long long cols, long long lds, long long ldd, float scale) {
  const size_t total = (size_t)rows * cols;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / cols, c = i - r * cols;
    st1_planes(hi, lo, r * ldd + c, src[r * lds + c] * scale);
  }
}

// ---------------------------------------------------------------- embedding gather into planes
__global__ void embed_gather_f16_kernel(const int* __restrict__ ids, const float* __restrict__ table, __half* __restrict__ hi,
                                        __half* __restrict__ lo, int M, int E, int V, int ldo) {
  int m = blockIdx.x * blockDim.y + threadIdx.y;
  if (m >= M) return;
  int id = ids[m];
  id = min(max(id, 0), V - 1);
  const float* src = table + (size_t)id * E;
  if ((E & 3) == 0 && (ldo & 3) == 0) {
    for (int c = threadIdx.x * 4; c < E; c += blockDim.x * 4)
      st4_planes(hi, lo, (size_t)m * ldo + c, __ldg(reinterpret_cast<const float4*>(src + c)));
  } else {
    for (int c = threadIdx.x; c < E; c += blockDim.x) st1_planes(hi, lo, (size_t)m * ldo + c, __ldg(src + c));
  }
}

// ---------------------------------------------------------------- generic LayerNorm forward (any H), warp per row
__global__ void layernorm_fwd_f16_kernel(const float* __restrict__ x, const float* __restrict__ add, float* __restrict__ sum,
                                         const float* __restrict__ g, const float* __restrict__ b, float* __restrict__ y,
                                         __half* __restrict__ y_hi, __half* __restrict__ y_lo, int ldy, float* __restrict__ mean,
                                         float* __restrict__ rstd, int M, int H, float eps) {
  int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const float* xr = x + (size_t)row * H;
  const float* ar = add ? add + (size_t)row * H : nullptr;
  float* sr = sum ? sum + (size_t)row * H : nullptr;
  float s = 0.f;
  for (int c = lane; c < H; c += 32) {
    float v = xr[c];
    if (ar) { v += ar[c]; if (sr) sr[c] = v; }
    s += v;
  }
  __syncwarp();
  const float* vr = (ar && sr) ? sr : xr;   // with add the summed row is re-read from `sum` (required then)
  float mu = warp_sum(s) / H;
  float v = 0.f;
  for (int c = lane; c < H; c += 32) { float d = vr[c] - mu; v += d * d; }
  float rs = rsqrtf(warp_sum(v) / H + eps);
  for (int c = lane; c < H; c += 32) {
    const float o = (vr[c] - mu) * rs * g[c] + b[c];
    if (y) y[(size_t)row * H + c] = o;
    if (y_hi) st1_planes(y_hi, y_lo, (size_t)row * ldy + c, o);
  }
  if (lane == 0) { if (mean) mean[row] = mu; if (rstd) rstd[row] = rs; }
}

// generic LayerNorm backward with the fp16 copy of the resulting dx (see layernorm_bwd_kernel in elementwise.cu)
template <int MAXC>
__global__ void layernorm_bwd_f16_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ g,
                                         const float* __restrict__ mean, const float* __restrict__ rstd, float* __restrict__ dx,
                                         __half* __restrict__ dx_hi, float* __restrict__ dg, float* __restrict__ db, int M, int H,
                                         int rows_per_cta, int accumulate) {
  __shared__ float red[32];
  int r0 = blockIdx.x * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float pg[MAXC], pb[MAXC];
#pragma unroll
  for (int i = 0; i < MAXC; ++i) { pg[i] = 0.f; pb[i] = 0.f; }
  for (int row = r0; row < r1; ++row) {
    const float* dyr = dy + (size_t)row * H;
    const float* xr = x + (size_t)row * H;
    float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      int c = threadIdx.x + i * blockDim.x;
      if (c < H) {
        float xh = (xr[c] - mu) * rs, dyg = dyr[c] * g[c];
        s1 += dyg; s2 += dyg * xh;
        pg[i] += dyr[c] * xh; pb[i] += dyr[c];
      }
    }
    s1 = block_sum(s1, red) / H;
    s2 = block_sum(s2, red) / H;
    float* dxr = dx + (size_t)row * H;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      int c = threadIdx.x + i * blockDim.x;
      if (c < H) {
        float xh = (xr[c] - mu) * rs;
        float v = rs * (dyr[c] * g[c] - s1 - xh * s2);
        v = accumulate ? dxr[c] + v : v;
        dxr[c] = v;
        if (dx_hi) dx_hi[(size_t)row * H + c] = __float2half_rn(v);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < MAXC; ++i) {
    int c = threadIdx.x + i * blockDim.x;
    if (c < H) { atomicAdd(dg + c, pg[i]); atomicAdd(db + c, pb[i]); }
  }
}

// generic relu backward into an fp16 plane (any N)
__global__ void relu_bwd_colsum_f16_kernel(const float* __restrict__ dy, const float* __restrict__ y, __half* __restrict__ dpre_hi,
                                           float* __restrict__ db, int M, int N, int rows_per_cta, float scale) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  int r0 = blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float s = 0.f;
  for (int r = r0; r < r1; ++r) {
    size_t i = (size_t)r * N + c;
    float v = y[i] > 0.f ? dy[i] * scale : 0.f;
    dpre_hi[i] = __float2half_rn(v);
    s += v;
  }
  if (db) atomicAdd(db + c, s);
}

// out[c] += sum_m x[m, c] over an fp16 plane.  block (32, 8): thread (tx, ty) owns 8 columns (one 16-byte load) of rows ty, ty+8, ...
__global__ void __launch_bounds__(256)
colsum_f16_vec_kernel(const __half* __restrict__ x, float* __restrict__ out, int M, int N, int ld, int rows_per_cta) {
  __shared__ float red[8][32][8];
  const int tx = threadIdx.x, ty = threadIdx.y, c = (blockIdx.x * 32 + tx) * 8;
  const int r0 = blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c < N) {
#pragma unroll 4
    for (int r = r0 + ty; r < r1; r += 8) {
      const uint4 q = __ldg(reinterpret_cast<const uint4*>(x + (size_t)r * ld + c));
      const __half2* h = reinterpret_cast<const __half2*>(&q);
#pragma unroll
      for (int j = 0; j < 4; ++j) { const float2 f = __half22float2(h[j]); s[2 * j] += f.x; s[2 * j + 1] += f.y; }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[ty][tx][j] = s[j];
  __syncthreads();
  if (ty == 0 && c < N) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float t = s[j];
#pragma unroll
      for (int w = 1; w < 8; ++w) t += red[w][tx][j];
      atomicAdd(out + c + j, t);
    }
  }
}
__global__ void colsum_f16_kernel(const __half* __restrict__ x, float* __restrict__ out, int M, int N, int ld, int rows_per_cta) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  int r0 = blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float s = 0.f;
  for (int r = r0; r < r1; ++r) s += __half2float(x[(size_t)r * ld + c]);
  atomicAdd(out + c, s);
}

// ---------------------------------------------------------------- cross-entropy gradient as an fp16 plane
// Same math as ce_rows_reg_kernel / ce_rows_kernel (losses.cu); the logits chunk is only read, the gradient rows
// coef * (softmax - onehot) / count go out as halves (2 B instead of 4 B per logit for the two backward GEMMs to read).
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
ce_rows_f16_reg_kernel(const float* __restrict__ logits, int ld, const int* __restrict__ targets, int V, int pad,
                       const float* __restrict__ count, float coef, float* __restrict__ loss_sum, __half* __restrict__ ghi, int ldg) {
  __shared__ float red[32];
  pdl_trigger();
  pdl_wait();
  const int r = blockIdx.x, V4 = V >> 2;
  const float4* row = reinterpret_cast<const float4*>(logits + (size_t)r * ld);
  __half* grow = ghi + (size_t)r * ldg;
  const int tgt = __ldg(targets + r);
  if (tgt == pad) {
    for (int c = threadIdx.x; c < V4; c += THREADS) *reinterpret_cast<uint2*>(grow + 4 * c) = make_uint2(0u, 0u);
    return;
  }
  float4 v[16];
  float mx = -3.0e38f;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int c = threadIdx.x + i * THREADS;
    v[i] = make_float4(-3.0e38f, -3.0e38f, -3.0e38f, -3.0e38f);
    if (c < V4) v[i] = __ldg(row + c);
    mx = fmaxf(fmaxf(mx, fmaxf(v[i].x, v[i].y)), fmaxf(v[i].z, v[i].w));
  }
  mx = block_max(mx, red);
  const float xt = logits[(size_t)r * ld + tgt];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    v[i].x = __expf(v[i].x - mx); v[i].y = __expf(v[i].y - mx); v[i].z = __expf(v[i].z - mx); v[i].w = __expf(v[i].w - mx);
    s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  }
  s = block_sum(s, red);
  const float inv_cnt = 1.f / __ldg(count);
  if (threadIdx.x == 0) atomicAdd(loss_sum, (mx + logf(s) - xt) * inv_cnt);
  const float k = coef * inv_cnt, ks = k / s;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int c = threadIdx.x + i * THREADS;
    if (c < V4) {
      float4 o = make_float4(ks * v[i].x, ks * v[i].y, ks * v[i].z, ks * v[i].w);
      if ((tgt >> 2) == c) { float* oa = &o.x; oa[tgt & 3] -= k; }
      st4_planes(grow, nullptr, (size_t)4 * c, o);
    }
  }
}
__global__ void __launch_bounds__(256)
ce_rows_f16_kernel(const float* __restrict__ logits, int ld, const int* __restrict__ targets, int V, int pad,
                   const float* __restrict__ count, float coef, float* __restrict__ loss_sum, __half* __restrict__ ghi, int ldg) {
  __shared__ float red[32];
  const int r = blockIdx.x;
  const float* row = logits + (size_t)r * ld;
  __half* grow = ghi + (size_t)r * ldg;
  const int tgt = targets[r];
  if (tgt == pad) {
    for (int c = threadIdx.x; c < V; c += blockDim.x) grow[c] = __float2half_rn(0.f);
    return;
  }
  float mx = -3.0e38f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) mx = fmaxf(mx, row[c]);
  mx = block_max(mx, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) s += __expf(row[c] - mx);
  s = block_sum(s, red);
  const float inv_cnt = 1.f / count[0];
  if (threadIdx.x == 0) atomicAdd(loss_sum, (mx + logf(s) - row[tgt]) * inv_cnt);
  const float k = coef * inv_cnt, inv_s = 1.f / s;
  for (int c = threadIdx.x; c < V; c += blockDim.x)
    grow[c] = __float2half_rn(k * (__expf(row[c] - mx) * inv_s - (c == tgt ? 1.f : 0.f)));
}

}  // namespace b200iq
using namespace b200iq;

#define ST ((cudaStream_t)stream)

extern "C" int b200iq_split_f16(const float* src, void* hi, void* lo, int rows, long long cols, long long lds, long long ldd,
                                float scale, void* stream) {
  B200IQ_CHECK_ARG(src && hi && rows > 0 && cols > 0 && lds >= cols && ldd >= cols, "split_f16: bad args");
  const bool vec = (cols & 3) == 0 && (lds & 3) == 0 && (ldd & 3) == 0 && (((uintptr_t)src) & 15) == 0 &&
                   ((((uintptr_t)hi) | ((uintptr_t)lo)) & 7) == 0;
  if (vec) {
    static int per_sm = 0;  // one resident wave (see b200iq_adam_step)
    if (!per_sm) { cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, split_f16_vec_kernel, 256, 0); if (per_sm < 1) per_sm = 1; }
    const size_t n4 = (size_t)rows * (size_t)(cols / 4), want = (n4 + 255) / 256, cap = (size_t)sm_count() * per_sm;
    split_f16_vec_kernel<<<(int)(want < cap ? (want < 1 ? 1 : want) : cap), 256, 0, ST>>>(src, (__half*)hi, (__half*)lo, rows, cols / 4, lds, ldd, scale);
  } else {
    split_f16_kernel<<<planes_grid_for((size_t)rows * cols, 256), 256, 0, ST>>>(src, (__half*)hi, (__half*)lo, rows, cols, lds, ldd, scale);
  }
  return launch_status("split_f16");
}

extern "C" int b200iq_embed_gather_f16(const int* ids, const float* table, void* out_hi, void* out_lo, int M, int E, int V, int ldo,
                                       void* stream) {
  B200IQ_CHECK_ARG(ids && table && out_hi && M > 0 && E > 0 && V > 0 && ldo >= E, "embed_gather_f16: bad args");
  B200IQ_CHECK_ARG((E & 3) != 0 || ((((uintptr_t)table) & 15) == 0 && ((((uintptr_t)out_hi) | ((uintptr_t)out_lo)) & 7) == 0),
                   "embed_gather_f16: table must be 16-byte, planes 8-byte aligned");
  dim3 blk(32, 8);
  embed_gather_f16_kernel<<<ceil_div(M, 8), blk, 0, ST>>>(ids, table, (__half*)out_hi, (__half*)out_lo, M, E, V, ldo);
  return launch_status("embed_gather_f16");
}

extern "C" int b200iq_layernorm_fwd_f16(const float* x, const float* add, float* sum, const float* gamma, const float* beta, float* y,
                                        void* y_hi, void* y_lo, int ldy, float* mean, float* rstd, int M, int H, float eps, void* stream) {
  B200IQ_CHECK_ARG(x && gamma && beta && (y || y_hi) && M > 0 && H > 0, "layernorm_fwd_f16: bad args");
  B200IQ_CHECK_ARG(!add || sum, "layernorm_fwd_f16: add needs the sum output");
  B200IQ_CHECK_ARG(!y_lo || y_hi, "layernorm_fwd_f16: y_lo without y_hi");
  B200IQ_CHECK_ARG(!y_hi || ldy >= H, "layernorm_fwd_f16: ldy too small");
  __half* hi = (__half*)y_hi; __half* lo = (__half*)y_lo;
  const bool al = (((uintptr_t)x | (uintptr_t)y | (uintptr_t)gamma | (uintptr_t)beta | (uintptr_t)add | (uintptr_t)sum) & 15) == 0 &&
                  (((uintptr_t)y_hi | (uintptr_t)y_lo) & 7) == 0 && (ldy & 3) == 0;
  if ((H == 128 || H == 256 || H == 512 || H == 1024) && al) {
    const int grid = min(ceil_div(M, 8), sm_count() * 8);
#define LNF(NV_) do { \
      if (add) launch_dep(ln_fwd_warp_kernel<NV_, true>, dim3(grid), dim3(256), 0, ST, x, add, sum, gamma, beta, y, mean, rstd, M, eps, 0, hi, lo, ldy); \
      else launch_dep(ln_fwd_warp_kernel<NV_, false>, dim3(grid), dim3(256), 0, ST, x, nullptr, nullptr, gamma, beta, y, mean, rstd, M, eps, 0, hi, lo, ldy); \
    } while (0)
    if (H == 128) LNF(1); else if (H == 256) LNF(2); else if (H == 512) LNF(4); else LNF(8);
#undef LNF
    return launch_status("layernorm_fwd_f16");
  }
  layernorm_fwd_f16_kernel<<<ceil_div(M, 8), 256, 0, ST>>>(x, add, sum, gamma, beta, y, hi, lo, ldy, mean, rstd, M, H, eps);
  return launch_status("layernorm_fwd_f16");
}

extern "C" int b200iq_layernorm_bwd_f16(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                                        float* dx, void* dx_hi, float* dgamma, float* dbeta, int M, int H, int accumulate_dx, void* stream) {
  B200IQ_CHECK_ARG(dy && x && gamma && mean && rstd && dx && dgamma && dbeta && M > 0 && H > 0, "layernorm_bwd_f16: bad args");
  B200IQ_CHECK_ARG(H <= 2048, "layernorm_bwd_f16: H=%d > 2048 unsupported", H);
  __half* hi = (__half*)dx_hi;
  if ((H == 128 || H == 256 || H == 512 || H == 1024) && (((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dx | (uintptr_t)gamma) & 15) == 0 &&
      (((uintptr_t)dx_hi) & 7) == 0) {
    static const int per_sm = [] { const char* v = getenv("B200IQ_LNB_CTAS"); return v ? atoi(v) : 2; }();
    const int grid = min(ceil_div(M, 8), sm_count() * per_sm);
    if (H == 128) launch_dep(ln_bwd_warp_kernel<1>, dim3(grid), dim3(256), 0, ST, dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, accumulate_dx, 0, hi);
    else if (H == 256) launch_dep(ln_bwd_warp_kernel<2>, dim3(grid), dim3(256), 0, ST, dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, accumulate_dx, 0, hi);
    else if (H == 512) launch_dep(ln_bwd_warp_kernel<4>, dim3(grid), dim3(256), 0, ST, dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, accumulate_dx, 0, hi);
    else launch_dep(ln_bwd_warp_kernel<8>, dim3(grid), dim3(256), 0, ST, dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, accumulate_dx, 0, hi);
    return launch_status("layernorm_bwd_f16");
  }
  int ctas = min(M, sm_count() * 4);
  int rows = ceil_div(M, ctas);
  ctas = ceil_div(M, rows);
  if (H <= 256) layernorm_bwd_f16_kernel<1><<<ctas, 256, 0, ST>>>(dy, x, gamma, mean, rstd, dx, hi, dgamma, dbeta, M, H, rows, accumulate_dx);
  else if (H <= 512) layernorm_bwd_f16_kernel<2><<<ctas, 256, 0, ST>>>(dy, x, gamma, mean, rstd, dx, hi, dgamma, dbeta, M, H, rows, accumulate_dx);
  else if (H <= 1024) layernorm_bwd_f16_kernel<4><<<ctas, 256, 0, ST>>>(dy, x, gamma, mean, rstd, dx, hi, dgamma, dbeta, M, H, rows, accumulate_dx);
  else layernorm_bwd_f16_kernel<8><<<ctas, 256, 0, ST>>>(dy, x, gamma, mean, rstd, dx, hi, dgamma, dbeta, M, H, rows, accumulate_dx);
  return launch_status("layernorm_bwd_f16");
}

extern "C" int b200iq_relu_bwd_colsum_f16(const float* dy, const float* y, void* dpre_hi, float* dbias, int M, int N, float scale, void* stream) {
  B200IQ_CHECK_ARG(dy && y && dpre_hi && M > 0 && N > 0, "relu_bwd_colsum_f16: bad args");
  if ((N & 3) == 0 && (((uintptr_t)dy | (uintptr_t)y) & 15) == 0 && (((uintptr_t)dpre_hi) & 7) == 0) {
    const int cx = ceil_div(N, 128), rpc = max(32, ceil_div(M, max(1, sm_count() * 4 / cx)));
    launch_dep(colsum4_kernel<true>, dim3(cx, ceil_div(M, rpc)), dim3(32, 8), 0, ST, dy, y, (float*)nullptr, dbias, M, N, N, rpc, scale, 0, (__half*)dpre_hi);
    return launch_status("relu_bwd_colsum_f16");
  }
  int rows = max(8, ceil_div(M, max(1, sm_count() * 4 / ceil_div(N, 128))));
  dim3 grid(ceil_div(N, 128), ceil_div(M, rows));
  relu_bwd_colsum_f16_kernel<<<grid, 128, 0, ST>>>(dy, y, (__half*)dpre_hi, dbias, M, N, rows, scale);
  return launch_status("relu_bwd_colsum_f16");
}

extern "C" int b200iq_colsum_add_f16(const void* x_hi, float* out, int M, int N, int ld, void* stream) {
  B200IQ_CHECK_ARG(x_hi && out && M > 0 && N > 0 && ld >= N, "colsum_add_f16: bad args");
  if ((N & 7) == 0 && (ld & 7) == 0 && (((uintptr_t)x_hi) & 15) == 0) {
    // >= 64 rows per CTA: narrow planes (N = 512: 2 column blocks) otherwise get ~590 CTAs of 35 rows and the per-CTA reduction +
    // 256 atomics dominate (24.3 -> 14.5 us at M = 10240, scripts/gpu_colsum_time.py)
    const int cx = ceil_div(N, 256), rpc = max(64, ceil_div(M, max(1, sm_count() * 4 / cx)));
    colsum_f16_vec_kernel<<<dim3(cx, ceil_div(M, rpc)), dim3(32, 8), 0, ST>>>((const __half*)x_hi, out, M, N, ld, rpc);
    return launch_status("colsum_add_f16");
  }
  int rows = max(8, ceil_div(M, max(1, sm_count() * 4 / ceil_div(N, 128))));
  dim3 grid(ceil_div(N, 128), ceil_div(M, rows));
  colsum_f16_kernel<<<grid, 128, 0, ST>>>((const __half*)x_hi, out, M, N, ld, rows);
  return launch_status("colsum_add_f16");
}

extern "C" int b200iq_ce_rows_f16(const float* logits, int ld, const int* targets, int rows, int V, int pad_id, const float* count,
                                  float coef, float* loss_sum, void* grad_hi, int ldg, void* stream) {
  B200IQ_CHECK_ARG(logits && targets && count && loss_sum && grad_hi && rows > 0 && V > 0 && ld >= V && ldg >= V, "ce_rows_f16: bad args");
  __half* g = (__half*)grad_hi;
  const bool al = (V & 3) == 0 && (ld & 3) == 0 && (ldg & 3) == 0 && (((uintptr_t)logits) & 15) == 0 && (((uintptr_t)grad_hi) & 7) == 0;
  if (al && V <= 16 * 4 * 256) {
    launch_dep(ce_rows_f16_reg_kernel<256>, dim3(rows), dim3(256), 0, ST, logits, ld, targets, V, pad_id, count, coef, loss_sum, g, ldg);
  } else if (al && V <= 16 * 4 * 512) {
    launch_dep(ce_rows_f16_reg_kernel<512>, dim3(rows), dim3(512), 0, ST, logits, ld, targets, V, pad_id, count, coef, loss_sum, g, ldg);
  } else {
    ce_rows_f16_kernel<<<rows, 256, 0, ST>>>(logits, ld, targets, V, pad_id, count, coef, loss_sum, g, ldg);
  }
  return launch_status("ce_rows_f16");
}
