// HBM-bound row/elementwise kernels of the IQ path: embedding gather / scatter-add, timing-signal +
// token-0 adds, LayerNorm fwd/bwd, column sums (bias grads), relu backward, Philox dropout / randn.
// All are coalesced along the feature dimension and use float4 accesses where the width allows.
#include "common.cuh"
#include "rowops.cuh"

namespace b200iq {

// ---------------------------------------------------------------- embedding gather
// out[m, :] = table[ids[m], :]            replaces: nn.Embedding forward, models/iq.py:71-78
__global__ void embed_gather_kernel(const int* __restrict__ ids, const float* __restrict__ table,
                                    float* __restrict__ out, int M, int E, int V, int rn) {
  int m = blockIdx.x * blockDim.y + threadIdx.y;
  if (m >= M) return;
  int id = ids[m];
  id = min(max(id, 0), V - 1);
  const float* src = table + (size_t)id * E;
  float* dst = out + (size_t)m * E;
  if ((E & 3) == 0) {
    for (int c = threadIdx.x * 4; c < E; c += blockDim.x * 4) {
      float4 v = __ldg(reinterpret_cast<const float4*>(src + c));
      v.x = rnd(v.x, rn); v.y = rnd(v.y, rn); v.z = rnd(v.z, rn); v.w = rnd(v.w, rn);
      *reinterpret_cast<float4*>(dst + c) = v;
    }
  } else {
    for (int c = threadIdx.x; c < E; c += blockDim.x) dst[c] = rnd(__ldg(src + c), rn);
  }
}

// dtable[ids[m], :] += dout[m, :] unless ids[m] == pad_id   (padding_idx blocks the gradient)
__global__ void embed_scatter_kernel(const int* __restrict__ ids, const float* __restrict__ dout,
                                     float* __restrict__ dtable, int M, int E, int V, int pad_id) {
  int m = blockIdx.x * blockDim.y + threadIdx.y;
  if (m >= M) return;
  int id = ids[m];
  if (id == pad_id || id < 0 || id >= V) return;
  const float* src = dout + (size_t)m * E;
  float* dst = dtable + (size_t)id * E;
  for (int c = threadIdx.x; c < E; c += blockDim.x) atomicAdd(dst + c, src[c]);
}

// x[b, t, :] += timing[t, :] (+ add0[b, :] + add1[b, :] when t == 0)
// replaces: transformer_layers.py:146,214 and encoder/decoder_transformer.py token-0 adds
__global__ void add_timing_kernel(float* __restrict__ x, const float* __restrict__ timing,
                                  const float* __restrict__ add0, const float* __restrict__ add1,
                                  int B, int n, int H, int rn) {
  size_t total = (size_t)B * n * H;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(i % H);
    size_t row = i / H;
    int t = (int)(row % n), b = (int)(row / n);
    // same association order as the reference: (emb + add0 + add1) first (decoder_transformer.py:31-34,46), timing after (:214)
    float v = x[i];
    if (t == 0) {
      if (add0) v += add0[(size_t)b * H + c];
      if (add1) v += add1[(size_t)b * H + c];
    }
    if (timing) v += timing[(size_t)t * H + c];
    x[i] = rnd(v, rn);
  }
}

// dst[b, :] (+)= src[b*n + 0, :]  — gathers the gradient of token 0 of every sequence
__global__ void tok0_gather_kernel(const float* __restrict__ src, float* __restrict__ dst, int B, int n, int H, int ldd, int accumulate, int rn) {
  size_t total = (size_t)B * H;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(i % H), b = (int)(i / H);
    float v = src[(size_t)b * n * H + c];
    size_t o = (size_t)b * ldd + c;
    dst[o] = rnd(accumulate ? dst[o] + v : v, rn);
  }
}

// dst[b*n + 0, :] += src[b, :]
__global__ void tok0_scatter_kernel(const float* __restrict__ src, float* __restrict__ dst, int B, int n, int H, int lds) {
  size_t total = (size_t)B * H;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(i % H), b = (int)(i / H);
    dst[(size_t)b * n * H + c] += src[(size_t)b * lds + c];
  }
}

// ---------------------------------------------------------------- LayerNorm (one warp per row)
// replaces: torch.nn.LayerNorm at transformer_layers.py:134,151,202,220,256-257,320-322
__global__ void layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ b,
                                     float* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd,
                                     int M, int H, float eps, int rn) {
  int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const float* xr = x + (size_t)row * H;
  float s = 0.f;
  for (int c = lane; c < H; c += 32) s += xr[c];
  float mu = warp_sum(s) / H;
  float v = 0.f;
  for (int c = lane; c < H; c += 32) { float d = xr[c] - mu; v += d * d; }
  float rs = rsqrtf(warp_sum(v) / H + eps);
  float* yr = y + (size_t)row * H;
  for (int c = lane; c < H; c += 32) yr[c] = rnd((xr[c] - mu) * rs * g[c] + b[c], rn);
  if (lane == 0) { if (mean) mean[row] = mu; if (rstd) rstd[row] = rs; }
}

// dx (+)= LN-backward(dy); dg += sum dy*xhat; db += sum dy.  Each CTA owns `rows_per_cta` rows and keeps
// per-column partial sums for dg/db in registers (column c handled by the same thread for every row).
template <int MAXC>
__global__ void layernorm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ g,
                                     const float* __restrict__ mean, const float* __restrict__ rstd,
                                     float* __restrict__ dx, float* __restrict__ dg, float* __restrict__ db,
                                     int M, int H, int rows_per_cta, int accumulate, int rn) {
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
        dxr[c] = rnd(accumulate ? dxr[c] + v : v, rn);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < MAXC; ++i) {
    int c = threadIdx.x + i * blockDim.x;
    if (c < H) { atomicAdd(dg + c, pg[i]); atomicAdd(db + c, pb[i]); }
  }
}

// ---------------------------------------------------------------- column sums / relu backward
// out[c] += sum_m x[m, c]   (bias gradients)
__global__ void colsum_kernel(const float* __restrict__ x, float* __restrict__ out, int M, int N, int ld, int rows_per_cta) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  int r0 = blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float s = 0.f;
  for (int r = r0; r < r1; ++r) s += x[(size_t)r * ld + c];
  atomicAdd(out + c, s);
}

// dpre = scale * dy * (y > 0), in place on dy allowed; db[c] += colsum(dpre).  scale = 1/(1-p) folds the backward of the
// dropout that follows the relu: a dropped element has y == 0 and is masked by the same test.
__global__ void relu_bwd_colsum_kernel(const float* dy, const float* __restrict__ y, float* dpre,
                                       float* __restrict__ db, int M, int N, int rows_per_cta, float scale, int rn) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  int r0 = blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float s = 0.f;
  for (int r = r0; r < r1; ++r) {
    size_t i = (size_t)r * N + c;
    float v = y[i] > 0.f ? rnd(dy[i] * scale, rn && scale != 1.f) : 0.f;
    dpre[i] = v;
    s += v;
  }
  if (db) atomicAdd(db + c, s);
}

// ---------------------------------------------------------------- Philox dropout / randn / axpy
// x *= mask / (1-p), mask = dropout_keep (common.cuh: 16 bits of Philox4x32-7(seed; offset + i/8) per element).  Calling
// it again with the same (seed, offset) on the gradient reproduces the mask, so no mask tensor is stored.
__global__ void dropout_kernel(float* __restrict__ x, size_t n, float p, uint64_t seed, uint64_t offset, int rn,
                               const unsigned long long* __restrict__ seed_dev) {
  if (seed_dev) seed += __ldg(seed_dev);
  const float scale = 1.f / (1.f - p);
  const uint32_t thr = dropout_threshold(p);
  size_t n8 = (n + 7) / 8;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n8; q += (size_t)gridDim.x * blockDim.x) {
    uint32_t w[4];
    dropout_words(seed, offset, q * 8, w);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      size_t i = q * 8 + j;
      if (i < n) x[i] = dropout_keep(w, i, thr) ? rnd(x[i] * scale, rn) : 0.f;
    }
  }
}

// out[i] ~ N(0,1): Philox4x32-10(seed; counter = offset + i/4) + Box-Muller (pairs -> cos/sin)
__global__ void randn_kernel(float* __restrict__ out, size_t n, uint64_t seed, uint64_t offset, const unsigned long long* seed_dev) {
  Philox ph(seed + (seed_dev ? __ldg(seed_dev) : 0ull));
  size_t n4 = (n + 3) / 4;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (size_t)gridDim.x * blockDim.x) {
    uint32_t r[4];
    uint64_t ctr = offset + q;
    ph((uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u, r);
    float v[4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float u1 = u32_to_unit(r[2 * h]), u2 = u32_to_unit(r[2 * h + 1]);
      float rad = sqrtf(-2.f * logf(u1)), ang = 6.283185307179586f * u2;
      v[2 * h] = rad * cosf(ang);
      v[2 * h + 1] = rad * sinf(ang);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) if (q * 4 + j < n) out[q * 4 + j] = v[j];
  }
}

// out = a*x + b*y  (out may alias x or y)
__global__ void axpby_kernel(const float* x, const float* y, float* out, size_t n, float a, float b, int rn) {
  pdl_trigger();
  pdl_wait();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = rnd(a * x[i] + (b == 0.f ? 0.f : b * y[i]), rn);
}

// dst = rn_tf32(src), float4-vectorised (n and both pointers 16-byte aligned) with a scalar tail
__global__ void round_tf32_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t n) {
  size_t n4 = n >> 2;
  const float4* s4 = reinterpret_cast<const float4*>(src);
  float4* d4 = reinterpret_cast<float4*>(dst);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n4; i += 4 * stride) {  // four loads in flight per thread: the weight arena streams from HBM
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = s4[i + u * stride];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      v[u].x = tf32_rn(v[u].x); v[u].y = tf32_rn(v[u].y); v[u].z = tf32_rn(v[u].z); v[u].w = tf32_rn(v[u].w);
      d4[i + u * stride] = v[u];
    }
  }
  for (; i < n4; i += stride) {
    float4 v = s4[i];
    v.x = tf32_rn(v.x); v.y = tf32_rn(v.y); v.z = tf32_rn(v.z); v.w = tf32_rn(v.w);
    d4[i] = v;
  }
  for (size_t j = (n4 << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) dst[j] = tf32_rn(src[j]);
}

static inline int grid_for(size_t n, int threads) {
  size_t g = (n + threads - 1) / threads;
  size_t cap = (size_t)sm_count() * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace b200iq
using namespace b200iq;

#define ST ((cudaStream_t)stream)

extern "C" int b200iq_embed_gather(const int* ids, const float* table, float* out, int M, int E, int V, void* stream) {
  B200IQ_CHECK_ARG(ids && table && out && M > 0 && E > 0 && V > 0, "embed_gather: bad args");
  dim3 blk(32, 8);
  embed_gather_kernel<<<ceil_div(M, 8), blk, 0, ST>>>(ids, table, out, M, E, V, round_mode());
  return launch_status("embed_gather");
}

extern "C" int b200iq_embed_scatter_add(const int* ids, const float* dout, float* dtable, int M, int E, int V, int pad_id, void* stream) {
  B200IQ_CHECK_ARG(ids && dout && dtable && M > 0 && E > 0 && V > 0, "embed_scatter_add: bad args");
  dim3 blk(32, 8);
  embed_scatter_kernel<<<ceil_div(M, 8), blk, 0, ST>>>(ids, dout, dtable, M, E, V, pad_id);
  return launch_status("embed_scatter_add");
}

extern "C" int b200iq_add_timing(float* x, const float* timing, const float* add0, const float* add1, int B, int n, int H, void* stream) {
  B200IQ_CHECK_ARG(x && B > 0 && n > 0 && H > 0, "add_timing: bad args");
  add_timing_kernel<<<grid_for((size_t)B * n * H, 256), 256, 0, ST>>>(x, timing, add0, add1, B, n, H, round_mode());
  return launch_status("add_timing");
}

extern "C" int b200iq_tok0_gather(const float* src, float* dst, int B, int n, int H, int ldd, int accumulate, void* stream) {
  B200IQ_CHECK_ARG(src && dst && B > 0 && n > 0 && H > 0 && ldd >= H, "tok0_gather: bad args");
  tok0_gather_kernel<<<grid_for((size_t)B * H, 256), 256, 0, ST>>>(src, dst, B, n, H, ldd, accumulate, round_mode());
  return launch_status("tok0_gather");
}

extern "C" int b200iq_tok0_scatter_add(const float* src, float* dst, int B, int n, int H, int lds, void* stream) {
  B200IQ_CHECK_ARG(src && dst && B > 0 && n > 0 && H > 0 && lds >= H, "tok0_scatter_add: bad args");
  tok0_scatter_kernel<<<grid_for((size_t)B * H, 256), 256, 0, ST>>>(src, dst, B, n, H, lds);
  return launch_status("tok0_scatter_add");
}

extern "C" int b200iq_layernorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* mean, float* rstd,
                                    int M, int H, float eps, void* stream) {
  B200IQ_CHECK_ARG(x && gamma && beta && y && M > 0 && H > 0, "layernorm_fwd: bad args");
  if ((H == 128 || H == 256 || H == 512 || H == 1024) && (((uintptr_t)x | (uintptr_t)y | (uintptr_t)gamma | (uintptr_t)beta) & 15) == 0) {
    const int grid = min(ceil_div(M, 8), sm_count() * 8), rn = round_mode();
    if (H == 128) launch_dep(ln_fwd_warp_kernel<1, false>, dim3(grid), dim3(256), 0, ST, x, nullptr, nullptr, gamma, beta, y, mean, rstd, M, eps, rn, (__half*)nullptr, (__half*)nullptr, 0);
    else if (H == 256) launch_dep(ln_fwd_warp_kernel<2, false>, dim3(grid), dim3(256), 0, ST, x, nullptr, nullptr, gamma, beta, y, mean, rstd, M, eps, rn, (__half*)nullptr, (__half*)nullptr, 0);
    else if (H == 512) launch_dep(ln_fwd_warp_kernel<4, false>, dim3(grid), dim3(256), 0, ST, x, nullptr, nullptr, gamma, beta, y, mean, rstd, M, eps, rn, (__half*)nullptr, (__half*)nullptr, 0);
    else launch_dep(ln_fwd_warp_kernel<8, false>, dim3(grid), dim3(256), 0, ST, x, nullptr, nullptr, gamma, beta, y, mean, rstd, M, eps, rn, (__half*)nullptr, (__half*)nullptr, 0);
    return launch_status("layernorm_fwd");
  }
  layernorm_fwd_kernel<<<ceil_div(M, 8), 256, 0, ST>>>(x, gamma, beta, y, mean, rstd, M, H, eps, round_mode());
  return launch_status("layernorm_fwd");
}

extern "C" int b200iq_add_layernorm_fwd(const float* x, const float* add, float* sum, const float* gamma, const float* beta, float* y,
                                        float* mean, float* rstd, int M, int H, float eps, void* stream) {
  B200IQ_CHECK_ARG(x && add && sum && gamma && beta && y && M > 0 && H > 0, "add_layernorm_fwd: bad args");
  B200IQ_CHECK_ARG(H == 128 || H == 256 || H == 512 || H == 1024, "add_layernorm_fwd: H=%d unsupported (128, 256, 512, 1024)", H);
  B200IQ_CHECK_ARG((((uintptr_t)x | (uintptr_t)add | (uintptr_t)sum | (uintptr_t)y | (uintptr_t)gamma | (uintptr_t)beta) & 15) == 0,
                   "add_layernorm_fwd: pointers must be 16-byte aligned");
  const int grid = min(ceil_div(M, 8), sm_count() * 8), rn = round_mode();
  if (H == 128) launch_dep(ln_fwd_warp_kernel<1, true>, dim3(grid), dim3(256), 0, ST, x, add, sum, gamma, beta, y, mean, rstd, M, eps, rn, (__half*)nullptr, (__half*)nullptr, 0);
  else if (H == 256) launch_dep(ln_fwd_warp_kernel<2, true>, dim3(grid), dim3(256), 0, ST, x, add, sum, gamma, beta, y, mean, rstd, M, eps, rn, (__half*)nullptr, (__half*)nullptr, 0);
  else if (H == 512) launch_dep(ln_fwd_warp_kernel<4, true>, dim3(grid), dim3(256), 0, ST, x, add, sum, gamma, beta, y, mean, rstd, M, eps, rn, (__half*)nullptr, (__half*)nullptr, 0);
  else launch_dep(ln_fwd_warp_kernel<8, true>, dim3(grid), dim3(256), 0, ST, x, add, sum, gamma, beta, y, mean, rstd, M, eps, rn, (__half*)nullptr, (__half*)nullptr, 0);
  return launch_status("add_layernorm_fwd");
}

extern "C" int b200iq_layernorm_bwd(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                                    float* dx, float* dgamma, float* dbeta, int M, int H, int accumulate_dx, void* stream) {
  B200IQ_CHECK_ARG(dy && x && gamma && mean && rstd && dx && dgamma && dbeta && M > 0 && H > 0, "layernorm_bwd: bad args");
  B200IQ_CHECK_ARG(H <= 2048, "layernorm_bwd: H=%d > 2048 unsupported", H);
  if ((H == 128 || H == 256 || H == 512 || H == 1024) && (((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dx | (uintptr_t)gamma) & 15) == 0) {
    const int grid = min(ceil_div(M, 8), sm_count() * 2), rn = round_mode();
    if (H == 128) launch_dep(ln_bwd_warp_kernel<1>, dim3(grid), dim3(256), 0, ST, dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, accumulate_dx, rn, (__half*)nullptr);
    else if (H == 256) launch_dep(ln_bwd_warp_kernel<2>, dim3(grid), dim3(256), 0, ST, dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, accumulate_dx, rn, (__half*)nullptr);
    else if (H == 512) launch_dep(ln_bwd_warp_kernel<4>, dim3(grid), dim3(256), 0, ST, dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, accumulate_dx, rn, (__half*)nullptr);
    else launch_dep(ln_bwd_warp_kernel<8>, dim3(grid), dim3(256), 0, ST, dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, accumulate_dx, rn, (__half*)nullptr);
    return launch_status("layernorm_bwd");
  }
  int ctas = min(M, sm_count() * 4);
  int rows = ceil_div(M, ctas);
  ctas = ceil_div(M, rows);
  if (H <= 256) layernorm_bwd_kernel<1><<<ctas, 256, 0, ST>>>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, H, rows, accumulate_dx, round_mode());
  else if (H <= 512) layernorm_bwd_kernel<2><<<ctas, 256, 0, ST>>>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, H, rows, accumulate_dx, round_mode());
  else if (H <= 1024) layernorm_bwd_kernel<4><<<ctas, 256, 0, ST>>>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, H, rows, accumulate_dx, round_mode());
  else layernorm_bwd_kernel<8><<<ctas, 256, 0, ST>>>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, M, H, rows, accumulate_dx, round_mode());
  return launch_status("layernorm_bwd");
}

extern "C" int b200iq_colsum_add(const float* x, float* out, int M, int N, int ld, void* stream) {
  B200IQ_CHECK_ARG(x && out && M > 0 && N > 0 && ld >= N, "colsum_add: bad args");
  if ((N & 3) == 0 && (ld & 3) == 0 && (((uintptr_t)x) & 15) == 0) {
    const int cx = ceil_div(N, 128), rpc = max(32, ceil_div(M, max(1, sm_count() * 4 / cx)));
    launch_dep(colsum4_kernel<false>, dim3(cx, ceil_div(M, rpc)), dim3(32, 8), 0, ST, x, nullptr, nullptr, out, M, N, ld, rpc, 1.f, 0, (__half*)nullptr);
    return launch_status("colsum_add");
  }
  int rows = max(8, ceil_div(M, max(1, sm_count() * 4 / ceil_div(N, 128))));
  dim3 grid(ceil_div(N, 128), ceil_div(M, rows));
  colsum_kernel<<<grid, 128, 0, ST>>>(x, out, M, N, ld, rows);
  return launch_status("colsum_add");
}

extern "C" int b200iq_relu_bwd_colsum(const float* dy, const float* y, float* dpre, float* dbias, int M, int N, float scale, void* stream) {
  B200IQ_CHECK_ARG(dy && y && dpre && M > 0 && N > 0, "relu_bwd_colsum: bad args");
  if ((N & 3) == 0 && (((uintptr_t)dy | (uintptr_t)y | (uintptr_t)dpre) & 15) == 0) {
    const int cx = ceil_div(N, 128), rpc = max(32, ceil_div(M, max(1, sm_count() * 4 / cx)));
    launch_dep(colsum4_kernel<true>, dim3(cx, ceil_div(M, rpc)), dim3(32, 8), 0, ST, dy, y, dpre, dbias, M, N, N, rpc, scale, round_mode(), (__half*)nullptr);
    return launch_status("relu_bwd_colsum");
  }
  int rows = max(8, ceil_div(M, max(1, sm_count() * 4 / ceil_div(N, 128))));
  dim3 grid(ceil_div(N, 128), ceil_div(M, rows));
  relu_bwd_colsum_kernel<<<grid, 128, 0, ST>>>(dy, y, dpre, dbias, M, N, rows, scale, round_mode());
  return launch_status("relu_bwd_colsum");
}

extern "C" int b200iq_dropout(float* x, size_t n, float p, uint64_t seed, uint64_t offset, void* stream) {
  B200IQ_CHECK_ARG(x && n > 0 && p >= 0.f && p < 1.f, "dropout: bad args");
  if (p == 0.f) return 0;
  dropout_kernel<<<grid_for((n + 7) / 8, 256), 256, 0, ST>>>(x, n, p, seed, offset, round_mode(), nullptr);
  return launch_status("dropout");
}

extern "C" int b200iq_dropout_dev(float* x, size_t n, float p, uint64_t seed, uint64_t offset, const unsigned long long* seed_dev, void* stream) {
  B200IQ_CHECK_ARG(x && n > 0 && p >= 0.f && p < 1.f, "dropout_dev: bad args");
  if (p == 0.f) return 0;
  dropout_kernel<<<grid_for((n + 7) / 8, 256), 256, 0, ST>>>(x, n, p, seed, offset, round_mode(), seed_dev);
  return launch_status("dropout_dev");
}

extern "C" int b200iq_randn(float* out, size_t n, uint64_t seed, uint64_t offset, void* stream) {
  B200IQ_CHECK_ARG(out && n > 0, "randn: bad args");
  randn_kernel<<<grid_for((n + 3) / 4, 256), 256, 0, ST>>>(out, n, seed, offset, nullptr);
  return launch_status("randn");
}

extern "C" int b200iq_randn_dev(float* out, size_t n, uint64_t seed, uint64_t offset, const unsigned long long* seed_dev, void* stream) {
  B200IQ_CHECK_ARG(out && n > 0, "randn_dev: bad args");
  randn_kernel<<<grid_for((n + 3) / 4, 256), 256, 0, ST>>>(out, n, seed, offset, seed_dev);
  return launch_status("randn_dev");
}

extern "C" int b200iq_axpby(const float* x, const float* y, float* out, size_t n, float a, float b, void* stream) {
  B200IQ_CHECK_ARG(x && y && out && n > 0, "axpby: bad args");
  launch_dep(axpby_kernel, dim3(grid_for(n, 256)), dim3(256), 0, ST, x, y, out, n, a, b, round_mode());
  return launch_status("axpby");
}

extern "C" int b200iq_round_tf32(const float* src, float* dst, size_t n, void* stream) {
  B200IQ_CHECK_ARG(src && dst && n > 0, "round_tf32: bad args");
  B200IQ_CHECK_ARG(((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0, "round_tf32: pointers must be 16-byte aligned");
  static int per_sm = 0;  // one resident wave (see b200iq_adam_step)
  if (!per_sm) { cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, round_tf32_kernel, 256, 0); if (per_sm < 1) per_sm = 1; }
  const size_t want = ((n + 3) / 4 + 255) / 256, cap = (size_t)sm_count() * per_sm;
  round_tf32_kernel<<<(int)(want < cap ? (want < 1 ? 1 : want) : cap), 256, 0, ST>>>(src, dst, n);
  return launch_status("round_tf32");
}
