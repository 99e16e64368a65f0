// Loss-side kernels: fused softmax cross-entropy (forward AND gradient in one pass over a logits chunk that the
// caller keeps L2-resident, so the full (B*T, V) logits tensor never exists in HBM), the repeated-z auxiliary CE,
// the latent tail (reparameterisation + two-Gaussian KL), BatchNorm1d head, MSE.
#include "common.cuh"

namespace b200iq {

// count[0] = number of targets != pad
__global__ void count_nonpad_kernel(const int* __restrict__ t, int n, int pad, float* __restrict__ count) {
  __shared__ float red[32];
  float c = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) c += (t[i] != pad) ? 1.f : 0.f;
  c = block_sum(c, red);
  if (threadIdx.x == 0 && c != 0.f) atomicAdd(count, c);
}

// One CTA per row.  In: logits row; out (in place): coef * (softmax - onehot) / count  (zero row when target == pad).
// loss_sum += (lse - logit[target]) / count.      replaces: train_iq.py:82-83 (F.cross_entropy, ignore_index=PAD)
__global__ void __launch_bounds__(256)
ce_rows_kernel(float* __restrict__ logits, int ld, const int* __restrict__ targets, int V, int pad,
               const float* __restrict__ count, float coef, float* __restrict__ loss_sum, int cache_row, int write_grad, int rn) {
  extern __shared__ float row_s[];
  __shared__ float red[32];
  const int r = blockIdx.x;
  float* row = logits + (size_t)r * ld;
  const int tgt = targets[r];
  if (tgt == pad) {
    if (write_grad) for (int c = threadIdx.x; c < V; c += blockDim.x) row[c] = 0.f;
    return;
  }
  float mx = -3.0e38f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) {
    float v = row[c];
    if (cache_row) row_s[c] = v;
    mx = fmaxf(mx, v);
  }
  mx = block_max(mx, red);
  const float* src = cache_row ? row_s : row;
  float s = 0.f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) s += __expf(src[c] - mx);
  s = block_sum(s, red);
  const float inv_cnt = 1.f / count[0];
  if (threadIdx.x == 0) atomicAdd(loss_sum, (mx + logf(s) - src[tgt]) * inv_cnt);
  if (write_grad) {
    const float k = coef * inv_cnt, inv_s = 1.f / s;
    for (int c = threadIdx.x; c < V; c += blockDim.x) {
      float p = __expf(src[c] - mx) * inv_s;
      row[c] = rnd(k * (p - (c == tgt ? 1.f : 0.f)), rn);
    }
  }
}

// Register-resident variant: the row lives in up to 16 float4 per thread (V <= 16 * 4 * THREADS), so the logits chunk is read
// once and written once with 16-byte accesses and exp() runs once per element.  Same math / same outputs as ce_rows_kernel.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
ce_rows_reg_kernel(float* __restrict__ logits, int ld, const int* __restrict__ targets, int V, int pad,
                   const float* __restrict__ count, float coef, float* __restrict__ loss_sum, int write_grad, int rn) {
  __shared__ float red[32];
  pdl_trigger();
  pdl_wait();
  const int r = blockIdx.x, V4 = V >> 2;
  float4* row = reinterpret_cast<float4*>(logits + (size_t)r * ld);
  const int tgt = __ldg(targets + r);
  if (tgt == pad) {
    if (write_grad) for (int c = threadIdx.x; c < V4; c += THREADS) row[c] = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  float4 v[16];
  float mx = -3.0e38f;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int c = threadIdx.x + i * THREADS;
    v[i] = make_float4(-3.0e38f, -3.0e38f, -3.0e38f, -3.0e38f);
    if (c < V4) v[i] = row[c];
    mx = fmaxf(fmaxf(mx, fmaxf(v[i].x, v[i].y)), fmaxf(v[i].z, v[i].w));
  }
  mx = block_max(mx, red);
  const float xt = logits[(size_t)r * ld + tgt];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    v[i].x = __expf(v[i].x - mx); v[i].y = __expf(v[i].y - mx); v[i].z = __expf(v[i].z - mx); v[i].w = __expf(v[i].w - mx);
    s += (v[i].x + v[i].y) + (v[i].z + v[i].w);   // padding slots hold exp(-huge) = 0
  }
  s = block_sum(s, red);
  const float inv_cnt = 1.f / __ldg(count);
  if (threadIdx.x == 0) atomicAdd(loss_sum, (mx + logf(s) - xt) * inv_cnt);
  if (write_grad) {
    const float k = coef * inv_cnt, ks = k / s;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int c = threadIdx.x + i * THREADS;
      if (c < V4) {
        float4 o = make_float4(ks * v[i].x, ks * v[i].y, ks * v[i].z, ks * v[i].w);
        if ((tgt >> 2) == c) { float* oa = &o.x; oa[tgt & 3] -= k; }
        row[c] = make_float4(rnd(o.x, rn), rnd(o.y, rn), rnd(o.z, rn), rnd(o.w, rn));
      }
    }
  }
}

// Auxiliary loss: CE(repeat(z_logit, T), target) without the repeat.  One CTA per sample b.
// loss_sum += (cnt_b * lse_b - sum_t z_logit[b, tgt_bt]) / count;  dz[b,:] = coef * (cnt_b * softmax - hist(tgt_b)) / count
// replaces: train_iq.py:92-94
__global__ void __launch_bounds__(256)
ce_aux_kernel(const float* __restrict__ zl, float* __restrict__ dz, const int* __restrict__ targets, int T, int V, int pad,
              const float* __restrict__ count, float coef, float* __restrict__ loss_sum, int rn) {
  __shared__ float red[32];
  const int b = blockIdx.x;
  const float* row = zl + (size_t)b * V;
  float* drow = dz + (size_t)b * V;
  float mx = -3.0e38f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) mx = fmaxf(mx, row[c]);
  mx = block_max(mx, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) s += __expf(row[c] - mx);
  s = block_sum(s, red);
  float cnt = 0.f, picked = 0.f;
  for (int t = threadIdx.x; t < T; t += blockDim.x) {
    int tg = targets[b * T + t];
    if (tg != pad) { cnt += 1.f; picked += row[tg]; }
  }
  cnt = block_sum(cnt, red);
  picked = block_sum(picked, red);
  const float inv_cnt = 1.f / count[0];
  if (threadIdx.x == 0) atomicAdd(loss_sum, (cnt * (mx + logf(s)) - picked) * inv_cnt);
  const float k = coef * inv_cnt, ks = k * cnt / s;
  for (int c = threadIdx.x; c < V; c += blockDim.x) drow[c] = ks * __expf(row[c] - mx);
  __syncthreads();
  for (int t = threadIdx.x; t < T; t += blockDim.x) {
    int tg = targets[b * T + t];
    if (tg != pad) atomicAdd(drow + tg, -k);
  }
  if (rn) {
    __syncthreads();
    for (int c = threadIdx.x; c < V; c += blockDim.x) drow[c] = tf32_rn(drow[c]);
  }
}

// ---------------------------------------------------------------- latent tail
// replaces: models/transformer_layers.py:43-58 + gaussian_kld :536-540 (mean over batch at :55)
__global__ void latent_fwd_kernel(const float* __restrict__ prior, const float* __restrict__ post, const float* __restrict__ eps,
                                  float* __restrict__ z, float* __restrict__ kld_sum, int B, int L, int rn) {
  __shared__ float red[32];
  float acc = 0.f;
  int n = B * L;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    int b = i / L, c = i - b * L;
    float mp = prior[b * 2 * L + c], lp = prior[b * 2 * L + L + c];
    if (post) {
      float mq = post[b * 2 * L + c], lq = post[b * 2 * L + L + c];
      z[i] = rnd(eps[i] * expf(0.5f * lq) + mq, rn);
      float dm = mp - mq;
      acc += -0.5f * (1.f + (lq - lp) - dm * dm / expf(lp) - expf(lq) / expf(lp));
    } else {
      z[i] = rnd(eps[i] * expf(0.5f * lp) + mp, rn);
    }
  }
  if (post && kld_sum) {
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) atomicAdd(kld_sum, acc / B);
  }
}

__global__ void latent_bwd_kernel(const float* __restrict__ prior, const float* __restrict__ post, const float* __restrict__ eps,
                                  const float* __restrict__ dz, float kld_coef, const float* __restrict__ kld_coef_dev,
                                  float* __restrict__ dprior, float* __restrict__ dpost, int B, int L, int rn) {
  int n = B * L;
  float kc = (kld_coef_dev ? __ldg(kld_coef_dev) : kld_coef) / B;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    int b = i / L, c = i - b * L;
    float mp = prior[b * 2 * L + c], lp = prior[b * 2 * L + L + c];
    float mq = post[b * 2 * L + c], lq = post[b * 2 * L + L + c];
    float g = dz ? dz[i] : 0.f;
    float ivp = expf(-lp), dm = mp - mq, r = expf(lq - lp);
    dpost[b * 2 * L + c] = rnd(g - kc * dm * ivp, rn);
    dpost[b * 2 * L + L + c] = rnd(g * eps[i] * 0.5f * expf(0.5f * lq) - 0.5f * kc * (1.f - r), rn);
    dprior[b * 2 * L + c] = rnd(kc * dm * ivp, rn);
    dprior[b * 2 * L + L + c] = rnd(-0.5f * kc * (-1.f + dm * dm * ivp + r), rn);
  }
}

// ---------------------------------------------------------------- BatchNorm1d head
// block = (32 columns, 32 row-lanes).   replaces: models/encoder_cnn.py:21,34 (nn.BatchNorm1d, momentum 0.01)
__global__ void __launch_bounds__(1024)
bn_fwd_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ bt, float* __restrict__ y,
              float* __restrict__ save_mean, float* __restrict__ save_rstd, float* __restrict__ run_mean, float* __restrict__ run_var,
              int B, int H, float eps, float momentum, int training) {
  __shared__ float red[32][33];
  int c = blockIdx.x * 32 + threadIdx.x, ty = threadIdx.y;
  float mu, rs;
  if (training) {
    float s = 0.f;
    if (c < H) for (int r = ty; r < B; r += 32) s += x[(size_t)r * H + c];
    red[ty][threadIdx.x] = s;
    __syncthreads();
    s = 0.f;
    for (int k = 0; k < 32; ++k) s += red[k][threadIdx.x];
    mu = s / B;
    __syncthreads();
    float v = 0.f;
    if (c < H) for (int r = ty; r < B; r += 32) { float d = x[(size_t)r * H + c] - mu; v += d * d; }
    red[ty][threadIdx.x] = v;
    __syncthreads();
    v = 0.f;
    for (int k = 0; k < 32; ++k) v += red[k][threadIdx.x];
    float var = v / B;
    rs = rsqrtf(var + eps);
    if (ty == 0 && c < H) {
      save_mean[c] = mu; save_rstd[c] = rs;
      if (run_mean) run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mu;
      if (run_var) run_var[c] = (1.f - momentum) * run_var[c] + momentum * (B > 1 ? v / (B - 1) : var);
    }
  } else {
    mu = (c < H) ? run_mean[c] : 0.f;
    rs = (c < H) ? rsqrtf(run_var[c] + eps) : 0.f;
  }
  if (c < H) {
    float gg = g[c], bb = bt[c];
    for (int r = ty; r < B; r += 32) y[(size_t)r * H + c] = (x[(size_t)r * H + c] - mu) * rs * gg + bb;
  }
}

__global__ void __launch_bounds__(1024)
bn_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ save_mean,
              const float* __restrict__ save_rstd, float* __restrict__ dx, float* __restrict__ dg, float* __restrict__ db, int B, int H, int rn) {
  __shared__ float red1[32][33], red2[32][33];
  int c = blockIdx.x * 32 + threadIdx.x, ty = threadIdx.y;
  float mu = c < H ? save_mean[c] : 0.f, rs = c < H ? save_rstd[c] : 0.f;
  float s1 = 0.f, s2 = 0.f;
  if (c < H) for (int r = ty; r < B; r += 32) {
    float d = dy[(size_t)r * H + c];
    s1 += d; s2 += d * (x[(size_t)r * H + c] - mu) * rs;
  }
  red1[ty][threadIdx.x] = s1; red2[ty][threadIdx.x] = s2;
  __syncthreads();
  s1 = 0.f; s2 = 0.f;
  for (int k = 0; k < 32; ++k) { s1 += red1[k][threadIdx.x]; s2 += red2[k][threadIdx.x]; }
  if (c < H) {
    if (ty == 0) { atomicAdd(dg + c, s2); atomicAdd(db + c, s1); }
    float gg = g[c] * rs, m1 = s1 / B, m2 = s2 / B;
    for (int r = ty; r < B; r += 32) {
      float xh = (x[(size_t)r * H + c] - mu) * rs;
      dx[(size_t)r * H + c] = rnd(gg * (dy[(size_t)r * H + c] - m1 - xh * m2), rn);
    }
  }
}

// ---------------------------------------------------------------- MSE (loss + both gradients)
// loss_sum += mean((a-b)^2);  da += coef*2(a-b)/n;  db (+)= -coef*2(a-b)/n      replaces: train_iq.py:84 (nn.MSELoss)
__global__ void mse_kernel(const float* __restrict__ a, const float* __restrict__ b, size_t n, float coef, float* __restrict__ loss_sum,
                           float* __restrict__ da, float* __restrict__ db, int accumulate_db, int rn) {
  __shared__ float red[32];
  float acc = 0.f, k = 2.f * coef / (float)n;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float d = a[i] - b[i];
    acc += d * d;
    if (da) da[i] += k * d;
    if (db) db[i] = rnd((accumulate_db ? db[i] : 0.f) - k * d, rn);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0 && loss_sum) atomicAdd(loss_sum, acc / (float)n);
}

// ---------------------------------------------------------------- decode-time helpers
// rows of logits -> log-probabilities in place.   replaces: F.log_softmax at models/decoder_rnn.py:107
__global__ void __launch_bounds__(256) log_softmax_rows_kernel(float* __restrict__ x, int ld, int V) {
  __shared__ float red[32];
  float* row = x + (size_t)blockIdx.x * ld;
  float mx = -3.0e38f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) mx = fmaxf(mx, row[c]);
  mx = block_max(mx, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) s += expf(row[c] - mx);
  s = block_sum(s, red);
  const float lse = mx + logf(s);
  for (int c = threadIdx.x; c < V; c += blockDim.x) row[c] -= lse;
}

// per row: the K largest logits (ties -> lowest index, like torch.max / torch.topk on distinct values) and their softmax
// probabilities.   replaces: torch.max + torch.topk(softmax(...), 6) at models/iq.py:136-137 and topk(1) at decoder_rnn.py:124
template <int K>
__global__ void __launch_bounds__(256) topk_softmax_rows_kernel(const float* __restrict__ x, int ld, int V, int* __restrict__ idx_out,
                                                                float* __restrict__ val_out) {
  __shared__ float red[32];
  __shared__ float sv[8];
  __shared__ int si[8];
  __shared__ float bv[32];
  __shared__ int bi[32];
  const float* row = x + (size_t)blockIdx.x * ld;
  float mx = -3.0e38f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) mx = fmaxf(mx, row[c]);
  mx = block_max(mx, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) s += expf(row[c] - mx);
  s = block_sum(s, red);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int k = 0; k < K; ++k) {
    float best = -3.0e38f;
    int besti = 0x7fffffff;
    for (int c = threadIdx.x; c < V; c += blockDim.x) {
      float v = row[c];
      bool taken = false;
      for (int j = 0; j < k; ++j) taken |= (si[j] == c);
      if (!taken && (v > best || (v == best && c < besti))) { best = v; besti = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ov = __shfl_xor_sync(0xffffffffu, best, o);
      int oi = __shfl_xor_sync(0xffffffffu, besti, o);
      if (ov > best || (ov == best && oi < besti)) { best = ov; besti = oi; }
    }
    if (lane == 0) { bv[w] = best; bi[w] = besti; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int j = 1; j < nw; ++j)
        if (bv[j] > bv[0] || (bv[j] == bv[0] && bi[j] < bi[0])) { bv[0] = bv[j]; bi[0] = bi[j]; }
      sv[k] = bv[0]; si[k] = bi[0];
    }
    __syncthreads();
  }
  if (threadIdx.x < K) {
    idx_out[(size_t)blockIdx.x * K + threadIdx.x] = si[threadIdx.x];
    if (val_out) val_out[(size_t)blockIdx.x * K + threadIdx.x] = expf(sv[threadIdx.x] - mx) / s;
  }
}

}  // namespace b200iq
using namespace b200iq;
#define ST ((cudaStream_t)stream)

extern "C" int b200iq_log_softmax_rows(float* x, int ld, int rows, int V, void* stream) {
  B200IQ_CHECK_ARG(x && rows > 0 && V > 0 && ld >= V, "log_softmax_rows: bad args");
  log_softmax_rows_kernel<<<rows, 256, 0, ST>>>(x, ld, V);
  return launch_status("log_softmax_rows");
}

extern "C" int b200iq_topk_softmax_rows(const float* x, int ld, int rows, int V, int k, int* idx_out, float* val_out, void* stream) {
  B200IQ_CHECK_ARG(x && idx_out && rows > 0 && V > 0 && ld >= V && (k == 1 || k == 6) && V >= k, "topk_softmax_rows: bad args (k must be 1 or 6)");
  if (k == 1) topk_softmax_rows_kernel<1><<<rows, 256, 0, ST>>>(x, ld, V, idx_out, val_out);
  else topk_softmax_rows_kernel<6><<<rows, 256, 0, ST>>>(x, ld, V, idx_out, val_out);
  return launch_status("topk_softmax_rows");
}

extern "C" int b200iq_count_nonpad(const int* targets, int n, int pad_id, float* count, void* stream) {
  B200IQ_CHECK_ARG(targets && count && n > 0, "count_nonpad: bad args");
  cudaMemsetAsync(count, 0, sizeof(float), ST);
  count_nonpad_kernel<<<min(ceil_div(n, 256), 64), 256, 0, ST>>>(targets, n, pad_id, count);
  return launch_status("count_nonpad");
}

extern "C" int b200iq_ce_rows(float* logits, int ld, const int* targets, int rows, int V, int pad_id, const float* count,
                              float coef, float* loss_sum, int write_grad, void* stream) {
  B200IQ_CHECK_ARG(logits && targets && count && loss_sum && rows > 0 && V > 0 && ld >= V, "ce_rows: bad args");
  size_t smem = (size_t)V * sizeof(float);
  if ((V & 3) == 0 && (ld & 3) == 0 && (((uintptr_t)logits) & 15) == 0 && V <= 16 * 4 * 512) {
    if (V <= 16 * 4 * 256) launch_dep(ce_rows_reg_kernel<256>, dim3(rows), dim3(256), 0, ST, logits, ld, targets, V, pad_id, count, coef, loss_sum, write_grad, round_mode());
    else launch_dep(ce_rows_reg_kernel<512>, dim3(rows), dim3(512), 0, ST, logits, ld, targets, V, pad_id, count, coef, loss_sum, write_grad, round_mode());
    return launch_status("ce_rows");
  }
  int cache = smem <= 200 * 1024;
  if (!cache) smem = 0;
  static size_t attr = 0;
  if (smem > 48 * 1024 && smem > attr) {
    cudaFuncSetAttribute(ce_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr = 200 * 1024;
  }
  ce_rows_kernel<<<rows, 256, smem, ST>>>(logits, ld, targets, V, pad_id, count, coef, loss_sum, cache, write_grad, round_mode());
  return launch_status("ce_rows");
}

extern "C" int b200iq_ce_aux(const float* z_logit, float* dz_logit, const int* targets, int B, int T, int V, int pad_id,
                             const float* count, float coef, float* loss_sum, void* stream) {
  B200IQ_CHECK_ARG(z_logit && dz_logit && targets && count && loss_sum && B > 0 && T > 0 && V > 0, "ce_aux: bad args");
  ce_aux_kernel<<<B, 256, 0, ST>>>(z_logit, dz_logit, targets, T, V, pad_id, count, coef, loss_sum, round_mode());
  return launch_status("ce_aux");
}

extern "C" int b200iq_latent_fwd(const float* prior, const float* post, const float* eps, float* z, float* kld_sum, int B, int L, void* stream) {
  B200IQ_CHECK_ARG(prior && eps && z && B > 0 && L > 0, "latent_fwd: bad args");
  latent_fwd_kernel<<<min(ceil_div(B * L, 256), 148), 256, 0, ST>>>(prior, post, eps, z, kld_sum, B, L, round_mode());
  return launch_status("latent_fwd");
}

extern "C" int b200iq_latent_bwd(const float* prior, const float* post, const float* eps, const float* dz, float kld_coef,
                                 float* dprior, float* dpost, int B, int L, void* stream) {
  B200IQ_CHECK_ARG(prior && post && eps && dprior && dpost && B > 0 && L > 0, "latent_bwd: bad args");
  latent_bwd_kernel<<<min(ceil_div(B * L, 256), 148), 256, 0, ST>>>(prior, post, eps, dz, kld_coef, nullptr, dprior, dpost, B, L, round_mode());
  return launch_status("latent_bwd");
}

extern "C" int b200iq_latent_bwd_dev(const float* prior, const float* post, const float* eps, const float* dz, const float* kld_coef_dev,
                                     float* dprior, float* dpost, int B, int L, void* stream) {
  B200IQ_CHECK_ARG(prior && post && eps && dprior && dpost && kld_coef_dev && B > 0 && L > 0, "latent_bwd_dev: bad args");
  latent_bwd_kernel<<<min(ceil_div(B * L, 256), 148), 256, 0, ST>>>(prior, post, eps, dz, 0.f, kld_coef_dev, dprior, dpost, B, L, round_mode());
  return launch_status("latent_bwd_dev");
}

extern "C" int b200iq_bn_fwd(const float* x, const float* gamma, const float* beta, float* y, float* save_mean, float* save_rstd,
                             float* running_mean, float* running_var, int B, int H, float eps, float momentum, int training, void* stream) {
  B200IQ_CHECK_ARG(x && gamma && beta && y && B > 0 && H > 0, "bn_fwd: bad args");
  B200IQ_CHECK_ARG(training ? (save_mean && save_rstd) : (running_mean && running_var), "bn_fwd: missing statistics buffers");
  bn_fwd_kernel<<<ceil_div(H, 32), dim3(32, 32), 0, ST>>>(x, gamma, beta, y, save_mean, save_rstd, running_mean, running_var, B, H, eps, momentum, training);
  return launch_status("bn_fwd");
}

extern "C" int b200iq_bn_bwd(const float* dy, const float* x, const float* gamma, const float* save_mean, const float* save_rstd,
                             float* dx, float* dgamma, float* dbeta, int B, int H, void* stream) {
  B200IQ_CHECK_ARG(dy && x && gamma && save_mean && save_rstd && dx && dgamma && dbeta && B > 0 && H > 0, "bn_bwd: bad args");
  bn_bwd_kernel<<<ceil_div(H, 32), dim3(32, 32), 0, ST>>>(dy, x, gamma, save_mean, save_rstd, dx, dgamma, dbeta, B, H, round_mode());
  return launch_status("bn_bwd");
}

extern "C" int b200iq_mse(const float* a, const float* b, size_t n, float coef, float* loss_sum, float* da, float* db,
                          int accumulate_db, void* stream) {
  B200IQ_CHECK_ARG(a && b && n > 0, "mse: bad args");
  mse_kernel<<<(int)min((n + 255) / 256, (size_t)296), 256, 0, ST>>>(a, b, n, coef, loss_sum, da, db, accumulate_db, round_mode());
  return launch_status("mse");
}
