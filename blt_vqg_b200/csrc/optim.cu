// Global-norm gradient clipping + Adam over ONE flat fp32 parameter arena (pure HBM streaming: 16 B read + 12 B written
// per parameter, float4 accesses).   replaces: Lightning gradient_clip_val=5 (train_iq.py:372) + torch.optim.Adam
// (train_iq.py:259-261) with the Noam learning rate of train_iq.py:252-257 supplied through `hyper`.
#include "common.cuh"

namespace b200iq {

__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ g, size_t n, float* __restrict__ out) {
  __shared__ float red[32];
  float acc = 0.f;
  size_t n4 = n >> 2;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 v = g4[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  for (size_t i = (n4 << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc += g[i] * g[i];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(out, acc);
}

// hyper (device): [0]=lr  [1]=1-beta1^t  [2]=1-beta2^t  [3]=grad_scale (1/world_size)
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, size_t n,
            const float* __restrict__ hyper, const float* __restrict__ normsq, float beta1, float beta2, float eps,
            float max_norm, int zero_grad) {
  const float lr = hyper[0], bc1 = hyper[1], bc2 = hyper[2], gs = hyper[3];
  float clip = 1.f;
  if (max_norm > 0.f) clip = fminf(1.f, max_norm / (sqrtf(normsq[0]) * gs + 1e-6f));
  const float gmul = gs * clip, step = lr / bc1, rbc2 = rsqrtf(bc2);
  size_t n4 = n >> 2;
  float4* p4 = reinterpret_cast<float4*>(p);
  float4* g4 = reinterpret_cast<float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 pp = p4[i], gg = g4[i], mm = m4[i], vv = v4[i];
    float* pa = &pp.x; float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float gr = ga[k] * gmul;
      ma[k] = beta1 * ma[k] + (1.f - beta1) * gr;
      va[k] = beta2 * va[k] + (1.f - beta2) * gr * gr;
      pa[k] -= step * ma[k] / (sqrtf(va[k]) * rbc2 + eps);
    }
    p4[i] = pp; m4[i] = mm; v4[i] = vv;
    if (zero_grad) g4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (size_t i = (n4 << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float gr = g[i] * gmul;
    m[i] = beta1 * m[i] + (1.f - beta1) * gr;
    v[i] = beta2 * v[i] + (1.f - beta2) * gr * gr;
    p[i] -= step * m[i] / (sqrtf(v[i]) * rbc2 + eps);
    if (zero_grad) g[i] = 0.f;
  }
}
}  // namespace b200iq
using namespace b200iq;

extern "C" int b200iq_grad_sumsq(const float* grad, size_t n, float* out, void* stream) {
  B200IQ_CHECK_ARG(grad && out && n > 0 && (((uintptr_t)grad) & 15) == 0, "grad_sumsq: bad args");
  cudaMemsetAsync(out, 0, sizeof(float), (cudaStream_t)stream);
  int grid = (int)min((size_t)sm_count() * 8, (n / 4 + 255) / 256 + 1);
  sumsq_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(grad, n, out);
  return launch_status("grad_sumsq");
}

extern "C" int b200iq_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, size_t n, const float* hyper,
                                const float* normsq, float beta1, float beta2, float eps, float max_norm, int zero_grad, void* stream) {
  B200IQ_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && hyper && n > 0, "adam_step: bad args");
  B200IQ_CHECK_ARG(max_norm <= 0.f || normsq, "adam_step: clipping needs normsq");
  B200IQ_CHECK_ARG(((((uintptr_t)param) | ((uintptr_t)grad) | ((uintptr_t)exp_avg) | ((uintptr_t)exp_avg_sq)) & 15) == 0, "adam_step: arenas must be 16-byte aligned");
  int grid = (int)min((size_t)sm_count() * 8, (n / 4 + 255) / 256 + 1);
  adam_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, hyper, normsq, beta1, beta2, eps, max_norm, zero_grad);
  return launch_status("adam_step");
}
