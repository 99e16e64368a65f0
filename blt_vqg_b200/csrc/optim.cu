// Global-norm gradient clipping + Adam over ONE flat fp32 parameter arena (pure HBM streaming: 16 B read + 12 B written
// per parameter, float4 accesses).   replaces: Lightning gradient_clip_val=5 (train_iq.py:372) + torch.optim.Adam
// (train_iq.py:259-261) with the Noam learning rate of train_iq.py:252-257 supplied through `hyper`.
#include "common.cuh"

namespace b200iq {

// Deterministic: every block writes its partial to scratch[1 + blockIdx.x]; the LAST block to finish (ticket in scratch[0],
// a fixed slot so that calls with different grids can share one scratch buffer) adds the partials in index order.  A float atomicAdd tree would make the clip factor differ in the last bit from run to run
// and from rank to rank, and data-parallel replicas would drift apart.
__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ g, size_t n, float* __restrict__ out, float* __restrict__ scratch) {
  __shared__ float red[32];
  __shared__ bool is_last;
  float acc = 0.f;
  size_t n4 = n >> 2;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  // four independent 16-byte loads in flight per thread (one was 2 TB/s: 32 KB in flight per SM does not cover HBM latency)
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n4; i += 4 * stride) {
    const float4 a = g4[i], b = g4[i + stride], c = g4[i + 2 * stride], d = g4[i + 3 * stride];
    acc += (a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w) + (b.x * b.x + b.y * b.y + b.z * b.z + b.w * b.w);
    acc += (c.x * c.x + c.y * c.y + c.z * c.z + c.w * c.w) + (d.x * d.x + d.y * d.y + d.z * d.z + d.w * d.w);
  }
  for (; i < n4; i += stride) {
    float4 v = g4[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  for (size_t i = (n4 << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc += g[i] * g[i];
  acc = block_sum(acc, red);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  scratch += 1;
  if (threadIdx.x == 0) {
    scratch[blockIdx.x] = acc;
    __threadfence();
    is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    float s = 0.f;
    for (unsigned int i = threadIdx.x; i < gridDim.x; i += blockDim.x) s += ((volatile float*)scratch)[i];
    // fixed-shape tree: thread t owns partials t, t+256, ...; block_sum's shuffle order is fixed
    s = block_sum(s, red);
    if (threadIdx.x == 0) { out[0] = s; *ticket = 0u; }
  }
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
// hyper[0..3] = {lr, 1-beta1^t, 1-beta2^t, grad_scale}: host scalars travel as kernel arguments, so there is no pinned
// staging buffer the host could overwrite while an earlier step's copy is still queued
__global__ void set_hyper_kernel(float* hyper, float lr, float bc1, float bc2, float gs) {
  hyper[0] = lr; hyper[1] = bc1; hyper[2] = bc2; hyper[3] = gs;
}
}  // namespace b200iq
using namespace b200iq;

extern "C" int b200iq_set_hyper(float* hyper, float lr, float bias_corr1, float bias_corr2, float grad_scale, void* stream) {
  B200IQ_CHECK_ARG(hyper, "set_hyper: null pointer");
  set_hyper_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(hyper, lr, bias_corr1, bias_corr2, grad_scale);
  return launch_status("set_hyper");
}

extern "C" int b200iq_grad_sumsq(const float* grad, size_t n, float* out, float* scratch, int scratch_floats, void* stream) {
  B200IQ_CHECK_ARG(grad && out && scratch && n > 0 && (((uintptr_t)grad) & 15) == 0, "grad_sumsq: bad args");
  static int per_sm = 0;  // one resident wave (see b200iq_adam_step)
  if (!per_sm) { cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sumsq_kernel, 256, 0); if (per_sm < 1) per_sm = 1; }
  int grid = (int)min((size_t)sm_count() * per_sm, (n / 4 + 255) / 256 + 1);
  if (grid > scratch_floats - 1) grid = scratch_floats - 1;
  B200IQ_CHECK_ARG(grid >= 1, "grad_sumsq: scratch too small");
  sumsq_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(grad, n, out, scratch);
  return launch_status("grad_sumsq");
}

extern "C" int b200iq_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, size_t n, const float* hyper,
                                const float* normsq, float beta1, float beta2, float eps, float max_norm, int zero_grad, void* stream) {
  B200IQ_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && hyper && n > 0, "adam_step: bad args");
  B200IQ_CHECK_ARG(max_norm <= 0.f || normsq, "adam_step: clipping needs normsq");
  B200IQ_CHECK_ARG(((((uintptr_t)param) | ((uintptr_t)grad) | ((uintptr_t)exp_avg) | ((uintptr_t)exp_avg_sq)) & 15) == 0, "adam_step: arenas must be 16-byte aligned");
  // one resident wave: the grid-stride loop deals every CTA of the grid an equal share, so CTAs beyond what fits on the SMs
  // (48 registers -> 5 per SM, not 8) ran as a second, mostly idle wave (433 us for 1.04 GB)
  static int per_sm = 0;
  if (!per_sm) { cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, adam_kernel, 256, 0); if (per_sm < 1) per_sm = 1; }
  int grid = (int)min((size_t)sm_count() * per_sm, (n / 4 + 255) / 256 + 1);
  adam_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, hyper, normsq, beta1, beta2, eps, max_norm, zero_grad);
  return launch_status("adam_step");
}
