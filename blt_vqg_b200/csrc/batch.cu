// Input side of the boundary: the frozen trunk's pooled features live in HBM once per IMAGE
// (bf16 or fp32 rows) and a batch names its images by row index — the per-step host->device traffic of a batch
// shrinks from B*FEAT*4 bytes of features to B*4 bytes of indices.  Both kernels are pure HBM streams: 16-byte
// loads, 16-byte stores, coalesced along the feature dimension, grid capped at a multiple of the SM count.
#include <cuda_bf16.h>
#include "common.cuh"

namespace b200iq {

__device__ __forceinline__ float bf16_bits_to_float(uint32_t b) { return __uint_as_float(b << 16); }

// out[b, :] = rn_tf32(float(cache[idx[b], :])); rows whose index is outside [0, num_images) become zeros and are counted
template <bool BF16>
__global__ void feature_gather_kernel(const void* __restrict__ cache, const int* __restrict__ idx, float* __restrict__ out,
                                      int B, int F, int num_images, int* __restrict__ bad, int rn) {
  constexpr int VEC = BF16 ? 8 : 4;                      // elements per 16-byte load
  const int per_row = F / VEC;
  const size_t total = (size_t)B * per_row;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / per_row), c = (int)(i % per_row) * VEC;
    const int row = __ldg(idx + b);
    float v[VEC];
    if (row < 0 || row >= num_images) {
#pragma unroll
      for (int u = 0; u < VEC; ++u) v[u] = 0.f;
      if (c == 0 && bad) atomicAdd(bad, 1);
    } else if (BF16) {
      const uint4 q = __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(cache) + (size_t)row * F + c));
      const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int u = 0; u < 4; ++u) { v[2 * u] = bf16_bits_to_float(w[u] & 0xFFFFu); v[2 * u + 1] = bf16_bits_to_float(w[u] >> 16); }
    } else {
      const float4 q = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(cache) + (size_t)row * F + c));
      v[0] = rnd(q.x, rn); v[1] = rnd(q.y, rn); v[2] = rnd(q.z, rn); v[3] = rnd(q.w, rn);   // a bf16 value is a TF32 value already
    }
    float4* dst = reinterpret_cast<float4*>(out + (size_t)b * F + c);
#pragma unroll
    for (int u = 0; u < VEC / 4; ++u) dst[u] = make_float4(v[4 * u], v[4 * u + 1], v[4 * u + 2], v[4 * u + 3]);
  }
}

// cache[idx[b], :] = feats[b, :] (round-to-nearest-even to bf16 when the cache is bf16); out-of-range rows are skipped and counted
template <bool BF16>
__global__ void feature_store_kernel(const float* __restrict__ feats, const int* __restrict__ idx, void* __restrict__ cache,
                                     int B, int F, int num_images, int* __restrict__ bad) {
  constexpr int VEC = BF16 ? 8 : 4;
  const int per_row = F / VEC;
  const size_t total = (size_t)B * per_row;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / per_row), c = (int)(i % per_row) * VEC;
    const int row = __ldg(idx + b);
    if (row < 0 || row >= num_images) {
      if (c == 0 && bad) atomicAdd(bad, 1);
      continue;
    }
    const float4* src = reinterpret_cast<const float4*>(feats + (size_t)b * F + c);
    if (BF16) {
      const float4 a = __ldg(src), d = __ldg(src + 1);
      const __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w);
      const __nv_bfloat162 p2 = __floats2bfloat162_rn(d.x, d.y), p3 = __floats2bfloat162_rn(d.z, d.w);
      uint4 q;
      q.x = *reinterpret_cast<const uint32_t*>(&p0); q.y = *reinterpret_cast<const uint32_t*>(&p1);
      q.z = *reinterpret_cast<const uint32_t*>(&p2); q.w = *reinterpret_cast<const uint32_t*>(&p3);
      *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(cache) + (size_t)row * F + c) = q;
    } else {
      *reinterpret_cast<float4*>(static_cast<float*>(cache) + (size_t)row * F + c) = __ldg(src);
    }
  }
}

// dst[i] = (int32) src[i]: the reference's collate_fn makes int64 token tensors (utils/data_loader.py:162-166)
__global__ void ids_narrow_kernel(const long long* __restrict__ src, int* __restrict__ dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = (int)src[i];
}

static inline int stream_grid(size_t items, int threads) {
  size_t g = (items + threads - 1) / threads, cap = (size_t)sm_count() * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace b200iq
using namespace b200iq;

#define ST ((cudaStream_t)stream)

static int check_cache_args(const char* what, const void* cache, const int* idx, const void* dense, int B, int F, int num_images,
                            int cache_bf16) {
  B200IQ_CHECK_ARG(cache && idx && dense && B > 0 && F > 0 && num_images > 0, "%s: bad args", what);
  B200IQ_CHECK_ARG(cache_bf16 == 0 || cache_bf16 == 1, "%s: cache_bf16 must be 0 (fp32 rows) or 1 (bf16 rows)", what);
  B200IQ_CHECK_ARG(F % (cache_bf16 ? 8 : 4) == 0, "%s: feature width %d must be a multiple of %d", what, F, cache_bf16 ? 8 : 4);
  B200IQ_CHECK_ARG(((((uintptr_t)cache) | ((uintptr_t)dense)) & 15) == 0, "%s: pointers must be 16-byte aligned", what);
  return 0;
}

extern "C" int b200iq_feature_gather(const void* cache, int cache_bf16, const int* idx, float* out, int B, int F, int num_images,
                                     int* bad_rows, void* stream) {
  if (int rc = check_cache_args("feature_gather", cache, idx, out, B, F, num_images, cache_bf16)) return rc;
  const size_t items = (size_t)B * (F / (cache_bf16 ? 8 : 4));
  if (cache_bf16) feature_gather_kernel<true><<<stream_grid(items, 256), 256, 0, ST>>>(cache, idx, out, B, F, num_images, bad_rows, round_mode());
  else feature_gather_kernel<false><<<stream_grid(items, 256), 256, 0, ST>>>(cache, idx, out, B, F, num_images, bad_rows, round_mode());
  return launch_status("feature_gather");
}

extern "C" int b200iq_feature_store(const float* feats, const int* idx, void* cache, int cache_bf16, int B, int F, int num_images,
                                    int* bad_rows, void* stream) {
  if (int rc = check_cache_args("feature_store", cache, idx, feats, B, F, num_images, cache_bf16)) return rc;
  const size_t items = (size_t)B * (F / (cache_bf16 ? 8 : 4));
  if (cache_bf16) feature_store_kernel<true><<<stream_grid(items, 256), 256, 0, ST>>>(feats, idx, cache, B, F, num_images, bad_rows);
  else feature_store_kernel<false><<<stream_grid(items, 256), 256, 0, ST>>>(feats, idx, cache, B, F, num_images, bad_rows);
  return launch_status("feature_store");
}

extern "C" int b200iq_ids_narrow(const int64_t* src, int* dst, size_t n, void* stream) {
  B200IQ_CHECK_ARG(src && dst && n > 0, "ids_narrow: bad args");
  ids_narrow_kernel<<<stream_grid(n, 256), 256, 0, ST>>>(reinterpret_cast<const long long*>(src), dst, n);
  return launch_status("ids_narrow");
}
