// fp16-operand GEMM (see include/b200iq.h: b200iq_gemm_f16): tcgen05.mma kind::f16, fp32 accumulate in TMEM.
//
// Operands are fp16 PLANES written by the producing kernels: `hi` = fp16(x) and, where the product must be accurate to
// ~2^-22, `lo` = fp16(x - hi).  TERMS = 1: C = A_hi.B_hi (one tensor-core pass: the backward products dX = dY.W and
// dW = dY^T.X, and the vocabulary projection).  TERMS = 3: C = A_hi.B_hi + A_lo.B_hi + A_hi.B_lo (three passes over the same
// smem stage: every forward product that feeds a ReLU gate downstream — profiles/r02a_cpu_precision_study.txt shows that
// 11-bit forward operands flip enough gates to put 1e-2 of error on the gradients, while 11-bit backward operands cost 3e-4).
// Same persistent, warp-specialised structure as gemm.cu (TMA producer warp, one MMA-issuing lane, 8 epilogue warps, two
// TMEM accumulators, CTA pairs with cta_group::2); BK = 64 halves = one 128-byte swizzle row, UMMA K = 16.
// The epilogue can write fp32 (TMA store / reduce-add) and / or fp16 planes (hi, hi + lo) for the next GEMM.
#include "common.cuh"
#include "tcgen.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace b200iq {

static constexpr int FBM = 128;
static constexpr int FBK = 64;          // 64 halves = 128 bytes = one swizzle row
static constexpr int FUMMA_K = 16;      // kind::f16
static constexpr int F16_THREADS = 320; // producer warp + MMA warp + 8 epilogue warps
static constexpr int F16_EPI_WARPS = 8;
static constexpr int F16_STG_BYTES = 4096;  // one 32 x 32 fp32 tile = two 32 x 32 fp16 tiles

struct F16Epi {
  float* C; __half* Chi; __half* Clo; const float* bias; const float* residual; const __half* gate;
  int M, N, ldc, ldh, ldr, ldg, flags; float alpha;
  float drop_p, drop_scale; uint64_t drop_seed, drop_offset; const unsigned long long* drop_seed_dev;
};

__device__ __forceinline__ void umma_f16(uint32_t tmem_c, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_c), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_c, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_c), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}

__device__ __forceinline__ bool half_bits_positive(uint32_t b16) { return (b16 & 0x7FFFu) != 0u && (b16 & 0x8000u) == 0u; }

// generic (slow) epilogue value: every flag, operands from global memory
__device__ __forceinline__ float f16_epi_value(float acc, const F16Epi& e, int row, int col, bool add_bias) {
  float v = acc * e.alpha;
  if ((e.flags & B200IQ_GEMM_BIAS) && add_bias) v += __ldg(e.bias + col);
  if (e.flags & B200IQ_GEMM_RESIDUAL) v += __ldg(e.residual + (size_t)row * e.ldr + col);
  if (e.flags & B200IQ_GEMM_RELU_MASK) v = (__half2float(e.gate[(size_t)row * e.ldg + col]) > 0.f) ? v : 0.f;
  if (e.flags & B200IQ_GEMM_RELU) v = fmaxf(v, 0.f);
  if (e.flags & B200IQ_GEMM_DROPOUT) {
    const size_t i = (size_t)row * e.N + col;
    uint32_t w[4];
    dropout_words(e.drop_seed + (e.drop_seed_dev ? __ldg(e.drop_seed_dev) : 0ull), e.drop_offset, i, w);
    v = dropout_keep(w, i, dropout_threshold(e.drop_p)) ? v * e.drop_scale : 0.f;
  }
  return v;
}

// TERMS: 1 (hi.hi) or 3 (hi.hi + lo.hi + hi.lo).  EPI: 0 lean, 1 residual (fp32 tile by TMA) / relu gate (fp16 plane), 2 dropout.
template <int BN, bool A_MN, bool B_MN, int TERMS, int STAGES, int EPI, int CTAS>
__global__ void __launch_bounds__(F16_THREADS, 1)
gemm_f16_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
                const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl,
                const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmR, F16Epi epi,
                int m_tiles, int n_tiles, int splits, int kblocks_total, int kblocks_per_split, int fast_epi, int gate_map) {
  constexpr int BNL = BN / CTAS;  // rows of B this CTA stages (the pair's MMA reads N/2 from each CTA)
  constexpr int NPL = TERMS == 3 ? 2 : 1;
  constexpr int A_BYTES = FBM * 128, B_BYTES = BNL * 128, STAGE_BYTES = NPL * (A_BYTES + B_BYTES);
  constexpr int NBUF = (TERMS == 3 && STAGE_BYTES > 49152) ? 1 : 2;  // staging buffers per epilogue warp (64 KB stages leave room for one)
  extern __shared__ uint8_t smem_raw[];
  pdl_trigger();
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* stg_base = smem + STAGES * STAGE_BYTES;  // [8 epilogue warps][NBUF][4096 B], 1024-aligned
  __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], tfull_bar[2], tempty_bar[2];
  __shared__ uint64_t gbar[EPI == 1 ? F16_EPI_WARPS : 1][4];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles = m_tiles * n_tiles, total = tiles * splits;
  const int rank = (CTAS == 2) ? (int)cluster_ctarank() : 0;
  const int w_first = blockIdx.x / CTAS, w_step = gridDim.x / CTAS;

  if (warp == 0 && lane == 0) {
    // the descriptors are fetched while the barriers / TMEM are being set up (the first TMA of each map otherwise waits for them)
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmAh) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmBh) : "memory");
    if (TERMS == 3) {
      asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmAl) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmBl) : "memory");
    }
    if (epi.C && fast_epi) asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmC) : "memory");
    if (gate_map) asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmR) : "memory");
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], F16_EPI_WARPS * CTAS); }
    if (EPI == 1) for (int a = 0; a < F16_EPI_WARPS; ++a) for (int k = 0; k < 4; ++k) mbar_init(&gbar[a][k], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    if (CTAS == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(2 * BN) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(2 * BN) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CTAS == 2) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      auto load = [&](void* dst, const CUtensorMap* tm, int c0, int c1, uint64_t* bar) {
        if (CTAS == 2) tma_load_2d_pair(dst, tm, c0, c1, bar); else tma_load_2d(dst, tm, c0, c1, bar);
      };
      // one plane of an operand tile: K-major = one box [rows x 64 halves]; MN-major = boxes of [64 k-rows x 64 halves]
      auto load_a = [&](uint8_t* dst, const CUtensorMap* tm, int m0, int k0, uint64_t* bar) {
        if (!A_MN) load(dst, tm, k0, m0, bar);
        else
          for (int i = 0; i < FBM / 64; ++i) load(dst + i * (FBK * 128), tm, m0 + 64 * i, k0, bar);
      };
      auto load_b = [&](uint8_t* dst, const CUtensorMap* tm, int n0, int k0, uint64_t* bar) {
        if (!B_MN) load(dst, tm, k0, n0, bar);
        else
          for (int i = 0; i < BNL / 64; ++i) load(dst + i * (FBK * 128), tm, n0 + 64 * i, k0, bar);
      };
      for (int w = w_first; w < total; w += w_step) {
        const int z = w / tiles, r = w - z * tiles, mt = r / n_tiles, nt = r - mt * n_tiles;
        const int m0 = (mt * CTAS + rank) * FBM, n0 = nt * BN + rank * BNL;
        const int kb0 = z * kblocks_per_split, kb1 = min(kblocks_total, kb0 + kblocks_per_split);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % STAGES, ph = (it / STAGES) & 1, k0 = kb * FBK;
          mbar_wait(&empty_bar[s], ph ^ 1);
          if (rank == 0) mbar_expect_tx(&full_bar[s], STAGE_BYTES * CTAS);
          uint8_t* sa = smem + s * STAGE_BYTES;
          uint8_t* sb = sa + NPL * A_BYTES;
          load_a(sa, &tmAh, m0, k0, &full_bar[s]);
          load_b(sb, &tmBh, n0, k0, &full_bar[s]);
          if (TERMS == 3) {
            load_a(sa + A_BYTES, &tmAl, m0, k0, &full_bar[s]);
            load_b(sb + B_BYTES, &tmBl, n0, k0, &full_bar[s]);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      // instruction descriptor: D = f32 (1 @4), A = B = f16 (0 @7, 0 @10), majors @15/@16, N>>3 @17, M>>4 @24
      const uint32_t idesc = (1u << 4) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                             ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((FBM * CTAS) >> 4) << 24);
      auto mma = [&](uint32_t tacc, uint64_t ad, uint64_t bd, uint32_t acc) {
        if (CTAS == 2) umma_f16_pair(tacc, ad, bd, idesc, acc); else umma_f16(tacc, ad, bd, idesc, acc);
      };
      // K-major: 128B-swizzled rows of 64 halves, 8-row groups 1024 B apart, a k-step of 16 halves = 32 B along the row.
      // MN-major: 64 mn-halves x 8 k-rows atoms (1024 B), k-groups 1024 B apart, 64-wide mn chunks FBK*128 B apart, a k-step = 2048 B.
      auto adesc = [&](uint32_t base, int k) { return A_MN ? umma_desc(base + k * 2048, FBK * 128, 1024, 2) : umma_desc(base + k * 32, 16, 1024, 2); };
      auto bdesc = [&](uint32_t base, int k) { return B_MN ? umma_desc(base + k * 2048, FBK * 128, 1024, 2) : umma_desc(base + k * 32, 16, 1024, 2); };
      uint32_t it = 0, ti = 0;
      for (int w = w_first; w < total; w += w_step, ++ti) {
        const int z = w / tiles;
        const int kb0 = z * kblocks_per_split, kb1 = min(kblocks_total, kb0 + kblocks_per_split);
        const uint32_t acc = ti & 1, aph = (ti >> 1) & 1;
        mbar_wait(&tempty_bar[acc], aph ^ 1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % STAGES, ph = (it / STAGES) & 1;
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * STAGE_BYTES), sb = sa + NPL * A_BYTES;
#pragma unroll
          for (int k = 0; k < FBK / FUMMA_K; ++k) {
            const uint64_t ah = adesc(sa, k), bh = bdesc(sb, k);
            mma(tacc, ah, bh, (kb > kb0 || k > 0) ? 1u : 0u);
            if (TERMS == 3) {
              mma(tacc, adesc(sa + A_BYTES, k), bh, 1u);
              mma(tacc, ah, bdesc(sb + B_BYTES, k), 1u);
            }
          }
          if (CTAS == 2) umma_commit_pair(&empty_bar[s]); else umma_commit(&empty_bar[s]);
        }
        if (CTAS == 2) umma_commit_pair(&tfull_bar[acc]); else umma_commit(&tfull_bar[acc]);
      }
    }
  } else {
    // warp w may read TMEM lanes 32*(w%4)..+31; two warps share a quadrant and split the BN columns between them
    const int q = warp & 3, chalf = (warp - 2) >> 2;
    constexpr int CHUNKS = BN / 32 / 2;
    uint8_t* stg = stg_base + (warp - 2) * NBUF * F16_STG_BYTES;
    const bool accum = epi.flags & B200IQ_GEMM_ACCUM;
    const bool has_r = EPI == 1 && (epi.flags & B200IQ_GEMM_RESIDUAL);
    const bool gate = EPI == 1 && (epi.flags & B200IQ_GEMM_RELU_MASK), relu = epi.flags & B200IQ_GEMM_RELU;
    const bool drop = EPI == 2;
    const bool drop_pairs = (epi.N & 7) == 0;
    const uint32_t drop_thr = dropout_threshold(epi.drop_p);
    const uint64_t drop_seed = epi.drop_seed + ((drop && epi.drop_seed_dev) ? __ldg(epi.drop_seed_dev) : 0ull);
    uint32_t ti = 0, chunk = 0;
    // EPI 1 + RESIDUAL: the 32 x 32 fp32 residual tile of a chunk is pulled by TMA straight into the chunk's staging buffer
    // (same 128B swizzle as the output tile), one chunk ahead of the math when there are two buffers
    auto fetch_r = [&](int row0, int col0, uint32_t ch) {
      if (lane == 0) {
        bulk_wait_read<NBUF - 1>();   // the store that last read this buffer has finished reading
        uint64_t* gb = &gbar[EPI == 1 ? warp - 2 : 0][ch & 1];
        mbar_expect_tx(gb, F16_STG_BYTES);
        tma_load_2d(stg + (ch % NBUF) * F16_STG_BYTES, &tmR, col0, row0, gb);
      }
    };
    // EPI 1 + RELU_MASK with a plane-only result (the FFN's dX through its first relu): the 32 x 32-half gate tile of a chunk comes
    // by TMA (64-byte swizzle) into one of three 2 KB slots behind the hi-plane transpose area, two chunks ahead.  Read
    // straight from global memory, every lane pulls 64 bytes of a different row: 128 L1 wavefronts per chunk for 2 KB, and with 8
    // epilogue warps that alone takes as long as the tile's MMAs (51 us against 27 us without a gate at M = 10240, N = 2048, K = 512)
    const bool gate_tma = gate && fast_epi && gate_map && NBUF == 2 && epi.C == nullptr && epi.Clo == nullptr;
    auto fetch_g = [&](int row0, int col0, uint32_t ch) {
      if (lane == 0) {
        fence_async_smem();   // the generic reads of this half-buffer (two chunks ago) are ordered before the async write
        uint64_t* gb = &gbar[EPI == 1 ? warp - 2 : 0][ch % 3];
        mbar_expect_tx(gb, 2048);
        tma_load_2d(stg + 2048 * (1 + ch % 3), &tmR, col0, row0, gb);
      }
    };
    for (int w = w_first; w < total; w += w_step, ++ti) {
      const int z = w / tiles, r = w - z * tiles, mt = r / n_tiles, nt = r - mt * n_tiles;
      const int m0 = (mt * CTAS + rank) * FBM, n0 = nt * BN;
      const uint32_t acc = ti & 1, aph = (ti >> 1) & 1;
      const int row0 = m0 + q * 32, row = row0 + lane;
      const bool add_bias = (z == 0);
      if (has_r && fast_epi && row0 < epi.M && n0 + chalf * CHUNKS * 32 < epi.N) fetch_r(row0, n0 + chalf * CHUNKS * 32, chunk);
      if (gate_tma && row0 < epi.M) {
        const int c0 = n0 + chalf * CHUNKS * 32;
        if (c0 < epi.N) fetch_g(row0, c0, chunk);
        if (CHUNKS > 1 && c0 + 32 < epi.N) fetch_g(row0, c0 + 32, chunk + 1);
      }
      mbar_wait(&tfull_bar[acc], aph);
      tc_fence_after();
#pragma unroll 1
      for (int c = chalf * CHUNKS; c < (chalf + 1) * CHUNKS; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= epi.N) break;  // warp-uniform
        const bool bias_on = (epi.flags & B200IQ_GEMM_BIAS) && add_bias;
        float4 bias4[8];
        if (bias_on) {
          if (col0 + 32 <= epi.N && ((((uintptr_t)(epi.bias + col0)) & 15) == 0)) {
            const float4* b4 = reinterpret_cast<const float4*>(epi.bias + col0);
#pragma unroll
            for (int j = 0; j < 8; ++j) bias4[j] = __ldg(b4 + j);
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float* b1 = &bias4[j].x;
#pragma unroll
              for (int t = 0; t < 4; ++t) b1[t] = (col0 + 4 * j + t < epi.N) ? __ldg(epi.bias + col0 + 4 * j + t) : 0.f;
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) bias4[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        // relu gate: this lane's 32 halves of the gate plane (fast path: N % 8 == 0, ldg % 8 == 0, 16-byte aligned base)
        uint4 gq[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) gq[j] = make_uint4(0u, 0u, 0u, 0u);
        if (gate && fast_epi && !gate_tma && row < epi.M) {
          const uint4* gp = reinterpret_cast<const uint4*>(epi.gate + (size_t)row * epi.ldg + col0);
#pragma unroll
          for (int j = 0; j < 4; ++j) if (col0 + 8 * j < epi.N) gq[j] = __ldg(gp + j);
        }
        uint32_t rr[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + c * 32, rr);
        if (row0 >= epi.M) continue;  // warp-uniform: nothing of this quadrant is inside the matrix
        if (gate_tma) {
          mbar_wait(&gbar[EPI == 1 ? warp - 2 : 0][chunk % 3], (chunk / 3) & 1);   // this chunk's gate tile has landed
          const uint4* gp = reinterpret_cast<const uint4*>(stg + 2048 * (1 + chunk % 3) + lane * 64);
          const int sw = (lane >> 1) & 3;   // 64-byte swizzle: 16-byte piece j of row r sits at piece j ^ ((r >> 1) & 3)
#pragma unroll
          for (int j = 0; j < 4; ++j) gq[j] = gp[j ^ sw];
          __syncwarp();
          if (c + 2 < (chalf + 1) * CHUNKS && col0 + 64 < epi.N) fetch_g(row0, col0 + 64, chunk + 2);   // slot of chunk - 1: read
        }
        if (fast_epi) {
          uint8_t* buf = gate_tma ? stg : stg + (chunk % NBUF) * F16_STG_BYTES;
          if (has_r) {
            mbar_wait(&gbar[EPI == 1 ? warp - 2 : 0][chunk & 1], (chunk >> 1) & 1);  // the residual tile has landed
          } else {
            if (lane == 0) bulk_wait_read<NBUF - 1>();
            __syncwarp();
          }
          uint32_t keep[4] = {0u, 0u, 0u, 0u};
          const uint32_t* gw = &gq[0].x;
          const bool gate_packed = gate && epi.C == nullptr && epi.Clo == nullptr;
          // ---- epilogue math, in place in rr[]
          // (a plain product — no bias, residual, gate on fp32 values, relu, dropout, alpha — leaves the accumulators as they are)
          const bool plain = !bias_on && !has_r && !(gate && !gate_packed) && !relu && !drop && epi.alpha == 1.f;
          if (!plain)
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 bb = bias4[j >> 2];
            float4 gg = make_float4(0.f, 0.f, 0.f, 0.f);
            if (has_r) gg = *reinterpret_cast<const float4*>(buf + lane * 128 + (((j >> 2) ^ (lane & 7)) << 4));
            const size_t ei = (size_t)row * epi.N + col0 + j;
            if (drop && (!drop_pairs || !(j & 4))) dropout_words(drop_seed, epi.drop_offset, ei, keep);
            const float* ga = &gg.x; const float* ba = &bb.x;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              float x = __uint_as_float(rr[j + t]) * epi.alpha + ba[t] + ga[t];
              if (gate && !gate_packed) {
                const uint32_t hw = gw[(j + t) >> 1];
                x = half_bits_positive((j + t) & 1 ? (hw >> 16) : (hw & 0xFFFFu)) ? x : 0.f;
              }
              if (relu) x = fmaxf(x, 0.f);
              if (drop) x = dropout_keep(keep, ei + t, drop_thr) ? x * epi.drop_scale : 0.f;
              rr[j + t] = __float_as_uint(x);
            }
          }
          // ---- fp16 planes: pack, transpose through the staging buffer, 16-byte stores (8 rows x 64 B per instruction)
          if (epi.Chi) {
            if (has_r) __syncwarp();  // every lane has read its residual row before the buffer is reused
            uint32_t hp[16], lp[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float a = __uint_as_float(rr[2 * j]), b = __uint_as_float(rr[2 * j + 1]);
              const __half2 h = __floats2half2_rn(a, b);
              hp[j] = *reinterpret_cast<const uint32_t*>(&h);
              // hi-plane-only result: the gate is applied to the packed pair (one compare + one AND per two elements instead of
              // five instructions per element on the fp32 values; these epilogues are instruction-bound)
              if (gate_packed) hp[j] &= __hgt2_mask(*reinterpret_cast<const __half2*>(&gw[j]), __float2half2_rn(0.f));
              lp[j] = 0u;
              if (epi.Clo) {   // warp-uniform: hi-only planes (every backward operand) skip the remainder's five instructions per pair
                const float2 hf = __half22float2(h);
                const __half2 l = __floats2half2_rn(a - hf.x, b - hf.y);
                lp[j] = *reinterpret_cast<const uint32_t*>(&l);
              }
            }
            const int sw = (lane >> 1) & 3;
#pragma unroll
            for (int c4 = 0; c4 < 4; ++c4) {
              *reinterpret_cast<uint4*>(buf + lane * 64 + ((c4 ^ sw) << 4)) = make_uint4(hp[4 * c4], hp[4 * c4 + 1], hp[4 * c4 + 2], hp[4 * c4 + 3]);
              if (epi.Clo)
                *reinterpret_cast<uint4*>(buf + 2048 + lane * 64 + ((c4 ^ sw) << 4)) = make_uint4(lp[4 * c4], lp[4 * c4 + 1], lp[4 * c4 + 2], lp[4 * c4 + 3]);
            }
            __syncwarp();
            const int c4 = lane & 3, gcol = col0 + c4 * 8;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int rl = 8 * i + (lane >> 2), grow = row0 + rl;
              if (grow < epi.M && gcol < epi.N) {
                const int so = rl * 64 + ((c4 ^ ((rl >> 1) & 3)) << 4);
                *reinterpret_cast<uint4*>(epi.Chi + (size_t)grow * epi.ldh + gcol) = *reinterpret_cast<const uint4*>(buf + so);
                if (epi.Clo)
                  *reinterpret_cast<uint4*>(epi.Clo + (size_t)grow * epi.ldh + gcol) = *reinterpret_cast<const uint4*>(buf + 2048 + so);
              }
            }
            __syncwarp();
          }
          // ---- fp32 C: 128B-swizzled staging tile -> TMA store (or reduce-add for split-K)
          if (epi.C) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              *reinterpret_cast<uint4*>(buf + lane * 128 + (((j >> 2) ^ (lane & 7)) << 4)) = make_uint4(rr[j], rr[j + 1], rr[j + 2], rr[j + 3]);
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (accum) tma_reduce_add_2d(&tmC, buf, col0, row0); else tma_store_2d(&tmC, buf, col0, row0);
              bulk_commit();
            }
          }
          ++chunk;
          if (has_r && c + 1 < (chalf + 1) * CHUNKS && col0 + 32 < epi.N) fetch_r(row0, col0 + 32, chunk);
        } else {
          // generic path (odd pitches / widths, the tiny parity configs): one row per lane, scalar loads and stores
          float* mine = reinterpret_cast<float*>(stg + lane * 128);
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<uint4*>(mine + j) = make_uint4(rr[j], rr[j + 1], rr[j + 2], rr[j + 3]);
          if (row < epi.M) {
#pragma unroll 1
            for (int j = 0; j < 32; ++j)
              if (col0 + j < epi.N) {
                const float v = f16_epi_value(mine[j], epi, row, col0 + j, add_bias);
                if (epi.C) {
                  float* dst = epi.C + (size_t)row * epi.ldc + col0 + j;
                  if (accum) atomicAdd(dst, v); else *dst = v;
                }
                if (epi.Chi) {
                  const __half h = __float2half_rn(v);
                  epi.Chi[(size_t)row * epi.ldh + col0 + j] = h;
                  if (epi.Clo) epi.Clo[(size_t)row * epi.ldh + col0 + j] = __float2half_rn(v - __half2float(h));
                }
              }
          }
          __syncwarp();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (CTAS == 2) mbar_arrive_leader(&tempty_bar[acc]); else mbar_arrive(&tempty_bar[acc]); }
    }
    // the staging buffers must outlive the TMA engine's READ of them; the writes themselves complete with the grid (what CUTLASS's
    // tma_store_wait does too) — waiting for full completion added the global write latency to every CTA's tail
    if (lane == 0) bulk_wait_read<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (CTAS == 2) cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    if (CTAS == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(2 * BN) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(2 * BN) : "memory");
  }
}

// ------------------------------------------------------------------ host side
// 2-D fp16 tensor map over a row-major [rows, cols] matrix with leading dimension ld (halves); box = [box_rows, 64 cols]
static int make_map_f16(CUtensorMap* tm, const void* base, int rows, int cols, int ld, int box_rows) {
  EncodeTiledFn enc = get_encode();
  B200IQ_CHECK_ARG(enc, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void*)base, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200IQ_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (fp16) failed (%d) rows=%d cols=%d ld=%d", (int)r, rows, cols, ld);
  return 0;
}
static int make_map_f32_tile(CUtensorMap* tm, const float* base, int rows, int cols, int ld) {
  EncodeTiledFn enc = get_encode();
  B200IQ_CHECK_ARG(enc, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200IQ_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (fp32 tile) failed (%d) rows=%d cols=%d ld=%d", (int)r, rows, cols, ld);
  return 0;
}

// 32 rows x 32 halves of an fp16 plane, 64-byte swizzle (the relu gate tile of one epilogue chunk)
static int make_map_f16_gate(CUtensorMap* tm, const void* base, int rows, int cols, int ld) {
  EncodeTiledFn enc = get_encode();
  B200IQ_CHECK_ARG(enc, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void*)base, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200IQ_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (gate tile) failed (%d) rows=%d cols=%d ld=%d", (int)r, rows, cols, ld);
  return 0;
}

struct F16Maps { CUtensorMap ah, al, bh, bl, c, r; };

template <int BN, bool A_MN, bool B_MN, int TERMS, int EPI, int CTAS>
static int launch_f16(const F16Maps& m, int fast_epi, int gate_map, const F16Epi& e, int K, int split, cudaStream_t st) {
  constexpr int NPL = TERMS == 3 ? 2 : 1;
  constexpr int STAGE = NPL * (FBM * 128 + (BN / CTAS) * 128);
  constexpr int NBUF = (TERMS == 3 && STAGE > 49152) ? 1 : 2;
  constexpr int ROOM = 232448 - 512 - 1024 - F16_EPI_WARPS * NBUF * F16_STG_BYTES;
  constexpr int STAGES = ROOM / STAGE > 6 ? 6 : ROOM / STAGE;
  static_assert(STAGES >= 3, "pipeline depth");
  constexpr int SMEM = STAGES * STAGE + F16_EPI_WARPS * NBUF * F16_STG_BYTES + 1024;
  static_assert(SMEM + 512 <= 232448, "shared memory budget");
  auto kern = gemm_f16_kernel<BN, A_MN, B_MN, TERMS, STAGES, EPI, CTAS>;
  static int attr_dev = -1;  // the opt-in is per device: redo it when the calling thread's device changes
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    attr_dev = dev;
  }
  int kblocks = ceil_div(K, FBK);
  if (split > kblocks) split = kblocks;
  int per = ceil_div(kblocks, split);
  split = ceil_div(kblocks, per);
  const int m_tiles = ceil_div(e.M, FBM * CTAS), n_tiles = ceil_div(e.N, BN);
  const int total = m_tiles * n_tiles * split;
  const int slots = sm_count() / CTAS;
  const int grid = (total < slots ? total : slots) * CTAS;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(F16_THREADS); cfg.dynamicSmemBytes = SMEM; cfg.stream = st;
  cudaLaunchAttribute at[2];
  int na = 0;
  if (CTAS == 2) {
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = 2; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl_mode()) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = at; cfg.numAttrs = na;
  cudaError_t err = cudaLaunchKernelEx(&cfg, kern, m.ah, m.al, m.bh, m.bl, m.c, m.r, e, m_tiles, n_tiles, split, kblocks, per, fast_epi, gate_map);
  if (err != cudaSuccess) { set_error("gemm_f16_kernel: %s", cudaGetErrorString(err)); return (int)err; }
  return launch_status("gemm_f16_kernel");
}

}  // namespace b200iq

using namespace b200iq;

extern "C" int b200iq_gemm_f16(const B200iqGemmF16* p, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B200IQ_CHECK_ARG(p && p->A_hi && p->B_hi && (p->C || p->C_hi), "gemm_f16: null pointer");
  B200IQ_CHECK_ARG(p->M > 0 && p->N > 0 && p->K > 0, "gemm_f16: bad shape M=%d N=%d K=%d", p->M, p->N, p->K);
  const int f = p->flags;
  const bool amn = f & B200IQ_GEMM_A_MN, bmn = f & B200IQ_GEMM_B_MN;
  const int terms = (p->A_lo && p->B_lo) ? 3 : 1;
  B200IQ_CHECK_ARG((p->A_lo != nullptr) == (p->B_lo != nullptr), "gemm_f16: give both lo planes (3 terms) or neither (1 term)");
  B200IQ_CHECK_ARG(!(f & (B200IQ_GEMM_SIMT | B200IQ_GEMM_FP32_OUT)), "gemm_f16: SIMT / FP32_OUT are flags of b200iq_gemm");
  B200IQ_CHECK_ARG(!(f & B200IQ_GEMM_BIAS) || p->bias, "gemm_f16: BIAS without bias pointer");
  B200IQ_CHECK_ARG(!(f & B200IQ_GEMM_RESIDUAL) || (p->residual && p->C), "gemm_f16: RESIDUAL needs the residual pointer and an fp32 C");
  B200IQ_CHECK_ARG(!(f & B200IQ_GEMM_RELU_MASK) || p->gate, "gemm_f16: RELU_MASK needs the fp16 gate plane");
  B200IQ_CHECK_ARG(!((f & B200IQ_GEMM_RESIDUAL) && (f & B200IQ_GEMM_RELU_MASK)), "gemm_f16: RESIDUAL and RELU_MASK are exclusive");
  B200IQ_CHECK_ARG(!p->C_lo || p->C_hi, "gemm_f16: C_lo without C_hi");
  int split = p->split_k < 1 ? 1 : p->split_k;
  B200IQ_CHECK_ARG(split == 1 || ((f & B200IQ_GEMM_ACCUM) && !p->C_hi && !(f & (B200IQ_GEMM_RELU | B200IQ_GEMM_RELU_MASK | B200IQ_GEMM_RESIDUAL))),
                   "gemm_f16: split_k>1 needs ACCUM, an fp32 C only and a linear epilogue");
  B200IQ_CHECK_ARG(!(f & B200IQ_GEMM_ACCUM) || (p->C && !p->C_hi), "gemm_f16: ACCUM writes fp32 C only");
  B200IQ_CHECK_ARG(p->lda >= (amn ? p->M : p->K) && p->ldb >= (bmn ? p->N : p->K), "gemm_f16: leading dimension too small");
  B200IQ_CHECK_ARG((p->lda & 7) == 0 && (p->ldb & 7) == 0, "gemm_f16: lda/ldb must be multiples of 8 halves (lda=%d ldb=%d)", p->lda, p->ldb);
  B200IQ_CHECK_ARG((((uintptr_t)p->A_hi | (uintptr_t)p->B_hi | (uintptr_t)p->A_lo | (uintptr_t)p->B_lo) & 15) == 0, "gemm_f16: operand planes must be 16-byte aligned (A_hi %p A_lo %p B_hi %p B_lo %p; M=%d N=%d K=%d)", p->A_hi, p->A_lo, p->B_hi, p->B_lo, p->M, p->N, p->K);
  B200IQ_CHECK_ARG(terms == 1 || (!amn && !bmn), "gemm_f16: the 3-term product takes K-major A and B");
  F16Epi e{p->C, (__half*)p->C_hi, (__half*)p->C_lo, p->bias, p->residual, (const __half*)p->gate,
           p->M, p->N, p->ldc, p->ldh, p->ldr, p->ldg, f, p->alpha, 0.f, 1.f, 0ull, 0ull, nullptr};
  if (f & B200IQ_GEMM_DROPOUT) {
    B200IQ_CHECK_ARG(p->drop_p > 0.f && p->drop_p < 1.f, "gemm_f16: DROPOUT needs 0 < drop_p < 1");
    B200IQ_CHECK_ARG((p->N & 3) == 0 && split == 1 && !(f & B200IQ_GEMM_ACCUM) && !amn && !bmn && !(f & (B200IQ_GEMM_RESIDUAL | B200IQ_GEMM_RELU_MASK)),
                     "gemm_f16: DROPOUT needs N %% 4 == 0, K-major operands, no split-K / ACCUM / residual / gate");
    e.drop_p = p->drop_p; e.drop_scale = 1.f / (1.f - p->drop_p); e.drop_seed = p->drop_seed; e.drop_offset = p->drop_offset;
    e.drop_seed_dev = p->drop_seed_dev;
  }
  // the fast epilogue needs TMA-storable fp32 tiles and 16-byte plane rows; anything else takes the generic per-lane path
  bool fast = true;
  if (p->C) fast = fast && p->ldc >= p->N && (p->ldc & 3) == 0 && (((uintptr_t)p->C) & 15) == 0;
  if (f & B200IQ_GEMM_RESIDUAL) fast = fast && (p->ldr & 3) == 0 && (((uintptr_t)p->residual) & 15) == 0;
  if (p->C_hi) fast = fast && (p->N & 7) == 0 && (p->ldh & 7) == 0 && p->ldh >= p->N && ((((uintptr_t)p->C_hi) | ((uintptr_t)p->C_lo)) & 15) == 0;
  if (f & B200IQ_GEMM_RELU_MASK) fast = fast && (p->N & 7) == 0 && (p->ldg & 7) == 0 && (((uintptr_t)p->gate) & 15) == 0;
  if ((f & B200IQ_GEMM_DROPOUT) && !fast) { set_error("gemm_f16: DROPOUT needs the aligned fast epilogue"); return -1; }
  B200IQ_CHECK_ARG(fast || !p->C || p->ldc >= p->N, "gemm_f16: ldc too small");

  // tile policy.  CTA pairs (cta_group::2, 256 rows per pair) wherever M > 128; the pair's N is 256, or 128 when that fills the
  // 74 pairs better: a persistent grid finishes in ceil(tiles / pairs) rounds, and e.g. M = 10240, N = 512 is 80 tiles of
  // 256 x 256 (2 rounds, the second with 6 pairs busy) but 160 tiles of 256 x 128 (3 rounds of half the work each)
  static const int force_bn = [] { const char* v = getenv("B200IQ_F16_TILE_N"); return v ? atoi(v) : 0; }();
  const bool pair = p->M > 128 && p->N > 64;
  int pair_bn = 256;
  if (pair) {
    const int slots = sm_count() / 2 > 0 ? sm_count() / 2 : 1, mt = ceil_div(p->M, 256);
    int kb = ceil_div(p->K, FBK), sp = split > kb ? kb : split;
    const long t256 = (long)mt * ceil_div(p->N, 256) * sp, t128 = (long)mt * ceil_div(p->N, 128) * sp;
    const double c256 = (double)((t256 + slots - 1) / slots) * 2.0, c128 = (double)((t128 + slots - 1) / slots) * 1.08;
    // measured (profiles/r02f_tile_policy.txt): the narrower tile re-reads A twice as often and only pays off on paper — K = 2048 and
    // N = 1536 shapes lose 10-25 %, N = K = 512 gains 8 % in isolation and nothing in the step; it stays the choice for N <= 128
    // and behind B200IQ_F16_TILE_N=128
    (void)c128; (void)c256;
    if (p->N <= 128) pair_bn = 128;
    if (force_bn == 128 || force_bn == 256) pair_bn = (p->N <= 128) ? 128 : force_bn;
  }
  const int BNL = pair ? pair_bn / 2 : 128;
  F16Maps m;
  memset(&m, 0, sizeof(m));
  int rc;
  // K-major plane [rows, K]: box = rows x 64 halves; MN-major plane [K, rows]: box = 64 k-rows x 64 halves
  if (!amn) rc = make_map_f16(&m.ah, p->A_hi, p->M, p->K, p->lda, FBM); else rc = make_map_f16(&m.ah, p->A_hi, p->K, p->M, p->lda, FBK);
  if (rc) return rc;
  if (!bmn) rc = make_map_f16(&m.bh, p->B_hi, p->N, p->K, p->ldb, BNL); else rc = make_map_f16(&m.bh, p->B_hi, p->K, p->N, p->ldb, FBK);
  if (rc) return rc;
  if (terms == 3) {
    rc = make_map_f16(&m.al, p->A_lo, p->M, p->K, p->lda, FBM); if (rc) return rc;
    rc = make_map_f16(&m.bl, p->B_lo, p->N, p->K, p->ldb, BNL); if (rc) return rc;
  }
  if (fast && p->C) { rc = make_map_f32_tile(&m.c, p->C, p->M, p->N, p->ldc); if (rc) return rc; }
  if (fast && (f & B200IQ_GEMM_RESIDUAL)) { rc = make_map_f32_tile(&m.r, p->residual, p->M, p->N, p->ldr); if (rc) return rc; }
  // gate tiles by TMA (RESIDUAL and RELU_MASK are exclusive, so the map slot is free): B200IQ_F16_GATE_TMA=0 keeps the per-lane loads
  static const bool gate_tma_on = [] { const char* v = getenv("B200IQ_F16_GATE_TMA"); return !(v && v[0] == '0'); }();
  int gm = 0;
  if (fast && gate_tma_on && (f & B200IQ_GEMM_RELU_MASK) && !p->C && !p->C_lo) {
    rc = make_map_f16_gate(&m.r, p->gate, p->M, p->N, p->ldg); if (rc) return rc;
    gm = 1;
  }
  const int fe = fast ? 1 : 0;
  const int epi_mode = (f & B200IQ_GEMM_DROPOUT) ? 2 : (f & (B200IQ_GEMM_RESIDUAL | B200IQ_GEMM_RELU_MASK)) ? 1 : 0;
#define F16_GO(BN_, AM_, BM_, T_, E_, C_) return launch_f16<BN_, AM_, BM_, T_, E_, C_>(m, fe, gm, e, p->K, split, st)
#define F16_TILE(AM_, BM_, T_, E_) do { if (pair && pair_bn == 256) F16_GO(256, AM_, BM_, T_, E_, 2); else if (pair) F16_GO(128, AM_, BM_, T_, E_, 2); else F16_GO(128, AM_, BM_, T_, E_, 1); } while (0)
  if (terms == 3) {
    if (epi_mode == 2) F16_TILE(false, false, 3, 2);
    if (epi_mode == 1) F16_TILE(false, false, 3, 1);
    F16_TILE(false, false, 3, 0);
  }
  if (!amn && !bmn) {
    if (epi_mode == 2) F16_TILE(false, false, 1, 2);
    if (epi_mode == 1) F16_TILE(false, false, 1, 1);
    F16_TILE(false, false, 1, 0);
  }
  if (!amn && bmn) {
    if (epi_mode == 1) F16_TILE(false, true, 1, 1);
    F16_TILE(false, true, 1, 0);
  }
  if (amn && bmn) {
    if (epi_mode == 1) F16_TILE(true, true, 1, 1);
    F16_TILE(true, true, 1, 0);
  }
  if (epi_mode == 1) F16_TILE(true, false, 1, 1);
  F16_TILE(true, false, 1, 0);
#undef F16_GO
#undef F16_TILE
}
