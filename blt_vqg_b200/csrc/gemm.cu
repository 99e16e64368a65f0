// GEMM for every dense projection of the IQ path (see include/b200iq.h: b200iq_gemm).
//
// tcgen05 path (default): PERSISTENT kernel, one CTA per SM looping over (split-K slice, 128 x BN tile) work items.
//   warp 0  : TMA producer   (cp.async.bulk.tensor.2d, 128B swizzle, mbarrier complete_tx)
//   warp 1  : MMA issuer     (one elected lane, tcgen05.mma.cta_group::1.kind::tf32, fp32 accum in 2 TMEM buffers)
//   warps 2-5: epilogue      (tcgen05.ld 32x32b -> bias / residual / relu / relu-mask -> swizzled smem -> TMA store / reduce-add)
// Operands may be K-major ([rows,K] row-major) or MN-major ([K,rows] row-major): the second form is
// what the backward GEMMs (dX = dY.W, dW = dY^T.X) read, so no tensor is ever transposed in HBM.
// Out-of-bounds rows/cols/k are zero-filled by TMA, so arbitrary M, N, K are handled without padding.
#include "common.cuh"
#include "tcgen.cuh"
#include <stdlib.h>

namespace b200iq {

static constexpr int BM = 128;
static constexpr int BK = 32;          // 32 fp32 = 128 bytes = one swizzle row
static constexpr int UMMA_K = 8;       // kind::tf32
static constexpr int GEMM_THREADS = 320;   // producer warp + MMA warp + 8 epilogue warps
static constexpr int EPI_WARPS = 8;

struct GemmEpi {
  float* C; const float* bias; const float* residual;
  int M, N, ldc, ldr, flags; float alpha;
  int round;  // round C to TF32 (activations / activation gradients); never with ACCUM or FP32_OUT
  float drop_p, drop_scale; uint64_t drop_seed, drop_offset;  // DROPOUT: keep-mask from Philox(seed; offset + (row*N+col)/4)
};


// epilogue math shared by both paths
__device__ __forceinline__ float epi_value(float acc, const GemmEpi& e, int row, int col, bool add_bias) {
  float v = acc * e.alpha;
  if ((e.flags & B200IQ_GEMM_BIAS) && add_bias) v += __ldg(e.bias + col);
  if (e.flags & B200IQ_GEMM_RESIDUAL) v += __ldg(e.residual + (size_t)row * e.ldr + col);
  if (e.flags & B200IQ_GEMM_RELU) v = fmaxf(v, 0.f);
  if (e.flags & B200IQ_GEMM_RELU_MASK) v = (__ldg(e.residual + (size_t)row * e.ldr + col) > 0.f) ? v : 0.f;
  if (e.flags & B200IQ_GEMM_DROPOUT) {
    const size_t i = (size_t)row * e.N + col;
    uint32_t w[4];
    dropout_words(e.drop_seed, e.drop_offset, i, w);
    v = dropout_keep(w, i, dropout_threshold(e.drop_p)) ? v * e.drop_scale : 0.f;
  }
  return rnd(v, e.round);
}


// Persistent GEMM: grid = min(#work items, #SMs); a work item = (split-K slice, 128 x BN output tile).
//   warp 0   : TMA producer — one continuous smem pipeline (STAGES deep) across all work items of the CTA
//   warp 1   : MMA issuer   — tcgen05.mma kind::tf32 into one of TWO TMEM accumulators (2 x BN columns)
//   warps 2-5: epilogue     — tcgen05.ld 32 columns at a time -> epilogue math -> 128B-swizzled smem staging tile ->
//                             TMA store (or TMA reduce-add for split-K) ; overlaps the next item's main loop
// The epilogue of the first version wrote one row per thread (32 sectors per store request, ncu profiles/r01_gemm_*); staging
// through shared memory + TMA makes every global write a full 128-byte line and clips M/N edges in hardware.
template <int BN, bool A_MN, bool B_MN, int STAGES, int EPI, int CTAS>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmR, GemmEpi epi, int m_tiles, int n_tiles, int splits,
                 int kblocks_total, int kblocks_per_split, int use_tma_c) {
  constexpr int BNL = BN / CTAS;  // rows of B this CTA stages (the pair's MMA reads N/2 from each CTA)
  constexpr int A_BYTES = BM * 128, B_BYTES = BNL * 128, STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int STG_BYTES = 32 * 128;  // one 32 x 32 fp32 staging tile
  extern __shared__ uint8_t smem_raw[];
  pdl_trigger();  // the next kernel of the stream may take SMs as this grid's CTAs retire (it blocks in its own pdl_wait)
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* stg_base = smem + STAGES * STAGE_BYTES;  // [8 epilogue warps][2 buffers][4096 B], 1024-aligned
  __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], tfull_bar[2], tempty_bar[2];
  __shared__ uint64_t gbar[EPI == 1 ? EPI_WARPS : 1][2];  // residual / gate tiles arriving by TMA, per epilogue warp and staging buffer
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles = m_tiles * n_tiles, total = tiles * splits;
  const int rank = (CTAS == 2) ? (int)cluster_ctarank() : 0;
  const int w_first = blockIdx.x / CTAS, w_step = gridDim.x / CTAS;  // work items are dealt to CTAs (or CTA pairs)

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], EPI_WARPS * CTAS); }
    if (EPI == 1) for (int a = 0; a < EPI_WARPS; ++a) { mbar_init(&gbar[a][0], 1); mbar_init(&gbar[a][1], 1); }
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
  if (CTAS == 2) cluster_sync_all();  // the peer's barriers are initialised before any remote arrive / TMA signal
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_wait();  // barriers, TMEM and the cluster handshake are set up under the previous kernel's tail; operands are read after it

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      auto load = [&](void* dst, const CUtensorMap* tm, int c0, int c1, uint64_t* bar) {
        if (CTAS == 2) tma_load_2d_pair(dst, tm, c0, c1, bar); else tma_load_2d(dst, tm, c0, c1, bar);
      };
      for (int w = w_first; w < total; w += w_step) {
        const int z = w / tiles, r = w - z * tiles, mt = r / n_tiles, nt = r - mt * n_tiles;
        const int m0 = (mt * CTAS + rank) * BM, n0 = nt * BN + rank * BNL;
        const int kb0 = z * kblocks_per_split, kb1 = min(kblocks_total, kb0 + kblocks_per_split);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % STAGES, ph = (it / STAGES) & 1, k0 = kb * BK;
          mbar_wait(&empty_bar[s], ph ^ 1);
          if (rank == 0) mbar_expect_tx(&full_bar[s], STAGE_BYTES * CTAS);  // the leader's barrier counts both CTAs' bytes
          uint8_t* sa = smem + s * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          if (!A_MN) load(sa, &tmA, k0, m0, &full_bar[s]);
          else
            for (int i = 0; i < BM / 32; ++i) load(sa + i * (BK * 128), &tmA, m0 + 32 * i, k0, &full_bar[s]);
          if (!B_MN) load(sb, &tmB, k0, n0, &full_bar[s]);
          else
            for (int i = 0; i < BNL / 32; ++i) load(sb + i * (BK * 128), &tmB, n0 + 32 * i, k0, &full_bar[s]);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      // instruction descriptor: D=f32, A=B=tf32, majors, N>>3 @17, M>>4 @24
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                             ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((BM * CTAS) >> 4) << 24);
      uint32_t it = 0, ti = 0;
      for (int w = w_first; w < total; w += w_step, ++ti) {
        const int z = w / tiles;
        const int kb0 = z * kblocks_per_split, kb1 = min(kblocks_total, kb0 + kblocks_per_split);
        const uint32_t acc = ti & 1, aph = (ti >> 1) & 1;
        mbar_wait(&tempty_bar[acc], aph ^ 1);  // epilogue has drained this accumulator (passes at first use)
        tc_fence_after();
        const uint32_t tacc = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % STAGES, ph = (it / STAGES) & 1;
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * STAGE_BYTES), sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t ad = A_MN ? umma_desc(sa + k * 1024, BK * 128, 512, 1) : umma_desc(sa + k * 32, 16, 1024, 2);
            const uint64_t bd = B_MN ? umma_desc(sb + k * 1024, BK * 128, 512, 1) : umma_desc(sb + k * 32, 16, 1024, 2);
            if (CTAS == 2) umma_tf32_pair(tacc, ad, bd, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            else umma_tf32(tacc, ad, bd, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
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
    uint8_t* stg = stg_base + (warp - 2) * 2 * STG_BYTES;
    const bool accum = epi.flags & B200IQ_GEMM_ACCUM;
    // EPI: 0 = lean (bias / relu / alpha / rounding), 1 = + residual or relu-gate tile, 2 = + Philox dropout.  Separate
    // instantiations keep each unrolled epilogue body inside the instruction cache: with everything in one body the dropout
    // kernel was 8.7k SASS instructions and its epilogue warps stalled on instruction fetch (ncu: stall_no_inst dominant).
    const bool has_r = EPI == 1 && (epi.flags & (B200IQ_GEMM_RESIDUAL | B200IQ_GEMM_RELU_MASK));
    const bool gate = EPI == 1 && (epi.flags & B200IQ_GEMM_RELU_MASK), relu = epi.flags & B200IQ_GEMM_RELU;
    const bool drop = EPI == 2;
    const bool drop_pairs = (epi.N & 7) == 0;
    const uint32_t drop_thr = dropout_threshold(epi.drop_p);
    uint32_t ti = 0, chunk = 0;
    // EPI 1: the 32 x 32 residual / relu-gate tile of a chunk is pulled by TMA straight into the chunk's staging buffer (same
    // 128B swizzle as the output tile, so lane r finds row r where it will write row r), one chunk ahead of the math.
    // `ch` = the chunk counter value the tile belongs to; its buffer was last read by the store of chunk ch - 2.
    auto fetch_r = [&](int row0, int col0, uint32_t ch) {
      if (lane == 0) {
        bulk_wait_read<1>();
        uint64_t* gb = &gbar[EPI == 1 ? warp - 2 : 0][ch & 1];
        mbar_expect_tx(gb, STG_BYTES);
        tma_load_2d(stg + (ch & 1) * STG_BYTES, &tmR, col0, row0, gb);
      }
    };
    for (int w = w_first; w < total; w += w_step, ++ti) {
      const int z = w / tiles, r = w - z * tiles, mt = r / n_tiles, nt = r - mt * n_tiles;
      const int m0 = (mt * CTAS + rank) * BM, n0 = nt * BN;
      const uint32_t acc = ti & 1, aph = (ti >> 1) & 1;
      const int row0 = m0 + q * 32, row = row0 + lane;
      const bool add_bias = (z == 0);
      // the first chunk's residual tile travels while the MMAs of this tile are still running
      if (has_r && use_tma_c && row0 < epi.M && n0 + chalf * CHUNKS * 32 < epi.N) fetch_r(row0, n0 + chalf * CHUNKS * 32, chunk);
      mbar_wait(&tfull_bar[acc], aph);
      tc_fence_after();
#pragma unroll 1
      for (int c = chalf * CHUNKS; c < (chalf + 1) * CHUNKS; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= epi.N) break;  // warp-uniform
        // the chunk's 32 bias values travel while tcgen05.ld is in flight (loaded inside the j-loop below they cost one
        // L1/L2 round trip per float4: ncu showed the epilogue FFMAs waiting on exactly these loads)
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
        uint32_t rr[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + c * 32, rr);
        if (row0 >= epi.M) continue;  // warp-uniform: nothing of this quadrant is inside the matrix
        if (use_tma_c) {
          uint8_t* buf = stg + (chunk & 1) * STG_BYTES;
          if (has_r) {
            mbar_wait(&gbar[EPI == 1 ? warp - 2 : 0][chunk & 1], (chunk >> 1) & 1);  // the tile has landed
          } else {
            if (lane == 0) bulk_wait_read<1>();  // the store that last read this buffer has finished reading
            __syncwarp();
          }
          uint32_t keep[4] = {0u, 0u, 0u, 0u};
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            float4* slot = reinterpret_cast<float4*>(buf + lane * 128 + (((j >> 2) ^ (lane & 7)) << 4));
            float4 bb = bias4[j >> 2], gg = make_float4(0.f, 0.f, 0.f, 0.f);
            if (has_r) gg = *slot;
            // one Philox call covers 8 consecutive elements = this float4 and its neighbour (N % 4 == 0, ldc == N); when
            // N % 8 == 0 every row starts a group of 8, so the odd float4s reuse the words of the even ones
            const size_t ei = (size_t)row * epi.N + col0 + j;
            if (drop && (!drop_pairs || !(j & 4))) dropout_words(epi.drop_seed, epi.drop_offset, ei, keep);
            float v[4];
            const float* ga = &gg.x; const float* ba = &bb.x;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              float x = __uint_as_float(rr[j + t]) * epi.alpha + ba[t];
              if (gate) x = ga[t] > 0.f ? x : 0.f; else x += ga[t];
              if (relu) x = fmaxf(x, 0.f);
              if (drop) x = dropout_keep(keep, ei + t, drop_thr) ? x * epi.drop_scale : 0.f;
              v[t] = rnd(x, epi.round);
            }
            // 128B swizzle: 16-byte chunk index XOR (row & 7), the layout the C tensor map expects; TMA clips M/N edges
            *slot = make_float4(v[0], v[1], v[2], v[3]);
          }
          fence_async_smem();
          __syncwarp();
          if (lane == 0) {
            if (accum) tma_reduce_add_2d(&tmC, buf, col0, row0); else tma_store_2d(&tmC, buf, col0, row0);
            bulk_commit();
          }
          ++chunk;
          if (has_r && c + 1 < (chalf + 1) * CHUNKS && col0 + 32 < epi.N) fetch_r(row0, col0 + 32, chunk);  // next chunk's tile
        } else {
          // C cannot go through TMA (pitch / base not 16-byte aligned): rare, so compact code matters more than speed —
          // park the row in the staging buffer and run a rolled loop over it
          float* mine = reinterpret_cast<float*>(stg + lane * 128);
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<uint4*>(mine + j) = make_uint4(rr[j], rr[j + 1], rr[j + 2], rr[j + 3]);
          if (row < epi.M) {
            float* crow = epi.C + (size_t)row * epi.ldc;
#pragma unroll 1
            for (int j = 0; j < 32; ++j)
              if (col0 + j < epi.N) {
                const float v = epi_value(mine[j], epi, row, col0 + j, add_bias);
                if (accum) atomicAdd(crow + col0 + j, v); else crow[col0 + j] = v;
              }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (CTAS == 2) mbar_arrive_leader(&tempty_bar[acc]); else mbar_arrive(&tempty_bar[acc]); }
    }
    if (lane == 0) bulk_wait_all();  // every queued TMA store has landed before the CTA exits
  }
  tc_fence_before();
  __syncthreads();
  if (CTAS == 2) cluster_sync_all();  // no CTA of the pair leaves (or frees TMEM) while the other may still touch it
  if (warp == 2) {
    tc_fence_after();
    if (CTAS == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(2 * BN) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(2 * BN) : "memory");
  }
}

// ------------------------------------------------------------------ exact-fp32 CUDA-core path
// 64x64 tile, 256 threads, 4x4 micro-tile; generic strides cover every operand layout.
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const float* __restrict__ A, const float* __restrict__ B, GemmEpi epi, int K,
                 long sam, long sak, long sbn, long sbk, int kchunk) {
  __shared__ float As[16][64 + 4], Bs[16][64 + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int kbeg = blockIdx.z * kchunk, kend = min(K, kbeg + kchunk);
  float acc[4][4] = {};
  for (int k0 = kbeg; k0 < kend; k0 += 16) {
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      int kk, mm;
      if (sak == 1) { kk = i & 15; mm = i >> 4; } else { mm = i & 63; kk = i >> 6; }
      int gm = m0 + mm, gk = k0 + kk;
      As[kk][mm] = (gm < epi.M && gk < kend) ? A[gm * sam + gk * sak] : 0.f;
      int nn;
      if (sbk == 1) { kk = i & 15; nn = i >> 4; } else { nn = i & 63; kk = i >> 6; }
      int gn = n0 + nn; gk = k0 + kk;
      Bs[kk][nn] = (gn < epi.N && gk < kend) ? B[gn * sbn + gk * sbk] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty * 4 + i]; b[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  for (int i = 0; i < 4; ++i) {
    int row = m0 + ty * 4 + i;
    if (row >= epi.M) continue;
    for (int j = 0; j < 4; ++j) {
      int col = n0 + tx * 4 + j;
      if (col >= epi.N) continue;
      float v = epi_value(acc[i][j], epi, row, col, blockIdx.z == 0);
      float* dst = epi.C + (size_t)row * epi.ldc + col;
      if (epi.flags & B200IQ_GEMM_ACCUM) atomicAdd(dst, v); else *dst = v;
    }
  }
}

// ------------------------------------------------------------------ host side
// 2-D fp32 tensor map over a row-major [rows, cols] matrix with leading dimension ld; box = [box_rows, 32 cols]
static int make_map(CUtensorMap* tm, const float* base, int rows, int cols, int ld, int box_rows, bool mn_major) {
  EncodeTiledFn enc = get_encode();
  B200IQ_CHECK_ARG(enc, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {32, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200IQ_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) rows=%d cols=%d ld=%d", (int)r, rows, cols, ld);
  return 0;
}

template <int BN, bool A_MN, bool B_MN, int EPI, int CTAS>
static int launch_tc(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tr, int use_tma_c, const GemmEpi& e, int K,
                     int split, cudaStream_t st) {
  constexpr int STAGES = (CTAS == 2) ? 5 : (BN == 256) ? 3 : 4;
  constexpr int SMEM = STAGES * (BM * 128 + (BN / CTAS) * 128) + EPI_WARPS * 2 * 32 * 128 + 1024;
  static int attr_dev = -1;  // the opt-in is per device: redo it when the calling thread's device changes
  auto kern = gemm_tf32_kernel<BN, A_MN, B_MN, STAGES, EPI, CTAS>;
  int cur_dev = 0;
  cudaGetDevice(&cur_dev);
  if (attr_dev != cur_dev) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    attr_dev = cur_dev;
  }
  int kblocks = ceil_div(K, BK);
  if (split > kblocks) split = kblocks;
  int per = ceil_div(kblocks, split);
  split = ceil_div(kblocks, per);
  const int m_tiles = ceil_div(e.M, BM * CTAS), n_tiles = ceil_div(e.N, BN);
  const int total = m_tiles * n_tiles * split;
  const int slots = sm_count() / CTAS;
  const int grid = (total < slots ? total : slots) * CTAS;
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(GEMM_THREADS); cfg.dynamicSmemBytes = SMEM; cfg.stream = st;
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
    cudaError_t err = cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, tr, e, m_tiles, n_tiles, split, kblocks, per, use_tma_c);
    if (err != cudaSuccess) { set_error("gemm_tf32_kernel: %s", cudaGetErrorString(err)); return (int)err; }
  }
  return launch_status("gemm_tf32_kernel");
}

// 128 x 256 tiles halve the B-operand traffic per flop but quantise worse: pick by rounds over the SMs x relative tile cost
// returns 128 / 256 (one CTA per tile) or 512 = CTA pair (cta_group::2, 256 x 256 per pair): per k-block a CTA pulls
// 32 KB from L2 for a 128x128 tile, 48 KB for 128x256, and 32 KB for its 128x256 half of a pair's tile.
static int g_force_tile = 0;  // B200IQ_GEMM_TILE=128|256|512 (tuning / tests)
static int pick_bn(int M, int N, int K, int split, bool heavy_epi, bool dropout_epi = false) {
  static bool env_done = false;
  if (!env_done) { const char* v = getenv("B200IQ_GEMM_TILE"); g_force_tile = v ? atoi(v) : 0; env_done = true; }
  const bool pair_ok = N > 128 && M > 128;
  if (g_force_tile == 512) return pair_ok ? 512 : 128;
  if (g_force_tile == 128 || g_force_tile == 256) return N <= 128 ? 128 : g_force_tile;
  if (N <= 128) return 128;
  // measured on the cfg3 shapes (scripts/gpu_gemm_sweep.py, profiles/r01g_gemm_sweep.txt):
  //  * residual / relu-gate epilogues (tile pulled in by TMA): pairs for wide outputs, 128 x 128 tiles for N = 512
  //  * dropout epilogues: pairs for long reductions or big outputs, else 128 x 128
  //  * everything else: CTA pairs (20-35 % faster than 128 x 256 at K >= 1536, 5-10 % at K = 512)
  if (heavy_epi) return (pair_ok && N >= 1024) ? 512 : 128;
  if (dropout_epi) return (pair_ok && (K / split >= 1024 || (long)M * N >= (8192L * 2048))) ? 512 : 128;
  if (pair_ok) return 512;
  const int sms = sm_count();
  const long t128 = (long)ceil_div(M, BM) * ceil_div(N, 128) * split, t256 = (long)ceil_div(M, BM) * ceil_div(N, 256) * split;
  const double c128 = (double)((t128 + sms - 1) / sms) * 1.15, c256 = (double)((t256 + sms - 1) / sms) * 2.0;
  return c256 < c128 ? 256 : 128;
}

}  // namespace b200iq

using namespace b200iq;

extern "C" int b200iq_set_gemm_tile(int tile) {
  pick_bn(256, 256, 256, 1, false);  // make sure the environment default has been read before it is overridden
  const int prev = g_force_tile;
  g_force_tile = tile;
  return prev;
}

extern "C" int b200iq_gemm(const B200iqGemm* p, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B200IQ_CHECK_ARG(p && p->A && p->B && p->C, "gemm: null pointer");
  B200IQ_CHECK_ARG(p->M > 0 && p->N > 0 && p->K > 0, "gemm: bad shape M=%d N=%d K=%d", p->M, p->N, p->K);
  const int f = p->flags;
  const bool amn = f & B200IQ_GEMM_A_MN, bmn = f & B200IQ_GEMM_B_MN;
  B200IQ_CHECK_ARG(!(f & B200IQ_GEMM_BIAS) || p->bias, "gemm: BIAS without bias pointer");
  B200IQ_CHECK_ARG(!(f & (B200IQ_GEMM_RESIDUAL | B200IQ_GEMM_RELU_MASK)) || p->residual, "gemm: residual/gate pointer missing");
  B200IQ_CHECK_ARG(!((f & B200IQ_GEMM_RESIDUAL) && (f & B200IQ_GEMM_RELU_MASK)), "gemm: RESIDUAL and RELU_MASK are exclusive");
  int split = p->split_k < 1 ? 1 : p->split_k;
  B200IQ_CHECK_ARG(split == 1 || ((f & B200IQ_GEMM_ACCUM) && !(f & (B200IQ_GEMM_RELU | B200IQ_GEMM_RELU_MASK | B200IQ_GEMM_RESIDUAL))),
                   "gemm: split_k>1 needs ACCUM and a linear epilogue");
  B200IQ_CHECK_ARG(p->lda >= (amn ? p->M : p->K) && p->ldb >= (bmn ? p->N : p->K) && p->ldc >= p->N, "gemm: leading dimension too small");
  const int round = (round_mode() && !(f & (B200IQ_GEMM_ACCUM | B200IQ_GEMM_FP32_OUT | B200IQ_GEMM_SIMT))) ? 1 : 0;
  GemmEpi e{p->C, p->bias, p->residual, p->M, p->N, p->ldc, p->ldr, f, p->alpha, round, 0.f, 1.f, 0ull, 0ull};
  if (f & B200IQ_GEMM_DROPOUT) {
    B200IQ_CHECK_ARG(p->drop_p > 0.f && p->drop_p < 1.f, "gemm: DROPOUT needs 0 < drop_p < 1");
    B200IQ_CHECK_ARG(p->ldc == p->N && (p->N & 3) == 0 && split == 1 && !(f & B200IQ_GEMM_ACCUM) && (((uintptr_t)p->C) & 15) == 0,
                     "gemm: DROPOUT needs a contiguous, 16-byte aligned C with N %% 4 == 0, no split-K / ACCUM");
    e.drop_p = p->drop_p; e.drop_scale = 1.f / (1.f - p->drop_p); e.drop_seed = p->drop_seed; e.drop_offset = p->drop_offset;
  }

  if (f & B200IQ_GEMM_SIMT) {
    long sam = amn ? 1 : p->lda, sak = amn ? p->lda : 1, sbn = bmn ? 1 : p->ldb, sbk = bmn ? p->ldb : 1;
    int kchunk = ceil_div(ceil_div(p->K, split), 16) * 16;
    dim3 grid(ceil_div(p->N, 64), ceil_div(p->M, 64), ceil_div(p->K, kchunk));
    gemm_simt_kernel<<<grid, 256, 0, st>>>(p->A, p->B, e, p->K, sam, sak, sbn, sbk, kchunk);
    return launch_status("gemm_simt_kernel");
  }

  B200IQ_CHECK_ARG((((uintptr_t)p->A) & 15) == 0 && (((uintptr_t)p->B) & 15) == 0, "gemm: A/B must be 16-byte aligned");
  B200IQ_CHECK_ARG((p->lda & 3) == 0 && (p->ldb & 3) == 0, "gemm: lda/ldb must be multiples of 4 (lda=%d ldb=%d)", p->lda, p->ldb);
  CUtensorMap ta, tb, tc;
  int rc;
  const int tile = pick_bn(p->M, p->N, p->K, split, (f & (B200IQ_GEMM_RESIDUAL | B200IQ_GEMM_RELU_MASK)) != 0, (f & B200IQ_GEMM_DROPOUT) != 0);
  const int BN = tile == 512 ? 256 : tile, BNL = tile == 512 ? 128 : tile;
  if (!amn) rc = make_map(&ta, p->A, p->M, p->K, p->lda, BM, false); else rc = make_map(&ta, p->A, p->K, p->M, p->lda, BK, true);
  if (rc) return rc;
  if (!bmn) rc = make_map(&tb, p->B, p->N, p->K, p->ldb, BNL, false); else rc = make_map(&tb, p->B, p->K, p->N, p->ldb, BK, true);
  if (rc) return rc;
  // C goes out through TMA (32 x 32 boxes, 128B swizzle) when its pitch / base allow it; else per-thread stores
  // C goes out through TMA (32 x 32 boxes, 128B swizzle) when its pitch / base allow it — and so does the residual / gate
  // tile coming in; otherwise the per-thread fallback path
  const bool has_res = (f & (B200IQ_GEMM_RESIDUAL | B200IQ_GEMM_RELU_MASK)) != 0;
  const int use_tma_c = ((p->ldc & 3) == 0 && (((uintptr_t)p->C) & 15) == 0 &&
                         (!has_res || ((p->ldr & 3) == 0 && (((uintptr_t)p->residual) & 15) == 0))) ? 1 : 0;
  CUtensorMap tr;
  memset(&tr, 0, sizeof(tr));
  if (use_tma_c && has_res) { rc = make_map(&tr, p->residual, p->M, p->N, p->ldr, 32, false); if (rc) return rc; }
  if (use_tma_c) rc = make_map(&tc, p->C, p->M, p->N, p->ldc, 32, false); else memset(&tc, 0, sizeof(tc));
  if (rc) return rc;
#define B200IQ_LAUNCH2(BN_, FE_, CT_) \
  do { \
    if (!amn && !bmn) return launch_tc<BN_, false, false, FE_, CT_>(ta, tb, tc, tr, use_tma_c, e, p->K, split, st); \
    if (!amn && bmn) return launch_tc<BN_, false, true, FE_, CT_>(ta, tb, tc, tr, use_tma_c, e, p->K, split, st); \
    if (amn && !bmn) return launch_tc<BN_, true, false, FE_, CT_>(ta, tb, tc, tr, use_tma_c, e, p->K, split, st); \
    return launch_tc<BN_, true, true, FE_, CT_>(ta, tb, tc, tr, use_tma_c, e, p->K, split, st); \
  } while (0)
#define B200IQ_LAUNCH(BN_, CT_) \
  do { \
    if (epi_mode == 2) return launch_tc<BN_, false, false, 2, CT_>(ta, tb, tc, tr, use_tma_c, e, p->K, split, st); \
    if (epi_mode == 1) B200IQ_LAUNCH2(BN_, 1, CT_); else B200IQ_LAUNCH2(BN_, 0, CT_); \
  } while (0)
  // without a TMA-storable C the per-thread fallback path handles every flag itself (epi_value)
  const int epi_mode = (f & B200IQ_GEMM_DROPOUT) ? 2 : (f & (B200IQ_GEMM_RESIDUAL | B200IQ_GEMM_RELU_MASK)) ? 1 : 0;
  B200IQ_CHECK_ARG(epi_mode != 2 || (!amn && !bmn && !(f & (B200IQ_GEMM_RESIDUAL | B200IQ_GEMM_RELU_MASK))),
                   "gemm: the tcgen05 DROPOUT epilogue takes K-major A and B and no residual / gate");
  if (tile == 512) B200IQ_LAUNCH(256, 2);
  if (BN == 256) B200IQ_LAUNCH(256, 1);
  B200IQ_LAUNCH(128, 1);
#undef B200IQ_LAUNCH2
#undef B200IQ_LAUNCH
}
