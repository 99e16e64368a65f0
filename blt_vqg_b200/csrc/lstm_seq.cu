// Persistent LSTM recurrence: the whole time loop of nn.LSTM(H, H, 1) in ONE launch (forward) / ONE launch (BPTT).
// replaces: the per-timestep cuDNN recurrence of torch.nn.LSTM at models/decoder_rnn.py:73,105 and models/encoder_rnn.py:51,83
//           (and, here, the 2 launches per timestep of the first version: one tcgen05 GEMM + one cell kernel).
//
// Decomposition: a thread-block CLUSTER of 8 CTAs owns 16 batch rows; CTA c of the cluster owns hidden units
// [c*U, (c+1)*U), U = ceil(H/8) rounded up to even, and keeps the 4U rows of W_hh that produce ITS gates (i,f,g,o of its
// units) resident in shared memory for all T steps (H = 300: 152 x 308 floats = 187 KB).  Per step a CTA
//   1. pulls h_{t-1} of its 16 rows (all H units) from the h sequence in global memory (L2 hits: 19 KB, float4),
//   2. computes gates[16, 4U] = h . W_slice^T with mma.sync.m16n8k8 TF32 (operands from shared memory, bank-conflict-free
//      pitch = 20 mod 32); the slice's rows are stored unit-major / gate-minor so the four gates of one (row, unit) land
//      in two ADJACENT lanes of a C fragment and meet through one shuffle,
//   3. adds the precomputed input projection, applies the cell, keeps c in REGISTERS across time, writes activated gates /
//      c_t / h_t (the backward pass and the vocabulary GEMM read them) and
//   4. meets its 7 peers at a hardware cluster barrier (release / acquire), after which h_t is visible to them.
// The algorithmic HBM traffic per token is what SURVEY 8(d) counts: read x-projection 16H B, write gates 16H + h,c 8H B.
#include <stdlib.h>
#include <cuda_fp16.h>
#include <type_traits>
#include "common.cuh"

namespace b200iq {

static constexpr int LS_CL = 8;        // CTAs per cluster
static constexpr int LS_ROWS = 16;     // batch rows per cluster (one MMA m-tile)
static constexpr int LS_THREADS = 256; // 8 warps
static constexpr int LS_MAXT = 3;      // n-tiles per warp in the forward: 8 warps x 3 x 8 columns >= 4U  (U <= 48)

__device__ __forceinline__ void ls_mma(float (&c)[4], float a0, float a1, float a2, float a3, float b0, float b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(__float_as_uint(a0)), "r"(__float_as_uint(a1)), "r"(__float_as_uint(a2)), "r"(__float_as_uint(a3)),
                 "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}
__device__ __forceinline__ uint32_t ls_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t ls_clusterid() { uint32_t r; asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r)); return r; }
__device__ __forceinline__ void ls_cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void ls_cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ void ls_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ls_sigmoid(float x) { return 1.f / (1.f + __expf(-x)); }

struct LstmSeqDims { int T, B, H, U, Hp, LD; };  // U units per CTA, Hp = H rounded up to 8, LD = shared-memory pitch

// W_slice[4*ul + gate][k] = W_hh[gate*H + (rank*U + ul)][k]   (zero for units / k beyond H)
// round: store round-to-nearest TF32 values (the single-pass products of the BPTT when the caller's tensors are unrounded fp32)
__device__ __forceinline__ void ls_load_wslice(float* Ws, const float* __restrict__ whh, const LstmSeqDims& d, int rank, bool round = false) {
  // one warp per slice row, lanes along k: coalesced 128-byte reads, 8 rows in flight per warp
  const int rows = 4 * d.U, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int n = warp; n < rows; n += LS_THREADS / 32) {
    const int ul = n >> 2, gate = n & 3, u = rank * d.U + ul;
    const float* src = whh + (size_t)(gate * d.H + u) * d.H;
#pragma unroll 4
    for (int k = lane; k < d.LD; k += 32) {
      const float w = (u < d.H && k < d.H) ? __ldg(src + k) : 0.f;
      Ws[n * d.LD + k] = round ? tf32_rn(w) : w;
    }
  }
}

// PRECISE (the fp16-plane / no-rounding mode of the library): h and W_hh are unrounded fp32 and every product h.W_hh^T is the 3-term
// TF32 split hi.hi + lo.hi + hi.lo (hi = round-to-nearest TF32, lo = x - hi), ~2^-21 per operand: the encoder's final state
// feeds ReLU MLPs, where 11-bit forward products flip gates (profiles/r02a_cpu_precision_study.txt).
template <bool PRECISE>
__global__ void __cluster_dims__(LS_CL, 1, 1) __launch_bounds__(LS_THREADS, 1)
lstm_seq_fwd_kernel(float* gates, const float* __restrict__ whh, float* hbuf, float* cbuf,
                    const int* __restrict__ lengths, LstmSeqDims d, int rn) {
  extern __shared__ float ls_smem[];
  float* Ws = ls_smem;                    // [4U][LD]
  float* hs = Ws + 4 * d.U * d.LD;        // [16][LD]
  const int rank = ls_ctarank(), b0 = ls_clusterid() * LS_ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int NT = d.U / 2, KS = d.Hp / 8;
  const size_t BH = (size_t)d.B * d.H, B4H = 4 * BH;
  ls_load_wslice(Ws, whh, d, rank);
  // zero the k-padding of the h tile once (columns H..LD-1 are never written again)
  for (int idx = threadIdx.x; idx < LS_ROWS * d.LD; idx += LS_THREADS) hs[idx] = 0.f;
  // each lane owns ONE (row, unit) pair per n-tile of its warp: even t4 -> row g, odd t4 -> row g+8; unit = 2*nt + (t4>>1)
  const int my_row = g + 8 * (t4 & 1), b = b0 + my_row;
  const bool row_ok = b < d.B;
  const int len = (lengths && row_ok) ? __ldg(lengths + b) : 0x7fffffff;
  float creg[LS_MAXT], xg[LS_MAXT][4];
  int unit[LS_MAXT];
#pragma unroll
  for (int i = 0; i < LS_MAXT; ++i) {
    const int nt = warp + 8 * i;
    unit[i] = (nt < NT) ? rank * d.U + 2 * nt + (t4 >> 1) : d.H;  // >= H: nothing to do
    creg[i] = (row_ok && unit[i] < d.H) ? cbuf[(size_t)b * d.H + unit[i]] : 0.f;
  }
  auto fetch_x = [&](int t) {
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i)
#pragma unroll
      for (int q = 0; q < 4; ++q)
        xg[i][q] = (row_ok && unit[i] < d.H && t < len) ? __ldcg(gates + (size_t)t * B4H + (size_t)b * 4 * d.H + q * d.H + unit[i]) : 0.f;
  };
  fetch_x(0);
  __syncthreads();
  for (int t = 0; t < d.T; ++t) {
    // 1. h_{t-1} of the cluster's 16 rows -> shared memory (float4 when H allows it)
    const float* hprev = hbuf + (size_t)t * BH;
    if ((d.H & 3) == 0) {
      const int h4 = d.H >> 2;
      for (int idx = threadIdx.x; idx < LS_ROWS * h4; idx += LS_THREADS) {
        const int r = idx / h4, c4 = idx - r * h4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (b0 + r < d.B) v = __ldcg(reinterpret_cast<const float4*>(hprev + (size_t)(b0 + r) * d.H + 4 * c4));  // L2: written by peer SMs
        *reinterpret_cast<float4*>(hs + r * d.LD + 4 * c4) = v;
      }
    } else {
      for (int idx = threadIdx.x; idx < LS_ROWS * d.H; idx += LS_THREADS) {
        const int r = idx / d.H, c = idx - r * d.H;
        hs[r * d.LD + c] = (b0 + r < d.B) ? __ldcg(hprev + (size_t)(b0 + r) * d.H + c) : 0.f;
      }
    }
    __syncthreads();
    // 2. gates[16, 4U] = h . W_slice^T
    float acc[LS_MAXT][4];
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
#pragma unroll 2
    for (int ks = 0; ks < KS; ++ks) {
      const float* hp = hs + 8 * ks + t4;
      float a0 = hp[g * d.LD], a1 = hp[(g + 8) * d.LD], a2 = hp[g * d.LD + 4], a3 = hp[(g + 8) * d.LD + 4];
      float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
      if (PRECISE) {
        const float h0 = tf32_rn(a0), h1 = tf32_rn(a1), h2 = tf32_rn(a2), h3 = tf32_rn(a3);
        l0 = a0 - h0; l1 = a1 - h1; l2 = a2 - h2; l3 = a3 - h3;
        a0 = h0; a1 = h1; a2 = h2; a3 = h3;
      }
#pragma unroll
      for (int i = 0; i < LS_MAXT; ++i) {
        const int nt = warp + 8 * i;
        if (nt < NT) {
          const float* wp = Ws + (8 * nt + g) * d.LD + 8 * ks + t4;
          if (PRECISE) {
            const float w0 = wp[0], w1 = wp[4], wh0 = tf32_rn(w0), wh1 = tf32_rn(w1);
            ls_mma(acc[i], l0, l1, l2, l3, wh0, wh1);
            ls_mma(acc[i], a0, a1, a2, a3, w0 - wh0, w1 - wh1);
            ls_mma(acc[i], a0, a1, a2, a3, wh0, wh1);
          } else {
            ls_mma(acc[i], a0, a1, a2, a3, wp[0], wp[4]);
          }
        }
      }
    }
    // 3. the four gates of (my_row, unit) meet: even lanes keep row g and take (g,o) from the odd neighbour, odd lanes keep
    //    row g+8 and take (i,f) from the even neighbour
    float hnew[LS_MAXT], ga[LS_MAXT][4];
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i) {
      const bool odd = t4 & 1;
      const float s0 = odd ? acc[i][0] : acc[i][2], s1 = odd ? acc[i][1] : acc[i][3];
      const float r0 = __shfl_xor_sync(0xffffffffu, s0, 1), r1 = __shfl_xor_sync(0xffffffffu, s1, 1);
      float gi = odd ? r0 : acc[i][0], gf = odd ? r1 : acc[i][1], gg = odd ? acc[i][2] : r0, go = odd ? acc[i][3] : r1;
      hnew[i] = 0.f;
      if (row_ok && unit[i] < d.H) {
        if (t < len) {
          gi = ls_sigmoid(gi + xg[i][0]); gf = ls_sigmoid(gf + xg[i][1]); gg = tanhf(gg + xg[i][2]); go = ls_sigmoid(go + xg[i][3]);
          creg[i] = gf * creg[i] + gi * gg;
          hnew[i] = rnd(go * tanhf(creg[i]), rn);
        } else {
          hnew[i] = hs[my_row * d.LD + unit[i]];  // packed-sequence semantics: the state stops at the row's length
        }
        hbuf[(size_t)(t + 1) * BH + (size_t)b * d.H + unit[i]] = hnew[i];
      }
      ga[i][0] = gi; ga[i][1] = gf; ga[i][2] = gg; ga[i][3] = go;
    }
    // 4. h_t first, then ARRIVE (release: only the h stores are outstanding), and let everything the peers do not need —
    //    activated gates, c_t, the next step's input projection — overlap the barrier latency before the WAIT
    ls_cluster_arrive();
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i)
      if (row_ok && unit[i] < d.H) {
        if (t < len) {
          float* gout = gates + (size_t)t * B4H + (size_t)b * 4 * d.H + unit[i];
          gout[0] = ga[i][0]; gout[d.H] = ga[i][1]; gout[2 * d.H] = ga[i][2]; gout[3 * d.H] = ga[i][3];
        }
        cbuf[(size_t)(t + 1) * BH + (size_t)b * d.H + unit[i]] = creg[i];
      }
    if (t + 1 < d.T) fetch_x(t + 1);
    ls_cluster_wait();  // h_t of all 8 unit slices is visible; also orders this step's reads of `hs` before the refill
  }
}


// ---------------------------------------------------------------- precise forward on fp16 mma.sync
// lstm_seq_fwd_kernel<true> splits h and W_hh into TF32 hi / lo terms in registers at every k-step of every time step: ~8 k
// issued instructions per warp and step for 114 MMAs' worth of math (ncu profiles/r02i_lstm_seq_fwd.txt: 30 % issue-active on 8
// warps, stall_wait on top; 29.6 us per time step).  Here the CTA's W_hh slice is split ONCE, when it is loaded, into fp16 hi / lo
// half-rows that stay in shared memory (the same bytes as the fp32 slice), h_{t-1} is split once per step as it is staged, and
// the 3-term product hi.hi + lo.hi + hi.lo runs as mma.sync.m16n8k16 with ldmatrix fragments: one ldmatrix.x4 per A term and
// k-step, one per n-tile and k-step for all four B registers (hi and lo sit side by side in a row).
// W is stored scaled by 2^6 (|W_hh| ~ H^-1/2: its lo term would otherwise sit among fp16 subnormals); the accumulators are
// scaled back, both exactly.
static constexpr float LS_WSCALE = 64.f;
struct LstmSeqHDims { int KP, P; };   // KP = H rounded up to 16 (k extent), P = row pitch in halves (hi: [0, KP), lo: [KP, 2 KP))

__device__ __forceinline__ void ls_ldsm4(uint32_t (&r)[4], const __half* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ void ls_mma_f16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __cluster_dims__(LS_CL, 1, 1) __launch_bounds__(LS_THREADS, 1)
lstm_seq_fwd_h_kernel(float* gates, const float* __restrict__ whh, float* hbuf, float* cbuf,
                      const int* __restrict__ lengths, LstmSeqDims d, LstmSeqHDims hd) {
  extern __shared__ float ls_smem[];
  __half* Ws = reinterpret_cast<__half*>(ls_smem);   // [4U][P]: row n = 4 * unit + gate
  __half* hs = Ws + (size_t)4 * d.U * hd.P;          // [16][P]
  const int KP = hd.KP, P = hd.P;
  const int rank = ls_ctarank(), b0 = ls_clusterid() * LS_ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3, r8 = lane & 7, mi = lane >> 3;
  const int NT = d.U / 2, KS = KP / 16;
  const size_t BH = (size_t)d.B * d.H, B4H = 4 * BH;
  // the slice, split once: one warp per row, lanes along k (coalesced 128-byte reads)
  if ((d.H & 3) == 0) {
    // 16-byte pieces; a warp covers 4 rows per pass of 8 lanes x 4 pieces each... simply: piece p of the slice = (row p / KP4, k4 = p % KP4)
    const int KP4 = KP >> 2, H4 = d.H >> 2, pieces = 4 * d.U * KP4;
    for (int p0 = threadIdx.x; p0 < pieces; p0 += 4 * LS_THREADS) {
      float4 v[4];
      int nn[4], kk[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int pidx = p0 + j * LS_THREADS;
        v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        nn[j] = -1;
        if (pidx < pieces) {
          nn[j] = pidx / KP4;
          kk[j] = pidx - nn[j] * KP4;
          const int u = rank * d.U + (nn[j] >> 2);
          if (u < d.H && kk[j] < H4) v[j] = __ldg(reinterpret_cast<const float4*>(whh + (size_t)((nn[j] & 3) * d.H + u) * d.H) + kk[j]);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (nn[j] >= 0) {
          const float x0 = v[j].x * LS_WSCALE, x1 = v[j].y * LS_WSCALE, x2 = v[j].z * LS_WSCALE, x3 = v[j].w * LS_WSCALE;
          const __half2 h0 = __floats2half2_rn(x0, x1), h1 = __floats2half2_rn(x2, x3);
          const float2 f0 = __half22float2(h0), f1 = __half22float2(h1);
          const __half2 l0 = __floats2half2_rn(x0 - f0.x, x1 - f0.y), l1 = __floats2half2_rn(x2 - f1.x, x3 - f1.y);
          __half* o = Ws + (size_t)nn[j] * P + 4 * kk[j];
          *reinterpret_cast<uint2*>(o) = make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
          *reinterpret_cast<uint2*>(o + KP) = make_uint2(*reinterpret_cast<const uint32_t*>(&l0), *reinterpret_cast<const uint32_t*>(&l1));
        }
    }
  } else {
    for (int n = warp; n < 4 * d.U; n += LS_THREADS / 32) {
      const int ul = n >> 2, gate = n & 3, u = rank * d.U + ul;
      const float* src = whh + (size_t)(gate * d.H + u) * d.H;
      for (int k = lane; k < KP; k += 32) {
        const float w = (u < d.H && k < d.H) ? __ldg(src + k) * LS_WSCALE : 0.f;
        const __half hi = __float2half_rn(w);
        Ws[(size_t)n * P + k] = hi;
        Ws[(size_t)n * P + KP + k] = __float2half_rn(w - __half2float(hi));
      }
    }
  }
  for (int idx = threadIdx.x; idx < LS_ROWS * P; idx += LS_THREADS) hs[idx] = __float2half_rn(0.f);   // k padding stays zero
  const int my_row = g + 8 * (t4 & 1), b = b0 + my_row;
  const bool row_ok = b < d.B;
  const int len = (lengths && row_ok) ? __ldg(lengths + b) : 0x7fffffff;
  float creg[LS_MAXT], xg[LS_MAXT][4];
  int unit[LS_MAXT];
#pragma unroll
  for (int i = 0; i < LS_MAXT; ++i) {
    const int nt = warp + 8 * i;
    unit[i] = (nt < NT) ? rank * d.U + 2 * nt + (t4 >> 1) : d.H;  // >= H: nothing to do
    creg[i] = (row_ok && unit[i] < d.H) ? cbuf[(size_t)b * d.H + unit[i]] : 0.f;
  }
  auto fetch_x = [&](int t) {
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i)
#pragma unroll
      for (int q = 0; q < 4; ++q)
        xg[i][q] = (row_ok && unit[i] < d.H && t < len) ? __ldcg(gates + (size_t)t * B4H + (size_t)b * 4 * d.H + q * d.H + unit[i]) : 0.f;
  };
  fetch_x(0);
  // per-lane fragment addresses: A = h rows (r8 + 8 (mi & 1)), k half (mi >> 1); B = slice rows 8 nt + r8, registers
  // {hi k 0-7, hi k 8-15, lo k 0-7, lo k 8-15}
  const __half* ap = hs + (size_t)(r8 + 8 * (mi & 1)) * P + 8 * (mi >> 1);
  const __half* bp[LS_MAXT];
#pragma unroll
  for (int i = 0; i < LS_MAXT; ++i) {
    const int nt = warp + 8 * i;
    bp[i] = Ws + (size_t)(8 * (nt < NT ? nt : 0) + r8) * P + 8 * (mi & 1) + (mi >> 1) * KP;
  }
  const float wsinv = 1.f / LS_WSCALE;
  int ntw = 0;   // n-tiles of this warp
#pragma unroll
  for (int i = 0; i < LS_MAXT; ++i) ntw += (warp + 8 * i < NT) ? 1 : 0;
  __syncthreads();
  for (int t = 0; t < d.T; ++t) {
    // 1. h_{t-1} of the cluster's 16 rows -> fp16 hi / lo half-rows
    const float* hprev = hbuf + (size_t)t * BH;
    if ((d.H & 3) == 0) {
      // thread -> row tid / 16, 16-byte pieces (tid % 16) + 16 j of that row: all of a thread's loads are issued before the first use
      const int h4 = d.H >> 2, r = threadIdx.x >> 4, c0 = threadIdx.x & 15;
      const bool live = b0 + r < d.B;
      const float4* src = reinterpret_cast<const float4*>(hprev + (size_t)(b0 + r) * d.H);
      __half* orow = hs + (size_t)r * P;
      for (int cb = c0; cb < h4; cb += 64) {
        float4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (live && cb + 16 * j < h4) v[j] = __ldcg(src + cb + 16 * j);   // L2: written by peer SMs
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (cb + 16 * j < h4) {
            const __half2 h0 = __floats2half2_rn(v[j].x, v[j].y), h1 = __floats2half2_rn(v[j].z, v[j].w);
            const float2 f0 = __half22float2(h0), f1 = __half22float2(h1);
            const __half2 l0 = __floats2half2_rn(v[j].x - f0.x, v[j].y - f0.y), l1 = __floats2half2_rn(v[j].z - f1.x, v[j].w - f1.y);
            __half* o = orow + 4 * (cb + 16 * j);
            *reinterpret_cast<uint2*>(o) = make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
            *reinterpret_cast<uint2*>(o + KP) = make_uint2(*reinterpret_cast<const uint32_t*>(&l0), *reinterpret_cast<const uint32_t*>(&l1));
          }
      }
    } else {
      for (int idx = threadIdx.x; idx < LS_ROWS * d.H; idx += LS_THREADS) {
        const int r = idx / d.H, c = idx - r * d.H;
        const float v = (b0 + r < d.B) ? __ldcg(hprev + (size_t)(b0 + r) * d.H + c) : 0.f;
        const __half hi = __float2half_rn(v);
        hs[(size_t)r * P + c] = hi;
        hs[(size_t)r * P + KP + c] = __float2half_rn(v - __half2float(hi));
      }
    }
    __syncthreads();
    // 2. gates[16, 4U] = h . W_slice^T, three terms; consecutive MMAs go to different accumulators
    float acc[LS_MAXT][4];
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
    // (the warp's tile count is warp-uniform: one branch per step, no predicates around the .sync instructions)
    auto product = [&](auto ntw_c) {
      constexpr int NTW = decltype(ntw_c)::value;
#pragma unroll 2
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t ah[4], al[4], w[NTW > 0 ? NTW : 1][4];
        ls_ldsm4(ah, ap + 16 * ks);
        ls_ldsm4(al, ap + KP + 16 * ks);
#pragma unroll
        for (int i = 0; i < NTW; ++i) ls_ldsm4(w[i], bp[i] + 16 * ks);
#pragma unroll
        for (int i = 0; i < NTW; ++i) ls_mma_f16(acc[i], al, w[i][0], w[i][1]);
#pragma unroll
        for (int i = 0; i < NTW; ++i) ls_mma_f16(acc[i], ah, w[i][2], w[i][3]);
#pragma unroll
        for (int i = 0; i < NTW; ++i) ls_mma_f16(acc[i], ah, w[i][0], w[i][1]);
      }
    };
    if (ntw == 3) product(std::integral_constant<int, 3>{});
    else if (ntw == 2) product(std::integral_constant<int, 2>{});
    else if (ntw == 1) product(std::integral_constant<int, 1>{});
    // 3. the four gates of (my_row, unit) meet (see lstm_seq_fwd_kernel)
    float hnew[LS_MAXT], ga[LS_MAXT][4];
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i) {
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[i][e] *= wsinv;
      const bool odd = t4 & 1;
      const float s0 = odd ? acc[i][0] : acc[i][2], s1 = odd ? acc[i][1] : acc[i][3];
      const float r0 = __shfl_xor_sync(0xffffffffu, s0, 1), r1 = __shfl_xor_sync(0xffffffffu, s1, 1);
      float gi = odd ? r0 : acc[i][0], gf = odd ? r1 : acc[i][1], gg = odd ? acc[i][2] : r0, go = odd ? acc[i][3] : r1;
      hnew[i] = 0.f;
      if (row_ok && unit[i] < d.H) {
        if (t < len) {
          gi = ls_sigmoid(gi + xg[i][0]); gf = ls_sigmoid(gf + xg[i][1]); gg = tanhf(gg + xg[i][2]); go = ls_sigmoid(go + xg[i][3]);
          creg[i] = gf * creg[i] + gi * gg;
          hnew[i] = go * tanhf(creg[i]);
        } else {
          hnew[i] = __ldcg(hprev + (size_t)b * d.H + unit[i]);  // packed-sequence semantics: the state stops at the row's length
        }
        hbuf[(size_t)(t + 1) * BH + (size_t)b * d.H + unit[i]] = hnew[i];
      }
      ga[i][0] = gi; ga[i][1] = gf; ga[i][2] = gg; ga[i][3] = go;
    }
    // 4. h_t first, then ARRIVE; what the peers do not need overlaps the barrier latency before the WAIT
    ls_cluster_arrive();
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i)
      if (row_ok && unit[i] < d.H) {
        if (t < len) {
          float* gout = gates + (size_t)t * B4H + (size_t)b * 4 * d.H + unit[i];
          gout[0] = ga[i][0]; gout[d.H] = ga[i][1]; gout[2 * d.H] = ga[i][2]; gout[3 * d.H] = ga[i][3];
        }
        cbuf[(size_t)(t + 1) * BH + (size_t)b * d.H + unit[i]] = creg[i];
      }
    if (t + 1 < d.T) fetch_x(t + 1);
    ls_cluster_wait();  // h_t of all 8 unit slices is visible; also orders this step's reads of `hs` before the refill
  }
}

// ---------------------------------------------------------------- BPTT, same decomposition
// Per step (t = T-1 .. 0) a CTA turns d h_t, d c_t of ITS units into the four pre-activation gate gradients (registers ->
// dgates[t] in global memory for the weight-gradient GEMMs, and a [16, 4U] shared-memory tile), multiplies that tile with
// its resident W_hh slice — a K-slice of  d h_{t-1} = dgates . W_hh  — and publishes the [16, Hp] partial product; after
// the cluster barrier every CTA sums the 8 partials of its own units (fixed order: deterministic).  Two partial buffers
// alternate so that a fast CTA's next write never races a slow CTA's read.
static constexpr int LS_MAXTB = 5;  // n-tiles per warp in the backward: 8 warps x 5 x 8 columns >= Hp  (Hp <= 320)

__global__ void __cluster_dims__(LS_CL, 1, 1) __launch_bounds__(LS_THREADS, 1)
lstm_seq_bwd_kernel(const float* __restrict__ gates, const float* __restrict__ cbuf, const float* __restrict__ whh,
                    const float* __restrict__ dh_last, const float* __restrict__ d_out, float* __restrict__ dgates,
                    float* __restrict__ dh0, float* __restrict__ dc0, float* part, const int* __restrict__ lengths,
                    LstmSeqDims d, int LDG, int rn, int rnd_ops) {
  // rnd_ops: the caller's tensors are unrounded fp32 -> round the operands of the (single-pass) products to TF32 here: the W_hh
  // slice once at load, the gate-gradient tile as it is written to shared memory (mma.sync would truncate instead)
  extern __shared__ float ls_smem[];
  float* Ws = ls_smem;                    // [4U][LD]
  float* dgs = Ws + 4 * d.U * d.LD;       // [16][LDG]  (unit-major / gate-minor columns, like the rows of Ws)
  const int rank = ls_ctarank(), cl = ls_clusterid(), b0 = cl * LS_ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int NT = d.U / 2, NTB = d.Hp / 8, KS = d.U / 2;
  const size_t BH = (size_t)d.B * d.H, B4H = 4 * BH;
  ls_load_wslice(Ws, whh, d, rank, rnd_ops != 0);
  for (int idx = threadIdx.x; idx < LS_ROWS * LDG; idx += LS_THREADS) dgs[idx] = 0.f;
  const int my_row = g + 8 * (t4 & 1), b = b0 + my_row;
  const bool row_ok = b < d.B;
  const int len = (lengths && row_ok) ? __ldg(lengths + b) : 0x7fffffff;
  int unit[LS_MAXT], ul[LS_MAXT];
  float dh_run[LS_MAXT], dc[LS_MAXT], gt[LS_MAXT][4], cprev[LS_MAXT], ct[LS_MAXT], dout[LS_MAXT];
#pragma unroll
  for (int i = 0; i < LS_MAXT; ++i) {
    const int nt = warp + 8 * i;
    ul[i] = 2 * nt + (t4 >> 1);
    unit[i] = (nt < NT) ? rank * d.U + ul[i] : d.H;
    const bool ok = row_ok && unit[i] < d.H;
    dh_run[i] = (ok && dh_last) ? __ldg(dh_last + (size_t)b * d.H + unit[i]) : 0.f;
    dc[i] = 0.f;
    ct[i] = ok ? __ldg(cbuf + (size_t)d.T * BH + (size_t)b * d.H + unit[i]) : 0.f;
  }
  auto fetch = [&](int t) {
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i) {
      const bool ok = row_ok && unit[i] < d.H;
      const size_t o = (size_t)b * d.H + unit[i];
#pragma unroll
      for (int q = 0; q < 4; ++q) gt[i][q] = (ok && t < len) ? __ldg(gates + (size_t)t * B4H + (size_t)b * 4 * d.H + q * d.H + unit[i]) : 0.f;
      cprev[i] = ok ? __ldg(cbuf + (size_t)t * BH + o) : 0.f;
      dout[i] = (ok && d_out) ? __ldg(d_out + (size_t)t * BH + o) : 0.f;
    }
  };
  fetch(d.T - 1);
  __syncthreads();
  const size_t part_cl = (size_t)LS_CL * LS_ROWS * d.Hp;  // floats per cluster and parity
  for (int t = d.T - 1; t >= 0; --t) {
    // a-c. cell backward for my (row, unit) pairs
    float dgk[LS_MAXT][4];
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i) {
      float (&dg4)[4] = dgk[i];
      dg4[0] = dg4[1] = dg4[2] = dg4[3] = 0.f;
      if (row_ok && unit[i] < d.H) {
        const float dh = dh_run[i] + dout[i];
        if (t < len) {
          const float gi = gt[i][0], gf = gt[i][1], gg = gt[i][2], go = gt[i][3];
          const float tc = tanhf(ct[i]);
          const float dct = dc[i] + dh * go * (1.f - tc * tc);
          dg4[0] = rnd(dct * gg * gi * (1.f - gi), rn);
          dg4[1] = rnd(dct * cprev[i] * gf * (1.f - gf), rn);
          dg4[2] = rnd(dct * gi * (1.f - gg * gg), rn);
          dg4[3] = rnd(dh * tc * go * (1.f - go), rn);
          dc[i] = dct * gf;
          dh_run[i] = 0.f;
        } else {
          dh_run[i] = dh;  // the state was copied at this step: the gradient passes straight through, dc unchanged
        }
        *reinterpret_cast<float4*>(dgs + my_row * LDG + 4 * ul[i]) =
            rnd_ops ? make_float4(tf32_rn(dg4[0]), tf32_rn(dg4[1]), tf32_rn(dg4[2]), tf32_rn(dg4[3])) : make_float4(dg4[0], dg4[1], dg4[2], dg4[3]);
        ct[i] = cprev[i];  // c_{t-1} is the next step's c_t
      }
    }
    __syncthreads();
    // e. partial d h_{t-1}[16, Hp] = dgates_slice[16, 4U] . W_slice[4U, Hp]
    float acc[LS_MAXTB][4];
#pragma unroll
    for (int i = 0; i < LS_MAXTB; ++i)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
#pragma unroll 2
    for (int ks = 0; ks < KS; ++ks) {
      const float* ap = dgs + 8 * ks + t4;
      const float a0 = ap[g * LDG], a1 = ap[(g + 8) * LDG], a2 = ap[g * LDG + 4], a3 = ap[(g + 8) * LDG + 4];
      const float* wp = Ws + (8 * ks + t4) * d.LD + g;
#pragma unroll
      for (int i = 0; i < LS_MAXTB; ++i) {
        const int nb = warp + 8 * i;
        if (nb < NTB) ls_mma(acc[i], a0, a1, a2, a3, wp[8 * nb], wp[4 * d.LD + 8 * nb]);
      }
    }
    // f. publish the partial product
    float* pp = part + ((size_t)(t & 1) * gridDim.x / LS_CL + cl) * part_cl + (size_t)rank * LS_ROWS * d.Hp;
#pragma unroll
    for (int i = 0; i < LS_MAXTB; ++i) {
      const int nb = warp + 8 * i;
      if (nb < NTB) {
        *reinterpret_cast<float2*>(pp + (size_t)g * d.Hp + 8 * nb + 2 * t4) = make_float2(acc[i][0], acc[i][1]);
        *reinterpret_cast<float2*>(pp + (size_t)(g + 8) * d.Hp + 8 * nb + 2 * t4) = make_float2(acc[i][2], acc[i][3]);
      }
    }
    ls_cluster_arrive();  // release covers the partial product only; dgates (not needed by the peers) goes out behind it
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i)
      if (row_ok && unit[i] < d.H) {
        float* gout = dgates + (size_t)t * B4H + (size_t)b * 4 * d.H + unit[i];
        gout[0] = dgk[i][0]; gout[d.H] = dgk[i][1]; gout[2 * d.H] = dgk[i][2]; gout[3 * d.H] = dgk[i][3];
      }
    if (t > 0) fetch(t - 1);  // in flight across the barrier
    ls_cluster_wait();
    // h. d h_{t-1} of my (row, unit) pairs = pass-through + the 8 K-slices, summed in rank order
    const float* pc = part + ((size_t)(t & 1) * gridDim.x / LS_CL + cl) * part_cl + (size_t)my_row * d.Hp;
#pragma unroll
    for (int i = 0; i < LS_MAXT; ++i)
      if (row_ok && unit[i] < d.H) {
        float s = 0.f;
#pragma unroll
        for (int src = 0; src < LS_CL; ++src) s += __ldcg(pc + (size_t)src * LS_ROWS * d.Hp + unit[i]);
        dh_run[i] += rnd(s, rn);
      }
  }
#pragma unroll
  for (int i = 0; i < LS_MAXT; ++i)
    if (row_ok && unit[i] < d.H) {
      dh0[(size_t)b * d.H + unit[i]] = dh_run[i];
      dc0[(size_t)b * d.H + unit[i]] = dc[i];
    }
}

}  // namespace b200iq
using namespace b200iq;

static int ls_dims(int T, int B, int H, LstmSeqDims* d, size_t* smem_fwd) {
  d->T = T; d->B = B; d->H = H;
  d->U = ((H + LS_CL - 1) / LS_CL + 1) & ~1;
  d->Hp = (H + 7) & ~7;
  d->LD = d->Hp + ((20 - d->Hp % 32) + 32) % 32;  // pitch = 20 (mod 32): the 8 x 4 fragment loads hit 32 distinct banks
  *smem_fwd = sizeof(float) * ((size_t)4 * d->U * d->LD + (size_t)LS_ROWS * d->LD);
  return 0;
}

// 1 when the persistent kernels can take this width (W_hh slice + h tile fit the 227 KB of one SM), else 0
extern "C" int b200iq_lstm_seq_supported(int H) {
  LstmSeqDims d; size_t smem;
  ls_dims(1, 1, H, &d, &smem);
  return (H >= 8 && d.U / 2 <= 8 * LS_MAXT && d.Hp <= 320 && smem + 1024 <= 227 * 1024) ? 1 : 0;
}

extern "C" int b200iq_lstm_seq_fwd(float* gates, const float* whh, float* hbuf, float* cbuf, const int* lengths,
                                   int T, int B, int H, void* stream) {
  B200IQ_CHECK_ARG(gates && whh && hbuf && cbuf && T > 0 && B > 0 && H > 0, "lstm_seq_fwd: bad args");
  B200IQ_CHECK_ARG(b200iq_lstm_seq_supported(H), "lstm_seq_fwd: H=%d does not fit the shared-memory resident slice", H);
  B200IQ_CHECK_ARG((((uintptr_t)hbuf) & 15) == 0, "lstm_seq_fwd: hbuf must be 16-byte aligned");
  LstmSeqDims d; size_t smem;
  ls_dims(T, B, H, &d, &smem);
  static int attr_dev = -1;  // the opt-in is per device: redo it when the calling thread's device changes
  int cur_dev = 0;
  cudaGetDevice(&cur_dev);
  if (attr_dev != cur_dev) {
    cudaFuncSetAttribute(lstm_seq_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    cudaFuncSetAttribute(lstm_seq_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    attr_dev = cur_dev;
  }
  const int clusters = (B + LS_ROWS - 1) / LS_ROWS;
  // rounding mode off = the caller keeps unrounded fp32 tensors (fp16-plane scheme): 3-term products, on the fp16 mma.sync kernel
  // when its tiles fit (B200IQ_LSTM_FWD_H=0: the TF32 register-split kernel, the A/B switch)
  if (!round_mode()) {
    static const bool use_h = [] { const char* e = getenv("B200IQ_LSTM_FWD_H"); return !(e && e[0] == '0'); }();
    LstmSeqHDims hd;
    hd.KP = (H + 15) & ~15;
    hd.P = 2 * hd.KP;
    while (((2 * hd.P) % 128) / 16 % 2 == 0) hd.P += 8;   // row pitch = odd multiple of 16 B (mod 128): conflict-free ldmatrix
    const size_t smem_h = sizeof(__half) * (size_t)(4 * d.U + LS_ROWS) * hd.P;
    if (use_h && smem_h + 1024 <= 227 * 1024) {
      static int attr_dev = -1;
      int dev = 0;
      cudaGetDevice(&dev);
      if (attr_dev != dev) { cudaFuncSetAttribute(lstm_seq_fwd_h_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024); attr_dev = dev; }
      lstm_seq_fwd_h_kernel<<<clusters * LS_CL, LS_THREADS, smem_h, (cudaStream_t)stream>>>(gates, whh, hbuf, cbuf, lengths, d, hd);
      return launch_status("lstm_seq_fwd_h");
    }
  }
  if (round_mode()) lstm_seq_fwd_kernel<false><<<clusters * LS_CL, LS_THREADS, smem, (cudaStream_t)stream>>>(gates, whh, hbuf, cbuf, lengths, d, 1);
  else lstm_seq_fwd_kernel<true><<<clusters * LS_CL, LS_THREADS, smem, (cudaStream_t)stream>>>(gates, whh, hbuf, cbuf, lengths, d, 0);
  return launch_status("lstm_seq_fwd");
}

static int ls_ldg(const LstmSeqDims& d) { const int w = 4 * d.U; return w + ((4 - w % 32) + 32) % 32; }  // pitch = 4 (mod 32)

extern "C" size_t b200iq_lstm_seq_bwd_workspace_bytes(int B, int H) {
  LstmSeqDims d; size_t smem;
  ls_dims(1, B, H, &d, &smem);
  const size_t clusters = (B + LS_ROWS - 1) / LS_ROWS;
  return sizeof(float) * 2 * clusters * LS_CL * LS_ROWS * d.Hp;
}

extern "C" int b200iq_lstm_seq_bwd(const float* gates, const float* cbuf, const float* whh, const float* dh_last, const float* d_out,
                                   float* dgates, float* dh0, float* dc0, float* workspace, const int* lengths,
                                   int T, int B, int H, void* stream) {
  B200IQ_CHECK_ARG(gates && cbuf && whh && dgates && dh0 && dc0 && workspace && T > 0 && B > 0 && H > 0, "lstm_seq_bwd: bad args");
  LstmSeqDims d; size_t smem;
  ls_dims(T, B, H, &d, &smem);
  const int LDG = ls_ldg(d);
  smem = sizeof(float) * ((size_t)4 * d.U * d.LD + (size_t)LS_ROWS * LDG);
  B200IQ_CHECK_ARG(b200iq_lstm_seq_supported(H) && d.Hp / 8 <= 8 * LS_MAXTB && smem + 1024 <= 227 * 1024,
                   "lstm_seq_bwd: H=%d does not fit the shared-memory resident slice", H);
  B200IQ_CHECK_ARG((((uintptr_t)workspace) & 15) == 0, "lstm_seq_bwd: workspace must be 16-byte aligned");
  static int attr_dev = -1;  // per device, like the forward
  int cur_dev = 0;
  cudaGetDevice(&cur_dev);
  if (attr_dev != cur_dev) { cudaFuncSetAttribute(lstm_seq_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024); attr_dev = cur_dev; }
  const int clusters = (B + LS_ROWS - 1) / LS_ROWS;
  lstm_seq_bwd_kernel<<<clusters * LS_CL, LS_THREADS, smem, (cudaStream_t)stream>>>(gates, cbuf, whh, dh_last, d_out, dgates, dh0, dc0,
                                                                                    workspace, lengths, d, LDG, round_mode(), round_mode() ? 0 : 1);
  return launch_status("lstm_seq_bwd");
}

// ------------------------------------------------------------------ allocation-free variant of b200iq_lstm_seq_bwd
// No counterpart in the reference: for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
// that forbid it).  The partial-product scratch of the BPTT kernel lives in a file-scope device array instead of a caller-owned
// workspace, so every GPU that loads the library holds B200IQ_LSTM_SEQ_BWD_STATIC_MAX_BYTES of HBM until the process exits.
// All calls on one device share the array: calls that overlap on the GPU (different streams) race on it.  Nothing in it has to
// start at zero — every (parity, cluster, rank) slot the kernel reads at step t was written by the same launch at step t before
// the cluster barrier — so the previous call's contents need no clearing.
static __device__ __align__(256) float g_lstm_bwd_scratch[B200IQ_LSTM_SEQ_BWD_STATIC_MAX_BYTES / sizeof(float)];

// `g_lstm_bwd_scratch` in HOST code is only the symbol's host-side stand-in, not a device address: resolve it once per device
// (with that device current) through cudaGetSymbolAddress and hand the kernel the result.
static int ls_static_scratch(float** out) {
  constexpr int MAXDEV = 64;
  static float* ptrs[MAXDEV] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_error("lstm_seq_bwd_static: cudaGetDevice: %s", cudaGetErrorString(e)); return (int)e; }
  if (dev < 0 || dev >= MAXDEV) { set_error("lstm_seq_bwd_static: device ordinal %d out of range", dev); return -1; }
  float* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    e = cudaGetSymbolAddress((void**)&p, g_lstm_bwd_scratch);
    if (e != cudaSuccess || !p) {
      set_error("lstm_seq_bwd_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : -1;
    }
    // function attributes are per device too: part of the same once-per-device set-up
    e = cudaFuncSetAttribute(lstm_seq_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) { set_error("lstm_seq_bwd_static: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);  // racing threads resolve the same address
  }
  *out = p;
  return 0;
}

extern "C" size_t b200iq_lstm_seq_bwd_static_max_bytes(void) { return B200IQ_LSTM_SEQ_BWD_STATIC_MAX_BYTES; }

extern "C" int b200iq_lstm_seq_bwd_static(const float* gates, const float* cbuf, const float* whh, const float* dh_last,
                                          const float* d_out, float* dgates, float* dh0, float* dc0, const int* lengths,
                                          int T, int B, int H, void* stream) {
  B200IQ_CHECK_ARG(gates && cbuf && whh && dgates && dh0 && dc0 && T > 0 && B > 0 && H > 0, "lstm_seq_bwd_static: bad args");
  LstmSeqDims d; size_t smem;
  ls_dims(T, B, H, &d, &smem);
  const int LDG = ls_ldg(d);
  smem = sizeof(float) * ((size_t)4 * d.U * d.LD + (size_t)LS_ROWS * LDG);
  B200IQ_CHECK_ARG(b200iq_lstm_seq_supported(H) && d.Hp / 8 <= 8 * LS_MAXTB && smem + 1024 <= 227 * 1024,
                   "lstm_seq_bwd_static: H=%d does not fit the shared-memory resident slice", H);
  const size_t need = b200iq_lstm_seq_bwd_workspace_bytes(B, H);
  B200IQ_CHECK_ARG(need <= (size_t)B200IQ_LSTM_SEQ_BWD_STATIC_MAX_BYTES,
                   "lstm_seq_bwd_static: B=%d, H=%d needs %zu bytes of scratch, the built-in array holds %zu "
                   "(use b200iq_lstm_seq_bwd with a caller-owned workspace)", B, H, need, (size_t)B200IQ_LSTM_SEQ_BWD_STATIC_MAX_BYTES);
  float* scratch = nullptr;
  if (int rc = ls_static_scratch(&scratch)) return rc;
  const int clusters = (B + LS_ROWS - 1) / LS_ROWS;
  lstm_seq_bwd_kernel<<<clusters * LS_CL, LS_THREADS, smem, (cudaStream_t)stream>>>(gates, cbuf, whh, dh_last, d_out, dgates, dh0, dc0,
                                                                                    scratch, lengths, d, LDG, round_mode(), round_mode() ? 0 : 1);
  return launch_status("lstm_seq_bwd_static");
}
