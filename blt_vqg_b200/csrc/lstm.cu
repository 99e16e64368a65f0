// LSTM cell kernels for the recurrence of the LSTM variant of IQ.
// replaces: torch.nn.LSTM(H, H, 1, batch_first=True) at models/decoder_rnn.py:73,105 and models/encoder_rnn.py:51,83
//           (cuDNN RNN forward + backward), including the packed-sequence length masking of encoder_rnn.py:80-86.
// Layout is TIME-MAJOR: gates[t] is a contiguous [B, 4H] slab (PyTorch gate order i|f|g|o along the 4H axis),
// h/c sequences are [T+1, B, H] with slot 0 = the initial state, so "h_{t-1} of every step" and "h_t of every step" are both
// contiguous [T*B, H] matrices for the weight-gradient / vocabulary GEMMs.  Per step the host launches ONE tcgen05 GEMM
// (gates_pre[t] = h[t] . W_hh^T + xproj[t], residual epilogue) and ONE of these elementwise kernels; everything that
// does not depend on the recurrence (input projection of all T steps, dW_ih, dW_hh, dx, bias grads) is a single big GEMM.
#include "common.cuh"

namespace b200iq {

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

// gates: in = pre-activations [B,4H]; out (in place) = activated (i,f,o sigmoid; g tanh), kept for the backward pass.
// c_next = f*c_prev + i*g ; h_next = o*tanh(c_next).  Rows with t >= lengths[b] copy their state (packed-sequence semantics).
__global__ void lstm_cell_fwd_kernel(float* __restrict__ gates, const float* __restrict__ c_prev, const float* __restrict__ h_prev,
                                     float* __restrict__ c_next, float* __restrict__ h_next, const int* __restrict__ lengths,
                                     int t, int B, int H, int rn) {
  const int n = B * H;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
    const int b = idx / H, j = idx - b * H;
    if (lengths && t >= lengths[b]) {
      c_next[idx] = c_prev[idx];
      h_next[idx] = h_prev[idx];
      continue;
    }
    float* g = gates + (size_t)b * 4 * H;
    const float i = sigmoidf_(g[j]), f = sigmoidf_(g[H + j]), gg = tanhf(g[2 * H + j]), o = sigmoidf_(g[3 * H + j]);
    const float c = f * c_prev[idx] + i * gg;
    g[j] = i; g[H + j] = f; g[2 * H + j] = gg; g[3 * H + j] = o;
    c_next[idx] = c;
    h_next[idx] = rnd(o * tanhf(c), rn);
  }
}

// Backward of one step.  dh_run (in): d loss / d h_t from the future (recurrent + pass-through); dh_out: d loss / d h_t from
// the consumers of the output sequence (may be NULL); dc (in/out): d loss / d c_t -> d loss / d c_{t-1}.
// Writes dgates (pre-activation gradients, [B,4H]) and leaves in dh_run ONLY the pass-through part (masked rows), so the
// following GEMM  dh_run = dgates . W_hh + dh_run  yields d loss / d h_{t-1}.
__global__ void lstm_cell_bwd_kernel(const float* __restrict__ gates, const float* __restrict__ c_prev, const float* __restrict__ c_t,
                                     float* __restrict__ dh_run, const float* __restrict__ dh_out, float* __restrict__ dc,
                                     float* __restrict__ dgates, const int* __restrict__ lengths, int t, int B, int H, int rn) {
  const int n = B * H;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
    const int b = idx / H, j = idx - b * H;
    float* dg = dgates + (size_t)b * 4 * H;
    const float dh = dh_run[idx] + (dh_out ? dh_out[idx] : 0.f);
    if (lengths && t >= lengths[b]) {
      dg[j] = 0.f; dg[H + j] = 0.f; dg[2 * H + j] = 0.f; dg[3 * H + j] = 0.f;
      dh_run[idx] = dh;  // state was copied: gradient passes straight through, dc unchanged
      continue;
    }
    const float* g = gates + (size_t)b * 4 * H;
    const float i = g[j], f = g[H + j], gg = g[2 * H + j], o = g[3 * H + j];
    const float tc = tanhf(c_t[idx]);
    const float dct = dc[idx] + dh * o * (1.f - tc * tc);
    dg[j] = rnd(dct * gg * i * (1.f - i), rn);
    dg[H + j] = rnd(dct * c_prev[idx] * f * (1.f - f), rn);
    dg[2 * H + j] = rnd(dct * i * (1.f - gg * gg), rn);
    dg[3 * H + j] = rnd(dh * tc * o * (1.f - o), rn);
    dc[idx] = dct * f;
    dh_run[idx] = 0.f;
  }
}

// out[t*B + b] = in[b*T + t]  (batch-major token ids -> time-major), optionally dropping the last `skip_last` columns
__global__ void transpose_ids_kernel(const int* __restrict__ in, int* __restrict__ out, int B, int T_in, int T_out, int col0) {
  int n = B * T_out;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
    int t = idx / B, b = idx - t * B;
    out[idx] = in[b * T_in + col0 + t];
  }
}

}  // namespace b200iq
using namespace b200iq;
#define ST ((cudaStream_t)stream)

extern "C" int b200iq_lstm_cell_fwd(float* gates, const float* c_prev, const float* h_prev, float* c_next, float* h_next,
                                    const int* lengths, int t, int B, int H, void* stream) {
  B200IQ_CHECK_ARG(gates && c_prev && h_prev && c_next && h_next && B > 0 && H > 0 && t >= 0, "lstm_cell_fwd: bad args");
  int grid = min(ceil_div(B * H, 256), sm_count() * 8);
  lstm_cell_fwd_kernel<<<grid, 256, 0, ST>>>(gates, c_prev, h_prev, c_next, h_next, lengths, t, B, H, round_mode());
  return launch_status("lstm_cell_fwd");
}

extern "C" int b200iq_lstm_cell_bwd(const float* gates, const float* c_prev, const float* c_t, float* dh_run, const float* dh_out,
                                    float* dc, float* dgates, const int* lengths, int t, int B, int H, void* stream) {
  B200IQ_CHECK_ARG(gates && c_prev && c_t && dh_run && dc && dgates && B > 0 && H > 0 && t >= 0, "lstm_cell_bwd: bad args");
  int grid = min(ceil_div(B * H, 256), sm_count() * 8);
  lstm_cell_bwd_kernel<<<grid, 256, 0, ST>>>(gates, c_prev, c_t, dh_run, dh_out, dc, dgates, lengths, t, B, H, round_mode());
  return launch_status("lstm_cell_bwd");
}

extern "C" int b200iq_transpose_ids(const int* in, int* out, int B, int T_in, int T_out, int col0, void* stream) {
  B200IQ_CHECK_ARG(in && out && B > 0 && T_out > 0 && col0 >= 0 && col0 + T_out <= T_in, "transpose_ids: bad args");
  transpose_ids_kernel<<<min(ceil_div(B * T_out, 256), sm_count() * 4), 256, 0, ST>>>(in, out, B, T_in, T_out, col0);
  return launch_status("transpose_ids");
}
