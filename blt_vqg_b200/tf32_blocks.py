"""The kind::tf32 block helpers the LSTM variant (lstm_iq.py) runs on: fp32 tensors rounded to TF32 by their producers,
weight operands from a rounded copy of the arena (b200iq_round_tf32), `ops.gemm` = tcgen05 kind::tf32.  The transformer path
moved to fp16 operand planes (transformer.py, gemm_f16.cu); this module keeps the first scheme's context and linear helpers."""
import torch

from . import _lib, ops
from .transformer import CONCURRENT_DW, Workspace  # noqa: F401  (Workspace re-exported for lstm_iq.py)


class Ctx:
    """everything a block schedule needs: parameter / gradient views, workspace, sizes, dropout keys"""

    def __init__(self, arena, gflat, ws, H, nh, F, E, attn_drop, relu_drop, training, seed, wflat=None):
        self.arena, self.ws = arena, ws
        self.P = arena.views      # fp32 masters: biases, LayerNorm / BatchNorm affine, embedding table
        # GEMM weight operands: the TF32-rounded copy of the arena (the masters themselves in exact mode)
        self.wflat = wflat if wflat is not None else arena.flat.detach()
        self.W = arena.views_over(wflat) if wflat is not None else arena.views
        self.gflat = gflat
        self.G = arena.views_over(gflat) if gflat is not None else None
        self.H, self.nh, self.F, self.E, self.d = H, nh, F, E, H // nh
        self.attn_drop = attn_drop if training else 0.0
        self.relu_drop = relu_drop if training else 0.0
        self.training = training
        self.seed = seed
        self.site = 0
        self.sm = 148
        self.on_ready = None  # callable(flat_offset): gradients at and after this arena offset are final (parallel.py)

    def side_stream(self):
        """the side stream paired with the CURRENT stream (the two encoders run on different streams, each gets its own)
        for work that is off the critical path: weight-gradient GEMMs, the decoder's context-side projections"""
        key = ("side_stream", torch.cuda.current_stream().cuda_stream)
        st = self.ws.meta.get(key)
        if st is None:
            st = self.ws.meta[key] = torch.cuda.Stream(device=self.ws.device)
        return st

    def mark(self):
        """an event at this point of the current stream: `on_side(fn, after=mark)` lets the caller issue its own
        critical-path kernel BEFORE the side work, so that kernel gets the SMs first and the side kernel fills its tail"""
        if not CONCURRENT_DW:
            return None
        key = ("side_event", torch.cuda.current_stream().cuda_stream)
        ev = self.ws.meta.get(key)
        if ev is None:
            ev = self.ws.meta[key] = torch.cuda.Event()
        ev.record()
        return ev

    def on_side(self, fn, after=None):
        """issue fn's launches on the side stream, ordered after `after` (a mark()) or after everything issued so far on
        the current stream.  The caller must join_side() before it overwrites anything fn reads and before it marks
        fn's gradients ready."""
        if not CONCURRENT_DW:
            return fn()
        side = self.side_stream()
        if after is not None:
            side.wait_event(after)
        else:
            side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), _lib.pinned_stream():
            fn()

    def join_side(self):
        if CONCURRENT_DW:
            torch.cuda.current_stream().wait_stream(self.side_stream())

    def side_mark(self):
        """an event after the side work issued so far: the current stream waits on it (wait_side_mark) before it overwrites
        what that work reads, without waiting for side work issued later"""
        if not CONCURRENT_DW:
            return None
        key = ("side_done_event", torch.cuda.current_stream().cuda_stream)
        ev = self.ws.meta.get(key)
        if ev is None:
            ev = self.ws.meta[key] = torch.cuda.Event()
        ev.record(self.side_stream())
        return ev

    def wait_side_mark(self, ev):
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)

    def ready(self, name):
        if self.on_ready is not None:
            self.on_ready(self.arena.offsets.get(name + "#pack", self.arena.offsets.get(name)))

    def pspan(self, first, last, cols):
        o0, o1 = self.arena.span(first, last)
        return self.wflat[o0:o1].view(-1, cols)

    def gspan(self, first, last, cols):
        o0, o1 = self.arena.span(first, last)
        return self.gflat[o0:o1].view(-1, cols)

    def key(self, tag):
        """(seed, offset) of a dropout site; remembered under `tag` so the backward re-creates the mask"""
        k = self.ws.meta.get(("key", tag))
        if k is None or k[0] != self.seed:
            self.site += 1
            k = (self.seed, self.site << 40)
            self.ws.meta[("key", tag)] = k
        return k

    def split_k(self, n_out, k_in, m_red):
        tiles = ((n_out + 127) // 128) * ((k_in + 127) // 128)
        s = max(1, (2 * self.sm) // tiles)
        return max(1, min(s, (m_red + 127) // 128))


def grad_w(c, dy, x, gw, gb=None):
    """dW[N,K] += dy[M,N]^T @ x[M,K];  db[N] += colsum(dy)"""
    M = dy.shape[0]
    ops.gemm(dy, x, gw, a_t=True, b_t=True, accumulate=True, split_k=c.split_k(gw.shape[0], gw.shape[1], M))
    if gb is not None:
        ops.colsum_add(dy, gb)



# ------------------------------------------------------------------ small MLPs
def linear_fwd(c, tag, pfx, x, relu=False, out=None, fp32_out=False):
    y = out if out is not None else c.ws.buf(tag, x.shape[0], c.W[pfx + ".weight"].shape[0])
    ops.gemm(x, c.W[pfx + ".weight"], y, bias=c.P[pfx + ".bias"], relu=relu, fp32_out=fp32_out)
    return y


def linear_bwd(c, tag, pfx, dy, x, relu_mask=None, need_dx=True, dx=None, accumulate_dx=False):
    """dy is the grad wrt the (pre-activation) output; relu_mask = post-relu INPUT x when x came out of a relu"""
    grad_w(c, dy, x, c.G[pfx + ".weight"], c.G[pfx + ".bias"])
    if not need_dx:
        return None
    if dx is None:
        dx = c.ws.buf(tag, dy.shape[0], c.W[pfx + ".weight"].shape[1])
    assert not (accumulate_dx and relu_mask is not None)
    ops.gemm(dy, c.W[pfx + ".weight"], dx, b_t=True, relu_mask=relu_mask, residual=dx if accumulate_dx else None,
             auto_split=True)   # e.g. z_classifier: 512 x 512 output over K = 15000 -> 4 tiles without split-K
    return dx
