"""Host-side forward / hand-written backward schedules of the transformer blocks.

Mirrors models/transformer_layers.py (Encoder :138-152, EncoderLayer :260-282, Decoder :205-221,
DecoderLayer :326-364, MultiHeadAttention :486-532, PositionwiseFeedForward :400-408) as explicit
kernel launch sequences over pre-allocated workspace buffers; there is no autograd underneath.
Every function only enqueues kernels on the current stream (CUDA-graph capturable).
"""
import os

import torch

from . import _lib, ops

PAD = 0
# small context-side GEMMs of the decoder (cross-attention K|V projections, their gradients) on a side stream
CONCURRENT = os.environ.get("B200IQ_CONCURRENT", "1") == "1"
# weight-gradient GEMMs of a block on the side stream, under the block's dX chain: the persistent GEMM grids leave SMs idle in
# their last wave (80 pair tiles on 74 SM pairs for every M=10240, N=512 product) and the side kernels' CTAs fill them
# B200IQ_PDL=3 (default): programmatic dependent launch for the decoder's forward only — the one phase with a single active
# stream, where a dependent kernel's waiting CTAs take SMs nobody else wants (12.47 -> 12.41 ms; =1 / =2 widen it to the whole
# step / the whole forward and lose 1.2 / 0.5 ms to the starved side streams, see iq.py; =0 off)
PDL_DECODER = os.environ.get("B200IQ_PDL", "3") == "3"
CONCURRENT_DW = CONCURRENT and os.environ.get("B200IQ_CONCURRENT_DW", "1") == "1"
# q | k | v (and the attention output's gradient) travel from the projection GEMMs to the attention kernels as fp16 planes
# (hi + lo: the same bytes as fp32; hi only where the backward reads them) whenever the fp16 mma.sync attention kernels take the
# shape (head_dim 64, n <= 32): those kernels then stage their operands without a conversion pass.  B200IQ_ATT_PLANES=0: fp32.
ATT_PLANES = os.environ.get("B200IQ_ATT_PLANES", "1") == "1"


def att_planes(c, nq, nk):
    return ATT_PLANES and not ops.exact_mode() and c.d == 64 and nq <= 32 and nk <= 32


class Workspace:
    """name -> persistent device buffer (stable addresses, so a captured CUDA graph can be replayed)"""

    def __init__(self, device):
        self.device = device
        self.bufs = {}
        self.meta = {}

    def buf(self, name, *shape, dtype=torch.float32):
        t = self.bufs.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.empty(shape, device=self.device, dtype=dtype)
            self.bufs[name] = t
        return t

    def __getitem__(self, name):
        return self.bufs[name]

    def act(self, name):
        """the operand made by ops.new_act(ws, name, ...) in the current mode"""
        if name in self.bufs:
            return ops.Act(f32=self.bufs[name])
        hi, lo = self.bufs[name + ".hi"], self.bufs.get(name + ".lo")
        K = self.meta[("actK", name)]
        return ops.Act(hi=hi[:, :K], lo=lo[:, :K] if lo is not None else None)

    def nbytes(self):
        return sum(t.numel() * t.element_size() for t in self.bufs.values())


class _PartMeta(dict):
    """meta of a WorkspacePart: private to the part (stack in / out tensors, dropout keys, side streams) except the operand widths
    ("actK", name), which describe the shared full-size buffers and are mirrored into the parent"""

    def __init__(self, parent_meta):
        super().__init__()
        self.parent_meta = parent_meta

    def __setitem__(self, k, v):
        if isinstance(k, tuple) and k and k[0] == "actK":
            self.parent_meta[k] = v
        super().__setitem__(k, v)


class WorkspacePart:
    """rows [part/nparts, (part+1)/nparts) of every buffer of a parent Workspace: the view a stack schedule gets when the batch
    is cut into `nparts` equal sample ranges that run as independent chains on their own streams (iq.py: the decoder's two
    half-batch passes).  Every buffer of a stack has the batch in its leading dimension ([B*n, ...], [B, ...]), so a part asks
    for `shape` and receives its slice of the parent's `nparts * shape[0]`-row buffer; whoever runs after the parts have joined
    reads the parent's full-size buffers as if one pass had written them."""

    def __init__(self, parent, part, nparts):
        self.parent, self.part, self.nparts = parent, part, nparts
        self.device = parent.device
        self.meta = _PartMeta(parent.meta)

    def _rows(self, t):
        n = t.shape[0] // self.nparts
        return t[self.part * n:(self.part + 1) * n]

    def buf(self, name, *shape, dtype=torch.float32):
        return self._rows(self.parent.buf(name, shape[0] * self.nparts, *shape[1:], dtype=dtype))

    def __getitem__(self, name):
        return self._rows(self.parent.bufs[name])

    def act(self, name):
        a = self.parent.act(name)
        return ops.Act(*(None if t is None else self._rows(t) for t in (a.f32, a.hi, a.lo)))


class WeightOperands:
    """name -> ops.Act of a weight matrix: the fp32 master view in exact mode, else views of the fp16 plane arenas (hi, lo) that
    mirror the parameter arena element for element; matrices whose row pitch is not a multiple of 8 halves (E = 300, L = 100)
    live in their own padded plane buffers (`odd`)"""

    def __init__(self, arena, whi=None, wlo=None, odd=None):
        self.arena, self.whi, self.wlo, self.odd = arena, whi, wlo, odd or {}
        self.vh = arena.views_over(whi) if whi is not None else None
        self.vl = arena.views_over(wlo) if wlo is not None else None
        self.cache = {}
        self.odd_spans = {}   # (o0, o1, cols) -> padded planes of a packed span whose pitch is not a multiple of 8 halves

    def __getitem__(self, name):
        a = self.cache.get(name)
        if a is None:
            if self.whi is None:
                a = ops.Act(f32=self.arena.views[name])
            elif name in self.odd:
                a = self.odd[name]
            else:
                a = ops.Act(hi=self.vh[name], lo=self.vl[name])
            self.cache[name] = a
        return a

    def span(self, first, last, cols):
        key = (first, last)
        a = self.cache.get(key)
        if a is None:
            o0, o1 = self.arena.span(first, last)
            if self.whi is None:
                a = ops.Act(f32=self.arena.flat.detach()[o0:o1].view(-1, cols))
            elif cols % 8:
                rows, ld = (o1 - o0) // cols, ops.pad8(cols)
                a = ops.Act(hi=torch.zeros(rows, ld, device=self.whi.device, dtype=torch.float16)[:, :cols],
                            lo=torch.zeros(rows, ld, device=self.whi.device, dtype=torch.float16)[:, :cols])
                self.odd_spans[(o0, o1, cols)] = a
                ops.split_f16(self.arena.flat.detach()[o0:o1].view(-1, cols), a.hi, a.lo)   # later steps: refresh_weight_operands
            else:
                a = ops.Act(hi=self.whi[o0:o1].view(-1, cols), lo=self.wlo[o0:o1].view(-1, cols))
            self.cache[key] = a
        return a


def build_weight_operands(arena, device, skip=()):
    """the fp16 plane arenas (hi, lo) mirroring a parameter arena, plus padded planes for 2-D weights whose row pitch is not a
    multiple of 8 halves (TMA wants 16-byte pitches); `skip`: 2-D tensors that never are GEMM operands (embedding tables)"""
    n = arena.flat.numel()
    whi = torch.zeros(n, device=device, dtype=torch.float16)
    wlo = torch.zeros(n, device=device, dtype=torch.float16)
    odd = {}
    for name, shape, _ in arena.specs:
        name = name.replace("#pack", "")
        if len(shape) == 2 and shape[1] % 8 and name not in skip:
            ld = ops.pad8(shape[1])
            odd[name] = ops.Act(hi=torch.zeros(shape[0], ld, device=device, dtype=torch.float16)[:, :shape[1]],
                                lo=torch.zeros(shape[0], ld, device=device, dtype=torch.float16)[:, :shape[1]])
    return WeightOperands(arena, whi, wlo, odd)


def refresh_weight_operands(w, arena, flat):
    """re-split the fp32 masters into the planes (once per step: 4 B read + 4 B written per parameter)"""
    ops.split_f16(flat, w.whi, w.wlo)
    for name, a in w.odd.items():
        ops.split_f16(arena.views[name], a.hi, a.lo)
    for (o0, o1, cols), a in w.odd_spans.items():
        ops.split_f16(flat[o0:o1].view(-1, cols), a.hi, a.lo)


class Ctx:
    """everything a block schedule needs: parameter / gradient views, workspace, sizes, dropout keys"""
    _count = 0

    def __init__(self, arena, gflat, ws, H, nh, F, E, attn_drop, relu_drop, training, seed, weights):
        self.arena, self.ws = arena, ws
        self.P = arena.views      # fp32 masters: biases, LayerNorm / BatchNorm affine, embedding table
        self.W = weights          # GEMM weight operands (WeightOperands)
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
        Ctx._count += 1
        self.seed_id = Ctx._count   # identity of this pass: blocks.act(cached=True) only reuses planes made under the same Ctx

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
        return self.W.span(first, last, cols)

    def gspan(self, first, last, cols):
        o0, o1 = self.arena.span(first, last)
        return self.gflat[o0:o1].view(-1, cols)

    def act(self, name, M, K, lo=True):
        """a forward operand buffer (planes hi + lo, or fp32 in exact mode)"""
        self.ws.meta[("actK", name)] = K
        return ops.new_act(self.ws, name, M, K, lo=lo)

    def gact(self, name, M, K):
        """a backward operand buffer: single fp16 plane (the gradients carry the loss scale), or fp32 in exact mode"""
        return self.act(name, M, K, lo=False)

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
    """dW[N,K] += dy[M,N]^T @ x[M,K];  db[N] += colsum(dy)   (dy, x: operands; single-pass product)"""
    M = dy.shape[0]
    ops.gemm2(dy, x, gw, a_t=True, b_t=True, accumulate=True, split_k=c.split_k(gw.shape[0], gw.shape[1], M), terms=1)
    if gb is not None:
        ops.colsum_add_act(dy, gb)


def ln_fwd(c, tag, pfx, x, kind="act", add=None, sum_out=None):
    """LayerNorm(x [+ add]) as a GEMM operand (kind "act": planes hi + lo, "act_hi": hi only) or as an fp32 tensor (kind "f32").
    With add, x + add is written to sum_out in the same pass (the residual add of the block in front of this LayerNorm)."""
    M, H = x.shape
    mean, rstd = c.ws.buf(tag + ".mean", M), c.ws.buf(tag + ".rstd", M)
    g, b = c.P[pfx + ".weight"], c.P[pfx + ".bias"]
    if kind == "f32":
        y = c.ws.buf(tag + ".y", M, H)
        ops.layernorm_fwd_act(x, g, b, mean, rstd, y=y, add=add, sum_out=sum_out)
        return y
    y = c.act(tag + ".y", M, H, lo=kind == "act")
    ops.layernorm_fwd_act(x, g, b, mean, rstd, act=y, add=add, sum_out=sum_out)
    return y


def ln_bwd(c, tag, pfx, dy, x, dx, accumulate, dx_act=None):
    """dx (+)= LayerNorm-backward(dy); dx_act (fp16-plane mode): also leave fp16(dx) in its hi plane for the GEMMs that read dx"""
    ops.layernorm_bwd_act(dy, x, c.P[pfx + ".weight"], c.ws[tag + ".mean"], c.ws[tag + ".rstd"], dx,
                          c.G[pfx + ".weight"], c.G[pfx + ".bias"], accumulate,
                          dx_hi=dx_act.hi if (dx_act is not None and dx_act.hi is not None) else None)


def dx_operand(c, name, dx):
    """the residual-stream gradient as a GEMM operand: the fp32 tensor itself in exact mode, else a single fp16 plane that the
    LayerNorm backward fills next to dx"""
    if ops.exact_mode():
        return ops.Act(f32=dx)
    return ops.Act(hi=c.ws.buf(name, dx.shape[0], dx.shape[1], dtype=torch.float16))


# ------------------------------------------------------------------ embedder (models/iq.py:57-79)
def embed_fwd(c, tag, ids, B, n, timing, add0=None, add1=None):
    M = B * n
    e = c.act(tag + ".e", M, c.E)
    ops.embed_gather_act(ids, c.P["embedding.0.weight"], e)
    x = c.ws.buf(tag + ".x0", M, c.H)
    ops.gemm2(e, c.W["embedding.1.weight"], x, bias=c.P["embedding.1.bias"])
    ops.add_timing(x, timing, add0, add1, B, n, c.H)
    return x


def embed_bwd(c, tag, ids, dxa):
    """dxa: the gradient wrt the embedder output as an operand (dx_operand)"""
    e = c.ws.act(tag + ".e")
    M = dxa.shape[0]
    ev = c.mark()
    de = c.ws.buf("tmp.de.%d" % M, M, c.E)
    ops.gemm2(dxa, c.W["embedding.1.weight"], de, b_t=True, terms=1)
    c.on_side(lambda: grad_w(c, dxa, e, c.G["embedding.1.weight"], c.G["embedding.1.bias"]), ev)
    ops.embed_scatter_add(ids, de, c.G["embedding.0.weight"], PAD)
    c.join_side()


# ------------------------------------------------------------------ attention blocks
def self_attn_fwd(c, tag, pfx, xn, ids, B, n, causal):
    """returns the block's output projection WITHOUT the residual: the caller's next LayerNorm adds it (ln_fwd add= / sum_out=),
    so the GEMM keeps its lean epilogue and the sum is written once, by the kernel that normalises it"""
    M, H = xn.shape
    wqkv = c.pspan(pfx + "query_linear.weight", pfx + "value_linear.weight", H)
    if att_planes(c, n, n):
        qkv = c.act(tag + ".qkv", M, 3 * H)
        ops.gemm2(xn, wqkv, planes=qkv)
    else:
        qkv = ops.Act(f32=c.ws.buf(tag + ".qkv", M, 3 * H))
        ops.gemm2(xn, wqkv, qkv.f32)
    ctx = c.act(tag + ".ctx", M, H)
    seed, off = c.key(tag + ".adrop")
    ops.attention_fwd(qkv.cols(0, H), qkv.cols(H, 2 * H), qkv.cols(2 * H, 3 * H), ctx, ids, PAD, causal, B, c.nh, n, n, c.d,
                      c.attn_drop, seed, off)
    proj = c.ws.buf(tag + ".proj", M, H)
    ops.gemm2(ctx, c.W[pfx + "output_linear.weight"], proj)
    return proj


def _dctx(c, dout, wo, M, H, planes):
    """gradient wrt the merged heads = dout . W_o: an fp16 hi plane when the attention backward takes planes, else fp32"""
    if planes:
        d = c.gact("tmp.dctx.%d" % M, M, H)
        ops.gemm2(dout, wo, planes=d, b_t=True, terms=1)
        return d
    d = c.ws.buf("tmp.dctx32.%d" % M, M, H)
    ops.gemm2(dout, wo, d, b_t=True, terms=1)
    return ops.Act(f32=d)


def _grad_parts(a, *col_ranges):
    """column slices of a gradient operand as the tensors attention_bwd writes (fp16 hi plane or fp32)"""
    t = a.f32 if a.f32 is not None else a.hi
    return [t[:, c0:c1] for c0, c1 in col_ranges]


def self_attn_bwd(c, tag, pfx, dout, xn, ids, B, n, causal):
    """dout: the block-output gradient as an operand; xn: the block's LayerNorm-ed input operand.  Returns fp32 grad wrt xn."""
    M, H = xn.shape
    qkv, ctx = c.ws.act(tag + ".qkv"), c.ws.act(tag + ".ctx")
    ev = c.mark()
    dctx = _dctx(c, dout, c.W[pfx + "output_linear.weight"], M, H, att_planes(c, n, n))
    c.on_side(lambda: grad_w(c, dout, ctx, c.G[pfx + "output_linear.weight"]), ev)
    dout_read = c.side_mark()
    dqkv = c.gact("tmp.dqkv.%d" % M, M, 3 * H)
    dq, dk, dv = _grad_parts(dqkv, (0, H), (H, 2 * H), (2 * H, 3 * H))
    seed, off = c.key(tag + ".adrop")
    ops.attention_bwd(qkv.cols(0, H), qkv.cols(H, 2 * H), qkv.cols(2 * H, 3 * H), dctx, dq, dk, dv,
                      ids, PAD, causal, B, c.nh, n, n, c.d, c.attn_drop, seed, off)
    ev = c.mark()
    dxn = c.ws.buf("tmp.dxn.%d" % M, M, H)
    ops.gemm2(dqkv, c.pspan(pfx + "query_linear.weight", pfx + "value_linear.weight", H), dxn, b_t=True, terms=1)
    c.on_side(lambda: grad_w(c, dqkv, xn, c.gspan(pfx + "query_linear.weight", pfx + "value_linear.weight", H)), ev)
    c.wait_side_mark(dout_read)   # the caller's LayerNorm backward accumulates into dout's buffer next; the stack joins
    return dxn                    # the rest of the side work at the end of the layer


def cross_kv_fwd(c, tag, pfx, enc_out, planes=False):
    """K | V of the context for one decoder layer: depends only on enc_out, so decoder_fwd computes all layers' up front"""
    wkv = c.pspan(pfx + "key_linear.weight", pfx + "value_linear.weight", c.H)
    if planes:
        kv = c.act(tag + ".kv", enc_out.shape[0], 2 * c.H)
        ops.gemm2(enc_out, wkv, planes=kv)
    else:
        kv = ops.Act(f32=c.ws.buf(tag + ".kv", enc_out.shape[0], 2 * c.H))
        ops.gemm2(enc_out, wkv, kv.f32)
    return kv


def cross_attn_fwd(c, tag, pfx, xn, enc_out, enc_ids, B, nq, nk, kv=None):
    """returns the output projection without the residual (see self_attn_fwd)"""
    M, H = xn.shape
    planes = att_planes(c, nq, nk)
    if planes:
        q = c.act(tag + ".q", M, H)
        ops.gemm2(xn, c.W[pfx + "query_linear.weight"], planes=q)
    else:
        q = ops.Act(f32=c.ws.buf(tag + ".q", M, H))
        ops.gemm2(xn, c.W[pfx + "query_linear.weight"], q.f32)
    if kv is None:
        kv = cross_kv_fwd(c, tag, pfx, enc_out, planes)
    ctx = c.act(tag + ".ctx", M, H)
    seed, off = c.key(tag + ".adrop")
    ops.attention_fwd(q, kv.cols(0, H), kv.cols(H, 2 * H), ctx, enc_ids, PAD, False, B, c.nh, nq, nk, c.d, c.attn_drop, seed, off)
    proj = c.ws.buf(tag + ".proj", M, H)
    ops.gemm2(ctx, c.W[pfx + "output_linear.weight"], proj)
    return proj


def cross_attn_bwd(c, tag, pfx, dout, xn, enc_out, d_enc_out, first, enc_ids, B, nq, nk):
    M, H = xn.shape
    Me = enc_out.shape[0]
    q, kv, ctx = c.ws.act(tag + ".q"), c.ws.act(tag + ".kv"), c.ws.act(tag + ".ctx")
    ev = c.mark()
    dctx = _dctx(c, dout, c.W[pfx + "output_linear.weight"], M, H, att_planes(c, nq, nk))
    c.on_side(lambda: grad_w(c, dout, ctx, c.G[pfx + "output_linear.weight"]), ev)
    dout_read = c.side_mark()
    dqa = c.gact("tmp.dq.%d" % M, M, H)
    dkva = c.gact("tmp.dkv.%d" % Me, Me, 2 * H)
    (dq,), (dk, dv) = _grad_parts(dqa, (0, H)), _grad_parts(dkva, (0, H), (H, 2 * H))
    seed, off = c.key(tag + ".adrop")
    ops.attention_bwd(q, kv.cols(0, H), kv.cols(H, 2 * H), dctx, dq, dk, dv, enc_ids, PAD, False, B, c.nh, nq, nk,
                      c.d, c.attn_drop, seed, off)
    def off_path():
        # the context-side 2560-row GEMMs (20-40 tiles) and the query weight gradient: none feeds the decoder's dX chain
        grad_w(c, dkva, enc_out, c.gspan(pfx + "key_linear.weight", pfx + "value_linear.weight", H))
        wkv = c.pspan(pfx + "key_linear.weight", pfx + "value_linear.weight", H)
        ops.gemm2(dkva, wkv, d_enc_out, b_t=True, residual=None if first else d_enc_out, terms=1)
        grad_w(c, dqa, xn, c.G[pfx + "query_linear.weight"])
    ev = c.mark()
    dxn = c.ws.buf("tmp.dxn.%d" % M, M, H)
    ops.gemm2(dqa, c.W[pfx + "query_linear.weight"], dxn, b_t=True, terms=1)
    c.on_side(off_path, ev)
    c.wait_side_mark(dout_read)
    return dxn


# ------------------------------------------------------------------ position-wise FFN (ReLU after BOTH linears, N1)
def ffn_fwd(c, tag, pfx, xn, x_res, next_ln=None):
    """relu + dropout after BOTH linears (N1) ride in the GEMM epilogues: h (operand planes) and f2 (fp32) are written once,
    post-dropout.  next_ln = (tag, prefix, kind) of the LayerNorm that reads this block's output (the next layer's first one, or
    the stack's final one): the residual add and that LayerNorm then run as one kernel and (out, LN(out)) is returned."""
    M, H = xn.shape
    h = c.act(tag + ".h", M, c.F)
    f2 = c.ws.buf(tag + ".f2", M, H)
    ops.gemm2(xn, c.W[pfx + "layers.0.weight"], planes=h, bias=c.P[pfx + "layers.0.bias"], relu=True,
              dropout=(c.relu_drop, *c.key(tag + ".d1")))
    ops.gemm2(h, c.W[pfx + "layers.1.weight"], f2, bias=c.P[pfx + "layers.1.bias"], relu=True,
              dropout=(c.relu_drop, *c.key(tag + ".d2")))
    out = c.ws.buf(tag + ".out", M, H)
    if next_ln is None:
        ops.axpby(x_res, f2, out)
        return out
    ln_tag, ln_pfx, kind = next_ln
    mean, rstd = c.ws.buf(ln_tag + ".mean", M), c.ws.buf(ln_tag + ".rstd", M)
    g, b = c.P[ln_pfx + ".weight"], c.P[ln_pfx + ".bias"]
    if kind == "f32":
        y = c.ws.buf(ln_tag + ".y", M, H)
        ops.layernorm_fwd_act(x_res, g, b, mean, rstd, y=y, add=f2, sum_out=out)
    else:
        y = c.act(ln_tag + ".y", M, H, lo=kind == "act")
        ops.layernorm_fwd_act(x_res, g, b, mean, rstd, act=y, add=f2, sum_out=out)
    return out, y


def ffn_bwd(c, tag, pfx, dout, xn):
    """dout: fp32 grad wrt (x_res + f2); returns fp32 grad wrt xn.  h and f2 are post-dropout, so (y > 0) is the relu mask AND
    the keep mask; the dropout backward is just the 1/(1-p) factor (no Philox in the backward)"""
    M, H = xn.shape
    h, f2 = c.ws.act(tag + ".h"), c.ws[tag + ".f2"]
    inv_keep = 1.0 / (1.0 - c.relu_drop)
    dpre2 = c.gact("tmp.dpre2.%d" % M, M, H)
    ops.relu_bwd_colsum_act(dout, f2, dpre2, c.G[pfx + "layers.1.bias"], scale=inv_keep)
    ev = c.mark()
    dh = c.gact("tmp.dh.%d" % M, M, c.F)
    ops.gemm2(dpre2, c.W[pfx + "layers.1.weight"], planes=dh, b_t=True, relu_mask=h, alpha=inv_keep, terms=1)
    c.on_side(lambda: grad_w(c, dpre2, h, c.G[pfx + "layers.1.weight"]), ev)
    ev = c.mark()
    dxn = c.ws.buf("tmp.dxn.%d" % M, M, H)
    ops.gemm2(dh, c.W[pfx + "layers.0.weight"], dxn, b_t=True, terms=1)
    c.on_side(lambda: grad_w(c, dh, xn, c.G[pfx + "layers.0.weight"], c.G[pfx + "layers.0.bias"]), ev)
    return dxn   # side work reads dpre2 / dh / saved activations only: joined by the stack at the end of the layer


# ------------------------------------------------------------------ stacks
def encoder_fwd(c, tag, pfx, x, ids, B, n, NL, out_kind="f32"):
    """models/transformer_layers.py:138-152 (timing already added by embed_fwd). Returns LN(x_NL) as `out_kind`."""
    c.ws.meta[tag + ".in"] = x
    xn = ln_fwd(c, tag + ".0.ln1", pfx + "enc.0.layer_norm_mha", x)
    for i in range(NL):
        lt, lp = "%s.%d" % (tag, i), "%senc.%d." % (pfx, i)
        y = self_attn_fwd(c, lt + ".sa", lp + "multi_head_attention.", xn, ids, B, n, False)
        a = c.ws.buf(lt + ".sa.out", x.shape[0], x.shape[1])
        xn2 = ln_fwd(c, lt + ".ln2", lp + "layer_norm_ffn", x, add=y, sum_out=a)   # a = x + attention block
        # the block's residual add and the LayerNorm that reads it (next layer's first, or the final one) are one kernel
        nxt = ("%s.%d.ln1" % (tag, i + 1), "%senc.%d.layer_norm_mha" % (pfx, i + 1), "act") if i + 1 < NL else (tag + ".lnf", pfx + "layer_norm", out_kind)
        x, xn = ffn_fwd(c, lt + ".ff", lp + "positionwise_feed_forward.", xn2, a, next_ln=nxt)
    c.ws.meta[tag + ".last"] = x
    return xn


def encoder_bwd(c, tag, pfx, dy, ids, B, n, NL):
    """dy: fp32 grad wrt the final LayerNorm output [M,H]. Returns the grad wrt the stack input as an operand (dx_operand)."""
    M, H = dy.shape
    dx = c.ws.buf(tag + ".dx", M, H)
    dxa = dx_operand(c, tag + ".dx.hi", dx)
    ln_bwd(c, tag + ".lnf", pfx + "layer_norm", dy, c.ws.meta[tag + ".last"], dx, False)
    for i in reversed(range(NL)):
        lt, lp = "%s.%d" % (tag, i), "%senc.%d." % (pfx, i)
        x_in = c.ws.meta[tag + ".in"] if i == 0 else c.ws["%s.%d.ff.out" % (tag, i - 1)]
        a = c.ws[lt + ".sa.out"]
        dxn2 = ffn_bwd(c, lt + ".ff", lp + "positionwise_feed_forward.", dx, c.ws.act(lt + ".ln2.y"))
        ln_bwd(c, lt + ".ln2", lp + "layer_norm_ffn", dxn2, a, dx, True, dx_act=dxa)
        dxn = self_attn_bwd(c, lt + ".sa", lp + "multi_head_attention.", dxa, c.ws.act(lt + ".ln1.y"), ids, B, n, False)
        ln_bwd(c, lt + ".ln1", lp + "layer_norm_mha", dxn, x_in, dx, True, dx_act=dxa if i == 0 else None)
        c.join_side()   # the layer's weight gradients are final, and the tmp.* buffers the side work read are free again
        c.ready(lp + "layer_norm_mha.weight")
    return dxa


def decoder_fwd(c, tag, pfx, x, ids, enc_out, enc_ids, B, T, S, NL, out_kind="act_hi", pdl=True):
    """models/transformer_layers.py:205-221; causal|pad mask of :207 is built inside the attention kernel.
    enc_out: the context as a GEMM operand."""
    c.ws.meta[tag + ".in"] = x
    # every layer's cross-attention K | V depends only on enc_out: six 2560-row GEMMs on the side stream, under layer 0's
    # self-attention block
    kvs = [None] * NL
    side = c.side_stream() if CONCURRENT else None
    if side is not None:
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), _lib.pinned_stream():
            for i in range(NL):
                kvs[i] = cross_kv_fwd(c, "%s.%d.ca" % (tag, i), "%sdec.%d.multi_head_attention_enc_dec." % (pfx, i), enc_out, att_planes(c, T, S))
    xn1 = ln_fwd(c, tag + ".0.ln1", pfx + "dec.0.layer_norm_mha_dec", x)
    pdl_prev = None
    for i in range(NL):
        lt, lp = "%s.%d" % (tag, i), "%sdec.%d." % (pfx, i)
        xn = xn1
        y = self_attn_fwd(c, lt + ".sa", lp + "multi_head_attention_dec.", xn, ids, B, T, True)
        a = c.ws.buf(lt + ".sa.out", x.shape[0], x.shape[1])
        xn = ln_fwd(c, lt + ".ln2", lp + "layer_norm_mha_enc", x, add=y, sum_out=a)
        if side is not None and i == 0:
            torch.cuda.current_stream().wait_stream(side)
        if PDL_DECODER and pdl and i == 0:
            pdl_prev = ops.set_pdl(True)   # from here to the end of the stack nothing runs beside the main stream
        y = cross_attn_fwd(c, lt + ".ca", lp + "multi_head_attention_enc_dec.", xn, enc_out, enc_ids, B, T, S, kv=kvs[i])
        b = c.ws.buf(lt + ".ca.out", x.shape[0], x.shape[1])
        xn = ln_fwd(c, lt + ".ln3", lp + "layer_norm_ffn", a, add=y, sum_out=b)
        nxt = ("%s.%d.ln1" % (tag, i + 1), "%sdec.%d.layer_norm_mha_dec" % (pfx, i + 1), "act") if i + 1 < NL else (tag + ".lnf", pfx + "layer_norm", out_kind)
        x, xn1 = ffn_fwd(c, lt + ".ff", lp + "positionwise_feed_forward.", xn, b, next_ln=nxt)
    c.ws.meta[tag + ".last"] = x
    if pdl_prev is not None:
        ops.set_pdl(pdl_prev)
    return xn1


def decoder_bwd(c, tag, pfx, dy, ids, enc_out, d_enc_out, enc_ids, B, T, S, NL):
    """returns the grad wrt the decoder input as an operand; writes (not accumulates) the sum of cross-attention grads into d_enc_out"""
    M, H = dy.shape
    dx = c.ws.buf(tag + ".dx", M, H)
    dxa = dx_operand(c, tag + ".dx.hi", dx)
    ln_bwd(c, tag + ".lnf", pfx + "layer_norm", dy, c.ws.meta[tag + ".last"], dx, False)
    for i in reversed(range(NL)):
        lt, lp = "%s.%d" % (tag, i), "%sdec.%d." % (pfx, i)
        x_in = c.ws.meta[tag + ".in"] if i == 0 else c.ws["%s.%d.ff.out" % (tag, i - 1)]
        a, b = c.ws[lt + ".sa.out"], c.ws[lt + ".ca.out"]
        dxn = ffn_bwd(c, lt + ".ff", lp + "positionwise_feed_forward.", dx, c.ws.act(lt + ".ln3.y"))
        ln_bwd(c, lt + ".ln3", lp + "layer_norm_ffn", dxn, b, dx, True, dx_act=dxa)
        dxn = cross_attn_bwd(c, lt + ".ca", lp + "multi_head_attention_enc_dec.", dxa, c.ws.act(lt + ".ln2.y"), enc_out,
                             d_enc_out, i == NL - 1, enc_ids, B, T, S)
        ln_bwd(c, lt + ".ln2", lp + "layer_norm_mha_enc", dxn, a, dx, True, dx_act=dxa)
        dxn = self_attn_bwd(c, lt + ".sa", lp + "multi_head_attention_dec.", dxa, c.ws.act(lt + ".ln1.y"), ids, B, T, True)
        ln_bwd(c, lt + ".ln1", lp + "layer_norm_mha_dec", dxn, x_in, dx, True, dx_act=dxa if i == 0 else None)
        c.join_side()
        c.ready(lp + "layer_norm_mha_dec.weight")
    return dx, dxa


# ------------------------------------------------------------------ small MLPs
def linear_fwd(c, tag, pfx, x, relu=False, out=None, planes=False, terms=3):
    """x: operand.  Returns the fp32 output tensor, or (planes=True) the output as an operand for the next GEMM"""
    N = c.P[pfx + ".weight"].shape[0]
    if planes:
        y = c.act(tag, x.shape[0], N)
        ops.gemm2(x, c.W[pfx + ".weight"], planes=y, bias=c.P[pfx + ".bias"], relu=relu, terms=terms)
        return y
    y = out if out is not None else c.ws.buf(tag, x.shape[0], N)
    ops.gemm2(x, c.W[pfx + ".weight"], y, bias=c.P[pfx + ".bias"], relu=relu, terms=terms)
    return y


def linear_bwd(c, tag, pfx, dy, x, relu_mask=None, need_dx=True, dx=None, accumulate_dx=False, planes=False):
    """dy: the grad wrt the (pre-activation) output as an operand; x: the layer's input operand; relu_mask = the post-relu INPUT
    operand when x came out of a relu.  Returns dx as an fp32 tensor, or (planes=True) as an operand for the next backward GEMM."""
    grad_w(c, dy, x, c.G[pfx + ".weight"], c.G[pfx + ".bias"])
    if not need_dx:
        return None
    K = c.P[pfx + ".weight"].shape[1]
    if planes:
        assert dx is None and not accumulate_dx
        dxa = c.gact(tag, dy.shape[0], K)
        ops.gemm2(dy, c.W[pfx + ".weight"], planes=dxa, b_t=True, relu_mask=relu_mask, terms=1)
        return dxa
    if dx is None:
        dx = c.ws.buf(tag, dy.shape[0], K)
    assert not (accumulate_dx and relu_mask is not None)
    ops.gemm2(dy, c.W[pfx + ".weight"], dx, b_t=True, relu_mask=relu_mask, residual=dx if accumulate_dx else None,
              auto_split=True, terms=1)   # e.g. z_classifier: 512 x 512 output over K = 15000 -> 4 tiles without split-K
    return dx
