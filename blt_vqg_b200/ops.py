"""Thin Python wrappers: one function per C-ABI entry point of libb200iq.so.

Every function launches on torch's current CUDA stream and returns immediately.
Tensors are fp32 CUDA tensors with unit inner stride; token ids are int32.
"""
import ctypes as C
import os

import torch

from . import _lib as L

# global switch: exact-fp32 CUDA-core GEMMs instead of tcgen05 TF32 (parity debugging, greedy decode)
_FORCE_SIMT = os.environ.get("B200IQ_GEMM", "") == "simt"
# fp32-in / fp32-out attention calls: 1 = precise mode (3-term forward products on unrounded inputs) — set by the transformer path
PRECISE_ATTENTION = 0
# device int64 word added to every Philox seed at run time (None: off).  A step captured in a CUDA graph keeps its launch
# arguments, so the per-step part of the dropout / eps keys lives here and the captured step increments it itself (train.py).
SEED_DEV = None
# number of kernel launches issued through the C ABI (bench.py reports it as gpu_launches)
LAUNCHES = 0
# when a list: gemm() brackets every launch with CUDA events on the launching stream and appends
# (start_event, end_event, flops, hbm_bytes, tag) — bench.py's per-kernel roofline pass
GEMM_TRACE = None


def exact_mode():
    """True when the library is in exact-fp32 mode (CUDA-core GEMMs, no TF32 rounding)"""
    return _FORCE_SIMT


# False: set_pdl(True) requests are ignored (train.TrainStep while it runs / captures a step for a CUDA graph)
PDL_ALLOWED = True


def set_pdl(on):
    """programmatic dependent launch for the hot kernels (b200iq_set_pdl); returns the previous setting"""
    return bool(L.lib().b200iq_set_pdl(1 if (on and PDL_ALLOWED) else 0))


class no_rounding:
    """context manager for the fp16-plane path (the transformer IQ): fp32 tensors stay unrounded — precision is decided per GEMM
    operand by its planes — so the library's TF32 write-rounding (the legacy kind::tf32 scheme the LSTM variant still uses) is off"""

    def __enter__(self):
        self.prev = L.lib().b200iq_set_tf32_rounding(0)

    def __exit__(self, *a):
        L.lib().b200iq_set_tf32_rounding(self.prev)


class Act:
    """a GEMM operand: exact mode -> `f32` tensor [M, K]; fp16-plane mode -> `hi` (+ `lo`) fp16 views [M, K] (row pitch a
    multiple of 8 halves) written by the producing kernel.  hi = fp16(x), lo = fp16(x - hi)."""
    __slots__ = ("f32", "hi", "lo")

    def __init__(self, f32=None, hi=None, lo=None):
        self.f32, self.hi, self.lo = f32, hi, lo

    @property
    def shape(self):
        return (self.f32 if self.f32 is not None else self.hi).shape

    def rows(self, r0, r1):
        return Act(*(None if t is None else t[r0:r1] for t in (self.f32, self.hi, self.lo)))

    def cols(self, c0, c1):
        return Act(*(None if t is None else t[:, c0:c1] for t in (self.f32, self.hi, self.lo)))


def pad8(n):
    return (n + 7) // 8 * 8


def new_act(ws, name, M, K, lo=True):
    """workspace-backed operand for the current mode (Workspace.buf keeps addresses stable)"""
    if _FORCE_SIMT:
        return Act(f32=ws.buf(name, M, K))
    ld = pad8(K)
    hi = ws.buf(name + ".hi", M, ld, dtype=torch.float16)[:, :K]
    return Act(hi=hi, lo=ws.buf(name + ".lo", M, ld, dtype=torch.float16)[:, :K] if lo else None)


def to_act(ws, name, x, lo=True, scale=1.0):
    """planes of an fp32 tensor [M, K] (b200iq_split_f16); exact mode: the tensor itself"""
    if _FORCE_SIMT:
        assert scale == 1.0
        return Act(f32=x)
    a = new_act(ws, name, x.shape[0], x.shape[1], lo=lo)
    split_f16(x, a.hi, a.lo, scale)
    return a


def split_f16(src, hi, lo=None, scale=1.0):
    """hi = fp16(scale*src), lo = fp16(scale*src - hi); src [rows, cols] (or flat), any row pitches"""
    if src.dim() == 1:
        rows, cols, lds, ldd = 1, src.numel(), src.numel(), hi.numel()
    else:
        rows, cols = src.shape
        lds, ldd = _ld(src), _ld(hi)
        assert lo is None or _ld(lo) == ldd
    _call("b200iq_split_f16", _p(src), _p(hi), _p(lo), rows, C.c_longlong(cols), C.c_longlong(lds), C.c_longlong(ldd), _f(scale), nbytes=_nb(src) + (2 if lo is not None else 1) * 2 * src.numel())


class GemmF16Params(C.Structure):
    _fields_ = [
        ("A_hi", C.c_void_p), ("A_lo", C.c_void_p), ("B_hi", C.c_void_p), ("B_lo", C.c_void_p),
        ("C", C.c_void_p), ("C_hi", C.c_void_p), ("C_lo", C.c_void_p),
        ("bias", C.c_void_p), ("residual", C.c_void_p), ("gate", C.c_void_p),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("lda", C.c_int), ("ldb", C.c_int), ("ldc", C.c_int), ("ldh", C.c_int), ("ldr", C.c_int), ("ldg", C.c_int),
        ("flags", C.c_int), ("split_k", C.c_int), ("alpha", C.c_float),
        ("drop_p", C.c_float), ("drop_seed", C.c_uint64), ("drop_offset", C.c_uint64), ("drop_seed_dev", C.c_void_p),
    ]


def _dp(t):
    return t.data_ptr() if t is not None else None


def gemm2(a, b, out=None, *, planes=None, a_t=False, b_t=False, bias=None, residual=None, relu=False, relu_mask=None,
          accumulate=False, split_k=1, alpha=1.0, dropout=None, auto_split=False, terms=3):
    """epi(alpha * A @ B^T) over operands `a`, `b` (Act).  out: fp32 tensor [M, N] and / or planes: Act receiving the result as
    operand planes for the next GEMM (exact mode: its fp32 tensor).  terms=3: the accurate 3-term product when both operands
    carry lo planes (forward products); terms=1: single pass (backward products, vocabulary projection).
    relu_mask: Act whose hi plane (> 0) gates the result."""
    if _FORCE_SIMT:
        dest = out if out is not None else planes.f32
        assert out is None or planes is None
        return gemm(a.f32, b.f32, dest, a_t=a_t, b_t=b_t, bias=bias, residual=residual, relu=relu,
                    relu_mask=relu_mask.f32 if relu_mask is not None else None, accumulate=accumulate, split_k=split_k,
                    alpha=alpha, fp32_out=True, dropout=dropout)
    ah, bh = a.hi, b.hi
    M, K = (ah.shape[1], ah.shape[0]) if a_t else ah.shape
    N, Kb = (bh.shape[1], bh.shape[0]) if b_t else bh.shape
    assert K == Kb, (ah.shape, bh.shape, a_t, b_t)
    three = terms == 3 and a.lo is not None and b.lo is not None and not a_t and not b_t
    if auto_split and not accumulate and residual is None and relu_mask is None and not relu and planes is None:
        # work units of the kernel's tile policy: CTA pairs of 256 x 256 wherever M > 128 and N > 64 (74 pairs), else 128 x 128 CTAs
        # (the vocabulary dX product, M = 1280, N = 512, K = 15000, is 10 pair tiles: 7 K-splits fill the pairs, 3 left 44 idle)
        if M > 128 and N > 64:
            tiles, slots = ((M + 255) // 256) * ((N + 255) // 256), 74
        else:
            tiles, slots = ((M + 127) // 128) * ((N + 127) // 128), 148
        sk = min(slots // tiles, K // 512) if tiles <= slots // 2 else 1
        if sk > 1:
            out.zero_()
            accumulate, split_k = True, sk
    flags = 0
    if a_t: flags |= L.GEMM_A_MN
    if b_t: flags |= L.GEMM_B_MN
    if bias is not None: flags |= L.GEMM_BIAS
    if relu: flags |= L.GEMM_RELU
    if residual is not None: flags |= L.GEMM_RESIDUAL
    if relu_mask is not None: flags |= L.GEMM_RELU_MASK
    if accumulate: flags |= L.GEMM_ACCUM
    dp, dseed, doff = dropout if dropout is not None and dropout[0] > 0 else (0.0, 0, 0)
    if dp > 0: flags |= L.GEMM_DROPOUT
    ph = planes.hi if planes is not None else None
    pl = planes.lo if planes is not None else None
    assert out is not None or ph is not None
    gate = relu_mask.hi if relu_mask is not None else None
    p = GemmF16Params(ah.data_ptr(), a.lo.data_ptr() if three else None, bh.data_ptr(), b.lo.data_ptr() if three else None,
                      _dp(out), _dp(ph), _dp(pl), _dp(bias), _dp(residual), _dp(gate), M, N, K,
                      _ld(ah), _ld(bh), _ld(out) if out is not None else 0, _ld(ph) if ph is not None else 0,
                      _ld(residual) if residual is not None else 0, _ld(gate) if gate is not None else 0,
                      flags, split_k, float(alpha), float(dp), dseed, doff, _dp(SEED_DEV) if dp > 0 else None)
    global LAUNCHES
    LAUNCHES += 1
    if GEMM_TRACE is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.check(L.lib().b200iq_gemm_f16(C.byref(p), L.stream_ptr()), "b200iq_gemm_f16")
        e1.record()
        nt = 3 if three else 1
        by = 2.0 * nt * (M * K + N * K) + (4.0 * M * N if out is not None else 0.0) + (2.0 * M * N * (2 if pl is not None else 1) if ph is not None else 0.0)
        # shared-memory fill of the launch (what the TMA producer pulls from L2): a CTA pair stages (256 + 256) x 64 halves per
        # k-block and plane (single CTAs: (128 + 128) x 64 per 128 x 128 tile), hi and lo planes for the 3-term product
        if M > 128 and N > 64:
            bn = 128 if N <= 128 else 256
            fill = ((M + 255) // 256) * ((N + bn - 1) // bn) * ((K + 63) // 64) * (256 + bn) * 64 * 2.0
        else:
            fill = ((M + 127) // 128) * ((N + 127) // 128) * ((K + 63) // 64) * 256 * 64 * 2.0
        fill *= 2 if three else 1
        GEMM_TRACE.append((e0, e1, 2.0 * M * N * K, by, (M, N, K, flags, split_k, nt), fill))
    else:
        L.check(L.lib().b200iq_gemm_f16(C.byref(p), L.stream_ptr()), "b200iq_gemm_f16")
    return out if out is not None else planes


class gemm_mode:
    """context manager: `with gemm_mode(simt=True): ...` runs everything in exact fp32 — CUDA-core GEMMs and no TF32
    rounding of activations (b200iq_set_tf32_rounding(0)); simt=False is the tensor-core path"""

    def __init__(self, simt):
        self.simt = simt

    def __enter__(self):
        global _FORCE_SIMT
        self.prev = _FORCE_SIMT
        _FORCE_SIMT = self.simt
        L.lib().b200iq_set_tf32_rounding(0 if self.simt else 1)

    def __exit__(self, *a):
        global _FORCE_SIMT
        _FORCE_SIMT = self.prev
        L.lib().b200iq_set_tf32_rounding(0 if self.prev else 1)


def _ld(t):
    assert t.dim() == 2 and t.stride(1) == 1, (t.shape, t.stride())
    return t.stride(0) if t.shape[0] > 1 else max(t.stride(0), t.shape[1])


def gemm(a, b, out=None, *, a_t=False, b_t=False, bias=None, residual=None, relu=False, relu_mask=None,
         accumulate=False, split_k=1, alpha=1.0, simt=None, fp32_out=False, auto_split=False, dropout=None):
    """out[M,N] = epi(alpha * A @ B^T); a is [M,K] ([K,M] if a_t), b is [N,K] ([K,N] if b_t).
    mirrors torch.nn.functional.linear(a, b, bias) (+ fused relu / residual / relu-backward mask)."""
    M, K = (a.shape[1], a.shape[0]) if a_t else a.shape
    N, Kb = (b.shape[1], b.shape[0]) if b_t else b.shape
    assert K == Kb, (a.shape, b.shape, a_t, b_t)
    if out is None:
        assert not accumulate
        out = torch.empty(M, N, device=a.device, dtype=torch.float32)
    assert out.shape == (M, N)
    if auto_split and not _FORCE_SIMT and not simt and not accumulate and residual is None and relu_mask is None and not relu:
        # skinny output, long reduction (dX of the vocabulary projection: 768 x 512 x 15000 = 24 tiles on 148 SMs):
        # zero C and let split-K slices TMA-reduce-add into it
        tiles = ((M + 127) // 128) * ((N + 127) // 128)
        sk = min(148 // tiles, K // 512) if tiles <= 74 else 1
        if sk > 1:
            out.zero_()
            accumulate, split_k = True, sk
    flags = 0
    if a_t: flags |= L.GEMM_A_MN
    if b_t: flags |= L.GEMM_B_MN
    if bias is not None: flags |= L.GEMM_BIAS
    if relu: flags |= L.GEMM_RELU
    if residual is not None: flags |= L.GEMM_RESIDUAL
    if relu_mask is not None:
        flags |= L.GEMM_RELU_MASK
        residual = relu_mask
    if accumulate: flags |= L.GEMM_ACCUM
    if fp32_out: flags |= L.GEMM_FP32_OUT
    dp, dseed, doff = dropout if dropout is not None and dropout[0] > 0 else (0.0, 0, 0)
    if dp > 0: flags |= L.GEMM_DROPOUT
    if simt if simt is not None else _FORCE_SIMT: flags |= L.GEMM_SIMT
    p = L.GemmParams(a.data_ptr(), b.data_ptr(), out.data_ptr(), bias.data_ptr() if bias is not None else None,
                     residual.data_ptr() if residual is not None else None, M, N, K,
                     _ld(a), _ld(b), _ld(out), _ld(residual) if residual is not None else 0,
                     flags, split_k, float(alpha), float(dp), dseed, doff)
    global LAUNCHES
    LAUNCHES += 1
    if GEMM_TRACE is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.check(L.lib().b200iq_gemm(C.byref(p), L.stream_ptr()), "b200iq_gemm")
        e1.record()
        GEMM_TRACE.append((e0, e1, 2.0 * M * N * K, 4.0 * (M * K + N * K + M * N), (M, N, K, flags, split_k)))
        return out
    L.check(L.lib().b200iq_gemm(C.byref(p), L.stream_ptr()), "b200iq_gemm")
    return out


# ------------------------------------------------------------------ helpers
def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _f(x):
    return C.c_float(float(x))


# when a list: every non-GEMM launch that states its algorithmic bytes is bracketed with CUDA events on the launching stream and
# appended as (start_event, end_event, kernel name, bytes) — bench.py's per-kernel roofline pass (HBM-bound kernels)
KERNEL_TRACE = None


def _nb(*ts):
    """bytes of the given tensors (None skipped): the algorithmic traffic of a streaming kernel = what it must read + write"""
    return sum(t.numel() * t.element_size() for t in ts if t is not None)


def _call(name, *args, nbytes=None):
    global LAUNCHES
    LAUNCHES += 1
    if KERNEL_TRACE is not None and nbytes is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.check(getattr(L.lib(), name)(*args, L.stream_ptr()), name)
        e1.record()
        KERNEL_TRACE.append((e0, e1, name, float(nbytes)))
        return
    L.check(getattr(L.lib(), name)(*args, L.stream_ptr()), name)


class AttnParams(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p), ("out", C.c_void_p),
        ("ldq", C.c_int), ("ldk", C.c_int), ("ldv", C.c_int), ("ldo", C.c_int),
        ("key_ids", C.c_void_p), ("pad_id", C.c_int), ("causal", C.c_int),
        ("B", C.c_int), ("num_heads", C.c_int), ("nq", C.c_int), ("nk", C.c_int), ("head_dim", C.c_int),
        ("scale", C.c_float), ("drop_p", C.c_float), ("seed", C.c_uint64), ("offset", C.c_uint64),
        ("kv_batch_stride", C.c_longlong),
        ("precise", C.c_int), ("out_hi", C.c_void_p), ("out_lo", C.c_void_p), ("seed_dev", C.c_void_p),
        ("planes_in", C.c_int), ("q_lo", C.c_void_p), ("k_lo", C.c_void_p), ("v_lo", C.c_void_p),
    ]


# ------------------------------------------------------------------ row / elementwise
def embed_gather(ids, table, out):
    _call("b200iq_embed_gather", _p(ids), _p(table), _p(out), ids.numel(), table.shape[1], table.shape[0])
    return out


def embed_scatter_add(ids, dout, dtable, pad_id):
    _call("b200iq_embed_scatter_add", _p(ids), _p(dout), _p(dtable), ids.numel(), dtable.shape[1], dtable.shape[0], int(pad_id))


def add_timing(x, timing, add0, add1, B, n, H):
    _call("b200iq_add_timing", _p(x), _p(timing), _p(add0), _p(add1), B, n, H)


def tok0_gather(src, dst, B, n, H, accumulate=False):
    """dst[b, :H] (+)= src[b*n, :H]; dst may be a column slice of a wider [B, ld] buffer"""
    _call("b200iq_tok0_gather", _p(src), _p(dst), B, n, H, dst.stride(0), int(accumulate))


def tok0_scatter_add(src, dst, B, n, H):
    """dst[b*n, :H] += src[b, :H]"""
    _call("b200iq_tok0_scatter_add", _p(src), _p(dst), B, n, H, src.stride(0))


def layernorm_fwd(x, g, b, y, mean, rstd, eps=1e-5):
    M, H = x.shape
    _call("b200iq_layernorm_fwd", _p(x), _p(g), _p(b), _p(y), _p(mean), _p(rstd), M, H, _f(eps), nbytes=_nb(x, y))
    return y


def add_layernorm_supported(x):
    return x.shape[1] in (128, 256, 512, 1024)


def add_layernorm_fwd(x, add, out, g, b, y, mean, rstd, eps=1e-5):
    """out = x + add (rounded like axpby), y = LayerNorm(out) in one pass"""
    M, H = x.shape
    _call("b200iq_add_layernorm_fwd", _p(x), _p(add), _p(out), _p(g), _p(b), _p(y), _p(mean), _p(rstd), M, H, _f(eps), nbytes=_nb(x, add, out, y))
    return y


def embed_gather_act(ids, table, out):
    """out: Act [M, E] (planes in fp16-plane mode)"""
    if out.f32 is not None:
        return embed_gather(ids, table, out.f32)
    _call("b200iq_embed_gather_f16", _p(ids), _p(table), _p(out.hi), _p(out.lo), ids.numel(), table.shape[1], table.shape[0], _ld(out.hi))
    return out


def layernorm_fwd_act(x, g, b, mean, rstd, y=None, act=None, add=None, sum_out=None, eps=1e-5):
    """LayerNorm(x [+ add]) with any of: fp32 y, operand planes `act`; with add the sum is written to sum_out.
    Exact mode (act.f32): the fp32 kernels."""
    M, H = x.shape
    if act is not None and act.f32 is not None:
        assert y is None
        y, act = act.f32, None
    if act is None:
        if add is not None:
            if add_layernorm_supported(x):
                return add_layernorm_fwd(x, add, sum_out, g, b, y, mean, rstd, eps)
            axpby(x, add, sum_out)
            return layernorm_fwd(sum_out, g, b, y, mean, rstd, eps)
        return layernorm_fwd(x, g, b, y, mean, rstd, eps)
    _call("b200iq_layernorm_fwd_f16", _p(x), _p(add), _p(sum_out), _p(g), _p(b), _p(y), _p(act.hi), _p(act.lo), _ld(act.hi),
          _p(mean), _p(rstd), M, H, _f(eps), nbytes=_nb(x, add, sum_out if add is not None else None, y, act.hi, act.lo))
    return y


def layernorm_bwd_act(dy, x, g, mean, rstd, dx, dg, db, accumulate_dx, dx_hi=None):
    """layernorm_bwd that also leaves fp16(dx) in dx_hi (fp16 [M, H]) for the GEMMs that read the residual-stream gradient"""
    if dx_hi is None:
        return layernorm_bwd(dy, x, g, mean, rstd, dx, dg, db, accumulate_dx)
    M, H = x.shape
    assert dx_hi.stride(0) == H
    _call("b200iq_layernorm_bwd_f16", _p(dy), _p(x), _p(g), _p(mean), _p(rstd), _p(dx), _p(dx_hi), _p(dg), _p(db), M, H, int(accumulate_dx), nbytes=_nb(dy, x, dx, dx if accumulate_dx else None, dx_hi))


def relu_bwd_colsum_act(dy, y, dpre, dbias, scale=1.0):
    """dpre: Act; planes mode: hi = fp16(scale * dy * (y > 0))"""
    if dpre.f32 is not None:
        return relu_bwd_colsum(dy, y, dpre.f32, dbias, scale)
    M, N = y.shape
    assert _ld(dpre.hi) == N
    _call("b200iq_relu_bwd_colsum_f16", _p(dy), _p(y), _p(dpre.hi), _p(dbias), M, N, _f(scale), nbytes=_nb(dy, y, dpre.hi))


def colsum_add_act(x, out):
    """out[c] += column sums of an operand (its hi plane in fp16-plane mode)"""
    if x.f32 is not None:
        return colsum_add(x.f32, out)
    M, N = x.hi.shape
    _call("b200iq_colsum_add_f16", _p(x.hi), _p(out), M, N, _ld(x.hi), nbytes=_nb(x.hi))


def ce_rows_act(logits, targets, pad_id, count, coef, loss_sum, grad):
    """cross-entropy of a logits chunk; grad: Act receiving coef * (softmax - onehot) / count (planes mode: hi plane; exact mode:
    grad.f32 must be the logits tensor itself — in place)"""
    if grad is None or grad.f32 is not None:
        return ce_rows(logits, targets, pad_id, count, coef, loss_sum, write_grad=grad is not None)
    rows, V = logits.shape
    _call("b200iq_ce_rows_f16", _p(logits), logits.stride(0), _p(targets), rows, V, int(pad_id), _p(count), _f(coef), _p(loss_sum),
          _p(grad.hi), _ld(grad.hi), nbytes=_nb(logits, grad.hi))


def layernorm_bwd(dy, x, g, mean, rstd, dx, dg, db, accumulate_dx):
    M, H = x.shape
    _call("b200iq_layernorm_bwd", _p(dy), _p(x), _p(g), _p(mean), _p(rstd), _p(dx), _p(dg), _p(db), M, H, int(accumulate_dx), nbytes=_nb(dy, x, dx, dx if accumulate_dx else None))


def colsum_add(x, out):
    M, N = x.shape
    _call("b200iq_colsum_add", _p(x), _p(out), M, N, x.stride(0), nbytes=_nb(x))


def relu_bwd_colsum(dy, y, dpre, dbias, scale=1.0):
    """dpre = scale * dy * (y > 0); dbias += colsum(dpre)"""
    M, N = y.shape
    _call("b200iq_relu_bwd_colsum", _p(dy), _p(y), _p(dpre), _p(dbias), M, N, _f(scale), nbytes=_nb(dy, y, dpre))


def dropout_(x, p, seed, offset):
    if p > 0:
        if SEED_DEV is not None:
            _call("b200iq_dropout_dev", _p(x), C.c_size_t(x.numel()), _f(p), C.c_uint64(seed), C.c_uint64(offset), _p(SEED_DEV))
        else:
            _call("b200iq_dropout", _p(x), C.c_size_t(x.numel()), _f(p), C.c_uint64(seed), C.c_uint64(offset))


def randn_(out, seed, offset=0):
    if SEED_DEV is not None:
        _call("b200iq_randn_dev", _p(out), C.c_size_t(out.numel()), C.c_uint64(seed), C.c_uint64(offset), _p(SEED_DEV))
    else:
        _call("b200iq_randn", _p(out), C.c_size_t(out.numel()), C.c_uint64(seed), C.c_uint64(offset))
    return out


def round_tf32(src, dst):
    """dst = nearest-TF32(src) (no-op copy in exact mode is NOT done here: callers skip it)"""
    _call("b200iq_round_tf32", _p(src), _p(dst), C.c_size_t(src.numel()))
    return dst


def axpby(x, y, out, a=1.0, b=1.0):
    """out = a*x + b*y"""
    _call("b200iq_axpby", _p(x), _p(y), _p(out), C.c_size_t(x.numel()), _f(a), _f(b))
    return out


# ------------------------------------------------------------------ attention
def _attn_params(q, k, v, out, key_ids, pad_id, causal, B, nh, nq, nk, d, drop_p, seed, offset, kv_batch_stride=0):
    """q / k / v: fp32 tensors, or Acts whose fp16 planes the projection GEMM wrote (planes_in: the hi planes go in as q / k / v,
    row pitches in halves)"""
    planes = isinstance(q, Act) and q.hi is not None
    if planes:
        qh, kh, vh = q.hi, k.hi, v.hi
        p = AttnParams(qh.data_ptr(), kh.data_ptr(), vh.data_ptr(), out.data_ptr() if out is not None else 0,
                       qh.stride(0), kh.stride(0), vh.stride(0), out.stride(0) if out is not None else 0,
                       key_ids.data_ptr() if key_ids is not None else 0, pad_id, int(causal),
                       B, nh, nq, nk, d, d ** -0.5, drop_p, seed, offset, kv_batch_stride, 0, None, None,
                       _dp(SEED_DEV) if drop_p > 0 else None, 1, _dp(q.lo), _dp(k.lo), _dp(v.lo))
        return p
    if isinstance(q, Act):
        q, k, v = q.f32, k.f32, v.f32
    return AttnParams(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr() if out is not None else 0,
                      q.stride(0), k.stride(0), v.stride(0), out.stride(0) if out is not None else 0,
                      key_ids.data_ptr() if key_ids is not None else 0, pad_id, int(causal),
                      B, nh, nq, nk, d, d ** -0.5, drop_p, seed, offset, kv_batch_stride, 0, None, None,
                      _dp(SEED_DEV) if drop_p > 0 else None, 0, None, None, None)


def attention_fwd(q, k, v, out, key_ids, pad_id, causal, B, nh, nq, nk, d, drop_p=0.0, seed=0, offset=0, kv_batch_stride=0):
    """kv_batch_stride (elements) lets k / v be the first nk positions of a [B, max_len, H] KV cache.
    out: fp32 tensor, or an Act — in fp16-plane mode the merged heads then go out as planes and the products run in the
    precise (3-term) mode on unrounded fp32 q / k / v"""
    if isinstance(out, Act):
        if out.f32 is not None:
            return attention_fwd(q, k, v, out.f32, key_ids, pad_id, causal, B, nh, nq, nk, d, drop_p, seed, offset, kv_batch_stride)
        p = _attn_params(q, k, v, None, key_ids, pad_id, causal, B, nh, nq, nk, d, drop_p, seed, offset, kv_batch_stride)
        p.precise, p.out_hi, p.out_lo, p.ldo = 1, out.hi.data_ptr(), _dp(out.lo), _ld(out.hi)
        # algorithmic bytes: q, k, v rows of every head once (fp32, or hi + lo planes: the same 4 bytes) + the output planes
        _call("b200iq_attention_fwd", C.byref(p), nbytes=4 * B * nh * d * (nq + 2 * nk) + 2 * B * nq * nh * d * (2 if out.lo is not None else 1))
        return out
    p = _attn_params(q, k, v, out, key_ids, pad_id, causal, B, nh, nq, nk, d, drop_p, seed, offset, kv_batch_stride)
    p.precise = 0 if _FORCE_SIMT else PRECISE_ATTENTION
    _call("b200iq_attention_fwd", C.byref(p), nbytes=4 * B * nh * d * (2 * nq + 2 * nk))
    return out


def attention_bwd(q, k, v, dout, dq, dk, dv, key_ids, pad_id, causal, B, nh, nq, nk, d, drop_p=0.0, seed=0, offset=0):
    """dq / dk / dv: fp32 tensors, or fp16 tensors (the hi planes of the gradients: operands of the projections' backward GEMMs).
    q / k / v / dout may be Acts with fp16 planes (planes_in: all four, dout as its hi plane)"""
    p = _attn_params(q, k, v, None, key_ids, pad_id, causal, B, nh, nq, nk, d, drop_p, seed, offset)
    if p.planes_in:
        assert isinstance(dout, Act) and dout.hi is not None and dq.dtype == torch.float16
        p.ldo, p.precise = dout.hi.stride(0), 1
        _call("b200iq_attention_bwd_f16", C.byref(p), _p(dout.hi), _p(dq), _p(dk), _p(dv), dq.stride(0), dk.stride(0), dv.stride(0),
              nbytes=2 * B * nh * d * (2 * nq + 2 * nk + nk + nq) + 2 * B * nh * d * (nq + 2 * nk))
        return
    if isinstance(dout, Act):
        dout = dout.f32
    p.ldo = dout.stride(0)
    if dq.dtype == torch.float16:
        p.precise = 1
        _call("b200iq_attention_bwd_f16", C.byref(p), _p(dout), _p(dq), _p(dk), _p(dv), dq.stride(0), dk.stride(0), dv.stride(0),
              nbytes=4 * B * nh * d * (2 * nq + 2 * nk) + 2 * B * nh * d * (nq + 2 * nk))
        return
    p.precise = 0 if _FORCE_SIMT else PRECISE_ATTENTION
    _call("b200iq_attention_bwd", C.byref(p), _p(dout), _p(dq), _p(dk), _p(dv), dq.stride(0), dk.stride(0), dv.stride(0),
          nbytes=4 * B * nh * d * (3 * nq + 4 * nk))


# ------------------------------------------------------------------ losses / heads / optimizer
def count_nonpad(targets, pad_id, count):
    _call("b200iq_count_nonpad", _p(targets), targets.numel(), int(pad_id), _p(count))


def ce_rows(logits, targets, pad_id, count, coef, loss_sum, write_grad=True):
    rows, V = logits.shape
    _call("b200iq_ce_rows", _p(logits), logits.stride(0), _p(targets), rows, V, int(pad_id), _p(count), _f(coef), _p(loss_sum), int(write_grad), nbytes=_nb(logits) * (2 if write_grad else 1))


def ce_aux(z_logit, dz_logit, targets, pad_id, count, coef, loss_sum):
    B, V = z_logit.shape
    _call("b200iq_ce_aux", _p(z_logit), _p(dz_logit), _p(targets), B, targets.numel() // B, V, int(pad_id), _p(count), _f(coef), _p(loss_sum), nbytes=_nb(z_logit, dz_logit))


def latent_fwd(prior, post, eps, z, kld_sum):
    B, L = eps.shape
    _call("b200iq_latent_fwd", _p(prior), _p(post), _p(eps), _p(z), _p(kld_sum), B, L)


def latent_bwd(prior, post, eps, dz, kld_coef, dprior, dpost):
    """kld_coef: a float, or a device tensor whose first element holds it (steps captured in a CUDA graph)"""
    B, L = eps.shape
    if isinstance(kld_coef, torch.Tensor):
        _call("b200iq_latent_bwd_dev", _p(prior), _p(post), _p(eps), _p(dz), _p(kld_coef), _p(dprior), _p(dpost), B, L)
    else:
        _call("b200iq_latent_bwd", _p(prior), _p(post), _p(eps), _p(dz), _f(kld_coef), _p(dprior), _p(dpost), B, L)


def bn_fwd(x, g, b, y, save_mean, save_rstd, run_mean, run_var, training, eps=1e-5, momentum=0.01):
    B, H = x.shape
    _call("b200iq_bn_fwd", _p(x), _p(g), _p(b), _p(y), _p(save_mean), _p(save_rstd), _p(run_mean), _p(run_var), B, H,
          _f(eps), _f(momentum), int(training))


def bn_bwd(dy, x, g, save_mean, save_rstd, dx, dg, db):
    B, H = x.shape
    _call("b200iq_bn_bwd", _p(dy), _p(x), _p(g), _p(save_mean), _p(save_rstd), _p(dx), _p(dg), _p(db), B, H)


def mse(a, b, coef, loss_sum, da, db, accumulate_db=False):
    _call("b200iq_mse", _p(a), _p(b), C.c_size_t(a.numel()), _f(coef), _p(loss_sum), _p(da), _p(db), int(accumulate_db))


def set_hyper(hyper, lr, bc1, bc2, grad_scale):
    _call("b200iq_set_hyper", _p(hyper), _f(lr), _f(bc1), _f(bc2), _f(grad_scale))


_SUMSQ_SCRATCH = {}


def grad_sumsq(grad, out):
    key = (grad.device, L.stream_ptr().value)
    scratch = _SUMSQ_SCRATCH.get(key)
    if scratch is None:
        scratch = _SUMSQ_SCRATCH[key] = torch.zeros(2048, device=grad.device)
    _call("b200iq_grad_sumsq", _p(grad), C.c_size_t(grad.numel()), _p(out), _p(scratch), scratch.numel(), nbytes=_nb(grad))


def adam_step(param, grad, m, v, hyper, normsq, beta1=0.9, beta2=0.999, eps=1e-8, max_norm=5.0, zero_grad=True):
    _call("b200iq_adam_step", _p(param), _p(grad), _p(m), _p(v), C.c_size_t(param.numel()), _p(hyper), _p(normsq),
          _f(beta1), _f(beta2), _f(eps), _f(max_norm), int(zero_grad), nbytes=_nb(param, grad, m, v) + _nb(param, m, v) + (_nb(grad) if zero_grad else 0))


# ------------------------------------------------------------------ LSTM recurrence
def lstm_seq_supported(H):
    return bool(L.lib().b200iq_lstm_seq_supported(int(H))) and not _FORCE_SIMT


def lstm_seq_fwd(gates, whh, hbuf, cbuf, lengths):
    """all T steps in one launch; gates [T,B,4H] in: x-projection + biases, out: activated gates"""
    T, B, H4 = gates.shape
    _call("b200iq_lstm_seq_fwd", _p(gates), _p(whh), _p(hbuf), _p(cbuf), _p(lengths), T, B, H4 // 4, nbytes=2 * _nb(gates) + _nb(hbuf, cbuf))


_LSTM_WS = {}


def lstm_seq_bwd(gates, cbuf, whh, dh_last, d_out, dgates, dh0, dc0, lengths):
    """BPTT of lstm_seq_fwd in one launch; dh_last / d_out may be None"""
    T, B, H4 = gates.shape
    H = H4 // 4
    key = (gates.device, B, H)
    ws = _LSTM_WS.get(key)
    if ws is None:
        L.lib().b200iq_lstm_seq_bwd_workspace_bytes.restype = C.c_size_t
        ws = _LSTM_WS[key] = torch.empty(L.lib().b200iq_lstm_seq_bwd_workspace_bytes(B, H) // 4, device=gates.device)
    _call("b200iq_lstm_seq_bwd", _p(gates), _p(cbuf), _p(whh), _p(dh_last), _p(d_out), _p(dgates), _p(dh0), _p(dc0), _p(ws),
          _p(lengths), T, B, H, nbytes=_nb(gates, cbuf, d_out, dgates))


def lstm_seq_bwd_static(gates, cbuf, whh, dh_last, d_out, dgates, dh0, dc0, lengths):
    """lstm_seq_bwd without a workspace argument and without any allocation: the scratch is a 5 MiB file-scope device array of
    the library (B <= 256; larger batches raise).  For callers that cannot allocate (graph capture, serving loops).  Calls that
    overlap on the GPU share — and race on — that array."""
    T, B, H4 = gates.shape
    _call("b200iq_lstm_seq_bwd_static", _p(gates), _p(cbuf), _p(whh), _p(dh_last), _p(d_out), _p(dgates), _p(dh0), _p(dc0),
          _p(lengths), T, B, H4 // 4, nbytes=_nb(gates, cbuf, d_out, dgates))


def lstm_cell_fwd(gates, c_prev, h_prev, c_next, h_next, lengths, t):
    B, H = c_prev.shape
    _call("b200iq_lstm_cell_fwd", _p(gates), _p(c_prev), _p(h_prev), _p(c_next), _p(h_next), _p(lengths), int(t), B, H)


def lstm_cell_bwd(gates, c_prev, c_t, dh_run, dh_out, dc, dgates, lengths, t):
    B, H = c_prev.shape
    _call("b200iq_lstm_cell_bwd", _p(gates), _p(c_prev), _p(c_t), _p(dh_run), _p(dh_out), _p(dc), _p(dgates), _p(lengths), int(t), B, H)


def transpose_ids(ids, out, col0, T_out):
    B, T_in = ids.shape
    _call("b200iq_transpose_ids", _p(ids), _p(out), B, T_in, int(T_out), int(col0))
    return out


# ------------------------------------------------------------------ decode-time helpers
def log_softmax_rows_(x):
    rows, V = x.shape
    _call("b200iq_log_softmax_rows", _p(x), x.stride(0), rows, V)
    return x


def topk_softmax_rows(x, k, idx_out, val_out=None):
    rows, V = x.shape
    _call("b200iq_topk_softmax_rows", _p(x), x.stride(0), rows, V, int(k), _p(idx_out), _p(val_out))


def feature_gather(cache, idx, out, bad_rows=None):
    """out[b,:] = fp32(cache[idx[b],:]) — cache is [num_images, F] bf16 or fp32 (b200iq_feature_gather)"""
    assert cache.dtype in (torch.bfloat16, torch.float32) and idx.dtype == torch.int32 and out.dtype == torch.float32
    _call("b200iq_feature_gather", _p(cache), C.c_int(int(cache.dtype == torch.bfloat16)), _p(idx), _p(out), idx.numel(),
          cache.shape[1], cache.shape[0], _p(bad_rows))
    return out


def feature_store(feats, idx, cache, bad_rows=None):
    """cache[idx[b],:] = feats[b,:] (b200iq_feature_store)"""
    assert cache.dtype in (torch.bfloat16, torch.float32) and idx.dtype == torch.int32 and feats.dtype == torch.float32
    _call("b200iq_feature_store", _p(feats), _p(idx), _p(cache), C.c_int(int(cache.dtype == torch.bfloat16)), idx.numel(),
          cache.shape[1], cache.shape[0], _p(bad_rows))


def ids_narrow(src, dst):
    """dst (int32) = src (int64), elementwise (b200iq_ids_narrow)"""
    assert src.dtype == torch.int64 and dst.dtype == torch.int32 and src.numel() == dst.numel()
    _call("b200iq_ids_narrow", _p(src), _p(dst), C.c_size_t(src.numel()))
    return dst
