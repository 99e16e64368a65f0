"""Linear-layer helpers over fp32 tensors for the LSTM variant (lstm_iq.py), on the same fp16-plane GEMM as the transformer.

The transformer schedule (transformer.py) lets every producer kernel write its output directly as operand planes.  The LSTM
variant's tensors are small (T*B = 4864 rows of H = 300), so here the fp32 tensor stays the interface and `act()` makes the
planes with one b200iq_split_f16 pass per use; each fp32 buffer owns a plane buffer of its own (keyed by address), so a
re-split on one stream while another stream still reads the planes rewrites identical bytes.
Forward products that feed a ReLU gate run 3-term (hi.hi + lo.hi + hi.lo), backward products single-pass under the loss scale.
"""
import torch

from . import ops
from .transformer import CONCURRENT_DW, Ctx, WeightOperands, Workspace, build_weight_operands, refresh_weight_operands  # noqa: F401


def act(c, x, lo=True, cached=False):
    """GEMM operand of the fp32 tensor x [M, K]: its planes (exact mode: x itself).  cached=True: x has not been written since
    its planes were last made (an activation the forward pass already split and the backward pass reads again) — reuse them."""
    if isinstance(x, ops.Act):
        return x
    if ops.exact_mode():
        return ops.Act(f32=x)
    key = ("planes", x.data_ptr(), tuple(x.shape), x.stride(0))
    a = c.ws.meta.get(key)
    if cached and a is not None and (a.lo is not None or not lo) and c.ws.meta.get(("planes_seed", key)) == c.seed_id:
        return a if lo or a.lo is None else ops.Act(hi=a.hi)
    c.ws.meta[("planes_seed", key)] = c.seed_id
    if a is None or (lo and a.lo is None):
        M, K = x.shape
        ld = ops.pad8(K)
        a = ops.Act(hi=torch.zeros(M, ld, device=x.device, dtype=torch.float16)[:, :K],
                    lo=torch.zeros(M, ld, device=x.device, dtype=torch.float16)[:, :K] if lo else None)
        c.ws.meta[key] = a
    ops.split_f16(x, a.hi, a.lo if lo else None)
    return a if lo or a.lo is None else ops.Act(hi=a.hi)


def mm(c, x, w, out, *, b_t=False, bias=None, residual=None, relu=False, relu_mask=None, auto_split=False, terms=3):
    """out = epi(x @ W^T) (b_t: x @ W, the dX form — single pass).  x, relu_mask: fp32 tensors or operands; w: a weight name,
    or an operand (Ctx.pspan)"""
    wa = c.W[w] if isinstance(w, str) else w
    terms = 1 if b_t else terms
    xa = act(c, x, lo=terms == 3)
    gate = act(c, relu_mask, lo=False) if relu_mask is not None else None
    return ops.gemm2(xa, wa, out, b_t=b_t, bias=bias, residual=residual, relu=relu, relu_mask=gate, auto_split=auto_split, terms=terms)


def grad_w(c, dy, x, gw, gb=None, dya=None, x_cached=False):
    """dW[N,K] += dy[M,N]^T @ x[M,K];  db[N] += colsum(dy).  dya: the planes of dy when the caller has made them already (several
    products read the same gradient); x_cached: x is a saved activation whose planes the forward pass made (act cached=)"""
    dya = dya if dya is not None else act(c, dy, lo=False)
    xa = act(c, x, lo=False, cached=x_cached)
    M = dya.shape[0]
    ops.gemm2(dya, xa, gw, a_t=True, b_t=True, accumulate=True, split_k=c.split_k(gw.shape[0], gw.shape[1], M), terms=1)
    if gb is not None:
        if isinstance(dy, ops.Act):
            ops.colsum_add_act(dy, gb)
        else:
            ops.colsum_add(dy, gb)


def linear_fwd(c, tag, pfx, x, relu=False, out=None, fp32_out=False, terms=3):
    y = out if out is not None else c.ws.buf(tag, x.shape[0], c.P[pfx + ".weight"].shape[0])
    mm(c, x, pfx + ".weight", y, bias=c.P[pfx + ".bias"], relu=relu, terms=terms)
    return y


def linear_bwd(c, tag, pfx, dy, x, relu_mask=None, need_dx=True, dx=None, accumulate_dx=False):
    """dy is the grad wrt the (pre-activation) output; relu_mask = post-relu INPUT x when x came out of a relu"""
    dya = act(c, dy, lo=False)   # one split for the weight-gradient and the input-gradient product
    grad_w(c, dy, x, c.G[pfx + ".weight"], c.G[pfx + ".bias"], dya=dya, x_cached=True)
    if not need_dx:
        return None
    if dx is None:
        dx = c.ws.buf(tag, dy.shape[0], c.P[pfx + ".weight"].shape[1])
    assert not (accumulate_dx and relu_mask is not None)
    gate = act(c, relu_mask, lo=False, cached=True) if relu_mask is not None else None
    mm(c, dya, pfx + ".weight", dx, b_t=True, relu_mask=gate, residual=dx if accumulate_dx else None, auto_split=True)
    return dx
