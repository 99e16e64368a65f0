"""CPU study (no GPU): which GEMM operand roundings cost how much gradient accuracy on the IQ transformer step.

The fp64 oracle is run with every `F.linear` (and the two attention products) replaced by an autograd function that rounds
its operands to a given significand width per product class:
  fwd   Y  = X W^T          dx   dX = dY W          dw   dW = dY^T X
  att   QK^T, PV and their four backward products
A policy gives, per class, the significand bits kept for the activation-side operand and for the weight-side operand
(0 = exact).  11 bits = TF32 / fp16 (round to nearest), 8 bits = bf16, 22 bits = a two-term fp16 (hi + lo) split.
Output: global and worst-tensor relative gradient error against the clean fp64 run.

    python scripts/cpu_precision_study.py [cfg] [B] [NL]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from blt_vqg_b200 import synth
from oracle import iq_oracle

REAL_LINEAR = F.linear


def rnd(x, bits):
    """round to nearest with `bits` significant bits (no exponent range limit)"""
    if not bits:
        return x
    m, e = torch.frexp(x)
    s = float(1 << bits)
    return torch.ldexp(torch.round(m * s) / s, e)


POLICY = {}
STATS = {}


class QLinear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        p = POLICY
        ctx.save_for_backward(x, w)
        ctx.has_b = b is not None
        y = rnd(x, p["fwd"][0]) @ rnd(w, p["fwd"][1]).t()
        return y + b if b is not None else y

    @staticmethod
    def backward(ctx, dy):
        p = POLICY
        x, w = ctx.saved_tensors
        dx = rnd(dy, p["dx"][0]) @ rnd(w, p["dx"][1])
        dy2, x2 = dy.reshape(-1, dy.shape[-1]), x.reshape(-1, x.shape[-1])
        dw = rnd(dy2, p["dw"][0]).t() @ rnd(x2, p["dw"][1])
        db = dy2.sum(0) if ctx.has_b else None
        return dx, dw, db


class QMatmul(torch.autograd.Function):
    """a @ b with both operands rounded (attention products), backward products rounded the same way"""

    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        n = POLICY["att"]
        return rnd(a, n) @ rnd(b, n)

    @staticmethod
    def backward(ctx, dc):
        a, b = ctx.saved_tensors
        n = POLICY["att"]
        n2 = POLICY.get("att_bwd", n)
        return rnd(dc, n2) @ rnd(b, n2).transpose(-1, -2), rnd(a, n2).transpose(-1, -2) @ rnd(dc, n2)


def qlinear(x, w, b=None):
    return QLinear.apply(x, w, b)


def mha(P, pre, q_in, kv_in, mask, nh, drop_p=0.0, training=False):
    q = F.linear(q_in, P[pre + "query_linear.weight"])
    k = F.linear(kv_in, P[pre + "key_linear.weight"])
    v = F.linear(kv_in, P[pre + "value_linear.weight"])
    B, nq, H = q.shape
    nk, d = k.shape[1], H // nh
    q = q.view(B, nq, nh, d).permute(0, 2, 1, 3) * (d ** -0.5)
    k = k.view(B, nk, nh, d).permute(0, 2, 1, 3)
    v = v.view(B, nk, nh, d).permute(0, 2, 1, 3)
    logits = QMatmul.apply(q, k.transpose(-1, -2))
    logits = logits.masked_fill(mask.unsqueeze(1), -1e18)
    w = torch.softmax(logits, dim=-1)
    ctx = QMatmul.apply(w, v).permute(0, 2, 1, 3).reshape(B, nq, H)
    return F.linear(ctx, P[pre + "output_linear.weight"])


def grads(sd, cfg, batch, eps, policy):
    global POLICY
    POLICY = policy
    F.linear = qlinear
    iq_oracle.mha = mha
    try:
        ocfg = dict(num_layers=cfg["NL"], num_heads=cfg["NH"], latent_dim=cfg["L"], image_recon_lambda=0.1, kl_ceiling=0.5,
                    aux_ceiling=1.0, full_kl_step=15000)
        P = {k: v.double().clone() if v.dtype.is_floating_point else v.clone() for k, v in sd.items()}
        for k, v in P.items():
            if v.dtype.is_floating_point and "running" not in k:
                v.requires_grad_(True)
        bn = (P["encoder_cnn.bn.running_mean"].clone(), P["encoder_cnn.bn.running_var"].clone())
        out, zl, kld, img, recon = iq_oracle.iq_forward(P, ocfg, batch["feats"].double(), batch["answers"], batch["posteriors"],
                                                        batch["questions"], eps=eps.double(), latent_mode=True, training=True, bn_state=bn)
        loss, tr = iq_oracle.iq_losses(ocfg, out, zl, kld, img, recon, batch["questions"], 7000, True)
        loss.backward()
        return {k: v.grad.detach() for k, v in P.items() if v.requires_grad and v.grad is not None}, {k: float(v) for k, v in tr.items()}
    finally:
        F.linear = REAL_LINEAR


def report(tag, g, truth, lt, l0):
    errs = []
    num = den = 0.0
    for n, gt in truth.items():
        if n not in g or n == "encoder_cnn.cnn.fc.bias":
            continue
        d = (g[n] - gt).norm().item()
        r = gt.norm().item()
        num += d * d
        den += r * r
        errs.append((d / (r + 1e-30), n))
    errs.sort(reverse=True)
    over = sum(1 for e, _ in errs if e > 2e-3)
    lerr = max(abs(lt[k] - l0[k]) / (abs(l0[k]) + 1e-30) for k in ("rec", "img", "kld", "aux"))
    short = lambda n: n.replace("answer_encoder.", "ae.").replace("multi_head_attention", "mha").replace("positionwise_feed_forward", "ffn").replace("decoder.decoder", "dec")[-40:]
    print("%-34s global %.2e  loss %.1e  >2e-3: %3d/%d | worst %s" % (tag, (num / den) ** 0.5, lerr, over, len(errs),
          ", ".join("%s %.1e" % (short(n), e) for e, n in errs[:3])), flush=True)


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
    cfg = dict(synth.CONFIGS[name])
    cfg["B"] = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    if len(sys.argv) > 3:
        cfg["NL"] = int(sys.argv[3])
    torch.set_num_threads(os.cpu_count())
    batch = synth.make_batch(cfg)
    from blt_vqg_b200.iq import IQ
    torch.manual_seed(3)
    m0 = IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg))
    sd = {k: v.clone() for k, v in m0.state_dict().items()}
    names = set(m0.arena.names())
    eps = torch.randn(cfg["B"], cfg["L"])
    E = (0, 0)
    exact = dict(fwd=E, dx=E, dw=E, att=0)
    truth, l0 = grads(sd, cfg, batch, eps, exact)
    truth = {k: v for k, v in truth.items() if k in names}
    print("%s B=%d NL=%d loss %s" % (name, cfg["B"], cfg["NL"], l0))
    T, Hh, S2 = (11, 11), (8, 8), (22, 22)
    pols = [
        ("all 11-bit (TF32 / fp16 single)", dict(fwd=T, dx=T, dw=T, att=11)),
        ("all bf16", dict(fwd=Hh, dx=Hh, dw=Hh, att=8)),
        ("only fwd 11", dict(fwd=T, dx=E, dw=E, att=0)),
        ("only dx 11", dict(fwd=E, dx=T, dw=E, att=0)),
        ("only dw 11", dict(fwd=E, dx=E, dw=T, att=0)),
        ("only att 11", dict(fwd=E, dx=E, dw=E, att=11)),
        ("only att fwd 11", dict(fwd=E, dx=E, dw=E, att=11, att_bwd=0)),
        ("fwd,dx act split (w 11), dw 11", dict(fwd=(0, 11), dx=(0, 11), dw=T, att=11)),
        ("fwd,dx w split (act 11), dw 11", dict(fwd=(11, 0), dx=(11, 0), dw=T, att=11)),
        ("fwd,dx exact; dw 11; att 11", dict(fwd=E, dx=E, dw=T, att=11)),
        ("fwd,dx exact; dw 11; att exact", dict(fwd=E, dx=E, dw=T, att=0)),
        ("fwd,dx exact; dw bf16; att exact", dict(fwd=E, dx=E, dw=Hh, att=0)),
        ("fwd 11; dx exact; dw 11; att 11", dict(fwd=T, dx=E, dw=T, att=11)),
        ("fwd exact; dx 11; dw 11; att 11", dict(fwd=E, dx=T, dw=T, att=11)),
    ]
    if os.environ.get("STUDY") == "target":
        S = (22, 22)
        pols = [
            ("fwd 22; dx 11; dw 11; att 22/11", dict(fwd=S, dx=T, dw=T, att=22, att_bwd=11)),
            ("fwd 22; dx 11; dw 11; att 11/11", dict(fwd=S, dx=T, dw=T, att=11, att_bwd=11)),
            ("fwd 22; dx bf16; dw bf16; att 22/11", dict(fwd=S, dx=Hh, dw=Hh, att=22, att_bwd=11)),
            ("fwd 16(bf16x2); dx 11; dw 11; att 22/11", dict(fwd=(16, 16), dx=T, dw=T, att=22, att_bwd=11)),
        ]
    for tag, p in pols:
        g, lt = grads(sd, cfg, batch, eps, p)
        report(tag, g, truth, lt, l0)


if __name__ == "__main__":
    main()
