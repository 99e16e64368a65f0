"""GPU: fused attention, cross-entropy, latent tail, BatchNorm head, MSE and clip+Adam kernels vs torch autograd."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(built_lib, request):
    from blt_vqg_b200 import _lib, ops as o
    _lib.require_device()
    prev = _lib.lib().b200iq_set_tf32_rounding(0)  # kernels are checked in exact mode; rounding has its own test
    request.addfinalizer(lambda: _lib.lib().b200iq_set_tf32_rounding(prev))
    torch.manual_seed(1)
    return o


def rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


def _ref_attention(q, k, v, key_ids, causal, B, nh, nq, nk, d):
    """models/transformer_layers.py:494-526 with the finite -1e18 fill"""
    qh = q.view(B, nq, nh, d).permute(0, 2, 1, 3) * d ** -0.5
    kh = k.view(B, nk, nh, d).permute(0, 2, 1, 3)
    vh = v.view(B, nk, nh, d).permute(0, 2, 1, 3)
    logits = qh @ kh.transpose(-1, -2)
    mask = key_ids.eq(0)[:, None, None, :].expand(B, 1, nq, nk)
    if causal:
        mask = mask | torch.triu(torch.ones(nq, nk, dtype=torch.bool, device=q.device), 1)[None, None]
    logits = logits.masked_fill(mask, -1e18)
    return (torch.softmax(logits, -1) @ vh).permute(0, 2, 1, 3).reshape(B * nq, nh * d)


@pytest.mark.parametrize("nq,nk,causal", [(5, 5, False), (21, 21, False), (20, 20, True), (20, 5, False), (64, 64, True), (1, 1, True)])
def test_attention_fwd_bwd(ops, nq, nk, causal):
    B, nh, d = 7, 4, 16
    H = nh * d
    q = torch.randn(B * nq, H, device="cuda", requires_grad=True)
    k = torch.randn(B * nk, H, device="cuda", requires_grad=True)
    v = torch.randn(B * nk, H, device="cuda", requires_grad=True)
    ids = torch.randint(1, 50, (B, nk), device="cuda", dtype=torch.int32)
    for b in range(B):
        ids[b, max(1, nk - b):] = 0  # ragged padding
    if nk > 1:
        ids[0, 0] = 0  # key 0 padded: with the causal mask query 0 is FULLY masked -> uniform weights
    out = torch.empty(B * nq, H, device="cuda")
    ops.attention_fwd(q.detach(), k.detach(), v.detach(), out, ids, 0, causal, B, nh, nq, nk, d)
    ref = _ref_attention(q, k, v, ids, causal, B, nh, nq, nk, d)
    assert torch.isfinite(out).all()
    assert rel(out, ref) < 2e-5
    do = torch.randn(B * nq, H, device="cuda")
    ref.backward(do)
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    ops.attention_bwd(q.detach(), k.detach(), v.detach(), do, dq, dk, dv, ids, 0, causal, B, nh, nq, nk, d)
    assert rel(dq, q.grad) < 1e-4 and rel(dk, k.grad) < 1e-4 and rel(dv, v.grad) < 1e-4


def test_attention_dropout_is_consistent_between_fwd_and_bwd(ops):
    """with dropout the op is out = (P*mask/(1-p)) V; check bwd against finite differences of fwd under the same key"""
    B, nh, d, n = 2, 2, 8, 6
    H = nh * d
    q, k, v = (torch.randn(B * n, H, device="cuda") for _ in range(3))
    ids = torch.ones(B, n, device="cuda", dtype=torch.int32)
    do = torch.randn(B * n, H, device="cuda")

    def f(q_, k_, v_):
        o = torch.empty(B * n, H, device="cuda")
        ops.attention_fwd(q_, k_, v_, o, ids, 0, False, B, nh, n, n, d, 0.3, 99, 5 << 40)
        return float((o.double() * do.double()).sum())
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    ops.attention_bwd(q, k, v, do, dq, dk, dv, ids, 0, False, B, nh, n, n, d, 0.3, 99, 5 << 40)
    h = 1e-2
    for t, g in ((q, dq), (k, dk), (v, dv)):
        for idx in ((0, 0), (3, 5), (B * n - 1, H - 1)):
            tp, tm = t.clone(), t.clone()
            tp[idx] += h
            tm[idx] -= h
            args_p = [tp if x is t else x for x in (q, k, v)]
            args_m = [tm if x is t else x for x in (q, k, v)]
            fd = (f(*args_p) - f(*args_m)) / (2 * h)
            assert fd == pytest.approx(float(g[idx]), rel=2e-2, abs=2e-3)


@pytest.mark.parametrize("nq,nk,causal", [(5, 5, False), (21, 21, False), (20, 20, True), (20, 5, False), (32, 32, True), (1, 1, True), (17, 9, False),
                                             # 32 < n <= 80: the row-block kernels (cfg5: T = 64, R = 65, S = 5)
                                             (64, 64, True), (65, 65, False), (64, 5, False), (5, 64, False), (80, 80, True), (33, 40, False),
                                             (48, 33, True), (72, 80, False)])
@pytest.mark.parametrize("drop", [0.0, 0.3])
def test_attention_tensor_core_path(ops, nq, nk, causal, drop):
    """head_dim 64 with TF32 rounding on -> the warp-per-head mma.sync kernels; checked against torch autograd (drop=0) and
    against the exact SIMT kernels under the same Philox key (drop>0: same mask by construction)"""
    from blt_vqg_b200 import _lib
    B, nh, d = 9, 3, 64
    H = nh * d
    rt = lambda x: ops.round_tf32(x, torch.empty_like(x))
    q = rt(torch.randn(B * nq, H, device="cuda"))
    kv = rt(torch.randn(B * nk, 2 * H, device="cuda"))
    k, v = kv[:, :H], kv[:, H:]  # column slices of a packed buffer (ld = 2H), like the workspace
    do = rt(torch.randn(B * nq, H, device="cuda"))
    ids = torch.randint(1, 50, (B, nk), device="cuda", dtype=torch.int32)
    for b in range(B):
        ids[b, max(1, nk - b):] = 0
    if nk > 1:
        ids[0, 0] = 0
    key = (drop, 1234, 9 << 40)

    def run():
        out = torch.empty(B * nq, H, device="cuda")
        dq, dkv = torch.empty_like(q), torch.empty_like(kv)
        ops.attention_fwd(q, k, v, out, ids, 0, causal, B, nh, nq, nk, d, *key)
        ops.attention_bwd(q, k, v, do, dq, dkv[:, :H], dkv[:, H:], ids, 0, causal, B, nh, nq, nk, d, *key)
        return out, dq, dkv[:, :H].clone(), dkv[:, H:].clone()
    exact = run()                                   # rounding off: SIMT kernels
    prev = _lib.lib().b200iq_set_tf32_rounding(1)
    try:
        tc = run()                                  # rounding on + d=64: tensor-core kernels (n <= 32 in registers, n <= 80 row blocks)
    finally:
        _lib.lib().b200iq_set_tf32_rounding(prev)
    for a, b in zip(tc, exact):
        assert torch.isfinite(a).all()
        assert rel(a, b) < 2e-3                     # P / dS are rounded to TF32 (2^-11 relative) before the second MMA
    if drop == 0.0:
        qr, kr, vr = (t.clone().requires_grad_(True) for t in (q, k.contiguous(), v.contiguous()))
        ref = _ref_attention(qr, kr, vr, ids, causal, B, nh, nq, nk, d)
        ref.backward(do)
        assert rel(tc[0], ref) < 2e-3 and rel(tc[1], qr.grad) < 2e-3 and rel(tc[2], kr.grad) < 2e-3 and rel(tc[3], vr.grad) < 2e-3


@pytest.mark.parametrize("nq,nk,causal", [(20, 20, True), (21, 21, False), (20, 5, False), (5, 5, False), (32, 32, True), (1, 1, True),
                                             (17, 9, False), (9, 30, False), (64, 64, True)])
@pytest.mark.parametrize("drop", [0.0, 0.3])
def test_attention_backward_into_fp16_planes(ops, nq, nk, causal, drop):
    """b200iq_attention_fwd in precise mode (fp16 hi + lo planes out) and b200iq_attention_bwd_f16 (precise mode: unrounded fp32 q / k / v / dO in, gradients out as fp16 hi planes — the fp16
    mma.sync kernel for n <= 32, the row-block kernel above that) against fp64 autograd (drop = 0) and against the exact CUDA-core
    kernels under the same Philox key (drop > 0: same mask by construction).  Bar 2e-3: 11-bit operands of the single-pass
    backward products plus the 2^-12 rounding of the fp16 result."""
    B, nh, d = 9, 3, 64
    H = nh * d
    q = torch.randn(B * nq, H, device="cuda")
    kv = torch.randn(B * nk, 2 * H, device="cuda") + 1.0   # a common component in the keys (LayerNorm + bias-free projection)
    k, v = kv[:, :H], kv[:, H:]
    do = torch.randn(B * nq, H, device="cuda") * 0.05
    ids = torch.randint(1, 50, (B, nk), device="cuda", dtype=torch.int32)
    for b in range(B):
        ids[b, max(1, nk - b):] = 0
    if nk > 1:
        ids[0, 0] = 0
    key = (drop, 4321, 11 << 40)
    # the precise forward (3-term products) into hi + lo planes against the exact kernels' fp32 output under the same key
    act = ops.Act(hi=torch.full((B * nq, H), float("nan"), device="cuda", dtype=torch.float16),
                  lo=torch.full((B * nq, H), float("nan"), device="cuda", dtype=torch.float16))
    ops.attention_fwd(q, k, v, act, ids, 0, causal, B, nh, nq, nk, d, *key)
    o32 = torch.empty(B * nq, H, device="cuda")
    ops.attention_fwd(q, k, v, o32, ids, 0, causal, B, nh, nq, nk, d, *key)
    assert torch.isfinite(act.hi).all() and torch.isfinite(act.lo).all()
    assert rel(act.hi.double() + act.lo.double(), o32) < 5e-6
    dq, dkv = torch.full((B * nq, H), float("nan"), device="cuda", dtype=torch.float16), torch.full((B * nk, 2 * H), float("nan"), device="cuda", dtype=torch.float16)
    ops.attention_bwd(q, k, v, do, dq, dkv[:, :H], dkv[:, H:], ids, 0, causal, B, nh, nq, nk, d, *key)
    eq, ekv = torch.empty(B * nq, H, device="cuda"), torch.empty(B * nk, 2 * H, device="cuda")
    ops.attention_bwd(q, k, v, do, eq, ekv[:, :H], ekv[:, H:], ids, 0, causal, B, nh, nq, nk, d, *key)   # rounding off, fp32 out: exact kernels
    for got, want in ((dq, eq), (dkv[:, :H], ekv[:, :H]), (dkv[:, H:], ekv[:, H:])):
        assert torch.isfinite(got).all()
        assert rel(got, want) < 2e-3
    if drop == 0.0:
        qr, kr, vr = (t.double().clone().requires_grad_(True) for t in (q, k.contiguous(), v.contiguous()))
        _ref_attention(qr, kr, vr, ids, causal, B, nh, nq, nk, d).backward(do.double())
        assert rel(dq, qr.grad) < 2e-3 and rel(dkv[:, :H], kr.grad) < 2e-3 and rel(dkv[:, H:], vr.grad) < 2e-3
    if nq <= 32 and nk <= 32:
        # planes_in: the same operands handed over as the fp16 planes a projection GEMM would have written (hi = fp16(x),
        # lo = fp16(x - hi)) — the kernels stage them without converting and must produce the same bits
        def planes(x, lo=True):
            hi = x.half()
            return ops.Act(hi=hi, lo=(x - hi.float()).half() if lo else None)
        qa, kva, doa = planes(q), planes(kv), planes(do, lo=False)
        ka, va = kva.cols(0, H), kva.cols(H, 2 * H)
        act2 = ops.Act(hi=torch.zeros_like(act.hi), lo=torch.zeros_like(act.lo))
        ops.attention_fwd(qa, ka, va, act2, ids, 0, causal, B, nh, nq, nk, d, *key)
        assert torch.equal(act2.hi, act.hi) and torch.equal(act2.lo, act.lo)
        dq2, dkv2 = torch.zeros_like(dq), torch.zeros_like(dkv)
        ops.attention_bwd(qa, ka, va, doa, dq2, dkv2[:, :H], dkv2[:, H:], ids, 0, causal, B, nh, nq, nk, d, *key)
        assert torch.equal(dq2, dq) and torch.equal(dkv2, dkv)


def test_ce_rows_and_aux(ops):
    rows, V, T = 300, 1000, 10
    logits = torch.randn(rows, V, device="cuda") * 3
    tg = torch.randint(1, V, (rows,), device="cuda", dtype=torch.int32)
    tg[::3] = 0
    lref = logits.clone().requires_grad_(True)
    loss_ref = F.cross_entropy(lref, tg.long(), ignore_index=0)
    (0.7 * loss_ref).backward()
    count, loss = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
    ops.count_nonpad(tg, 0, count)
    assert float(count) == float((tg != 0).sum())
    work = logits.clone()
    ops.ce_rows(work, tg, 0, count, 0.7, loss, write_grad=True)
    assert float(loss) == pytest.approx(float(loss_ref), rel=1e-5)
    assert rel(work, lref.grad) < 1e-5
    # V too large for the shared-memory row cache
    big = torch.randn(8, 70000, device="cuda")
    tb = torch.randint(1, 70000, (8,), device="cuda", dtype=torch.int32)
    cb, lb = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
    ops.count_nonpad(tb, 0, cb)
    ref_b = F.cross_entropy(big, tb.long())
    ops.ce_rows(big.clone(), tb, 0, cb, 1.0, lb, write_grad=False)
    assert float(lb) == pytest.approx(float(ref_b), rel=1e-5)
    # aux: CE(repeat(z_logit, T), target)   train_iq.py:92-94
    B = rows // T
    zl = (torch.randn(B, V, device="cuda") * 2).requires_grad_(True)
    aux_ref = F.cross_entropy(zl.unsqueeze(1).repeat(1, T, 1).reshape(-1, V), tg.long(), ignore_index=0)
    (1.3 * aux_ref).backward()
    dz, la = torch.empty(B, V, device="cuda"), torch.zeros(1, device="cuda")
    ops.ce_aux(zl.detach(), dz, tg.view(B, T), 0, count, 1.3, la)
    assert float(la) == pytest.approx(float(aux_ref), rel=1e-5)
    assert rel(dz, zl.grad) < 1e-5


def test_latent_tail(ops):
    B, L = 33, 20
    prior = (torch.randn(B, 2 * L, device="cuda") * 0.5).requires_grad_(True)
    post = (torch.randn(B, 2 * L, device="cuda") * 0.5).requires_grad_(True)
    eps = torch.randn(B, L, device="cuda")
    mp, lp, mq, lq = prior[:, :L], prior[:, L:], post[:, :L], post[:, L:]
    z_ref = eps * torch.exp(0.5 * lq) + mq
    kld_ref = (-0.5 * torch.sum(1 + (lq - lp) - (mp - mq) ** 2 / torch.exp(lp) - torch.exp(lq) / torch.exp(lp), -1)).mean()
    dz = torch.randn(B, L, device="cuda")
    ((z_ref * dz).sum() + 0.37 * kld_ref).backward()
    z, kld = torch.empty(B, L, device="cuda"), torch.zeros(1, device="cuda")
    ops.latent_fwd(prior.detach(), post.detach(), eps, z, kld)
    assert rel(z, z_ref) < 1e-6 and float(kld) == pytest.approx(float(kld_ref), rel=1e-5)
    dprior, dpost = torch.empty(B, 2 * L, device="cuda"), torch.empty(B, 2 * L, device="cuda")
    ops.latent_bwd(prior.detach(), post.detach(), eps, dz, 0.37, dprior, dpost)
    assert rel(dprior, prior.grad) < 1e-5 and rel(dpost, post.grad) < 1e-5
    zp = torch.empty(B, L, device="cuda")
    ops.latent_fwd(prior.detach(), None, eps, zp, None)  # prior-only sampling of decode_greedy (iq.py:126)
    assert rel(zp, eps * torch.exp(0.5 * lp) + mp) < 1e-6


def test_batchnorm_head(ops):
    B, H = 50, 70
    x = (torch.randn(B, H, device="cuda") * 2 + 1).requires_grad_(True)
    bn = torch.nn.BatchNorm1d(H, momentum=0.01).cuda()
    bn.weight.data.uniform_(0.5, 1.5)
    bn.bias.data.normal_()
    y_ref = bn(x)
    dy = torch.randn(B, H, device="cuda")
    y_ref.backward(dy)
    y, sm, sr = torch.empty(B, H, device="cuda"), torch.empty(H, device="cuda"), torch.empty(H, device="cuda")
    rm, rv = torch.zeros(H, device="cuda"), torch.ones(H, device="cuda")
    ops.bn_fwd(x.detach(), bn.weight.data, bn.bias.data, y, sm, sr, rm, rv, True)
    assert rel(y, y_ref) < 1e-5 and rel(rm, bn.running_mean) < 1e-5 and rel(rv, bn.running_var) < 1e-5
    dx, dg, db = torch.empty(B, H, device="cuda"), torch.zeros(H, device="cuda"), torch.zeros(H, device="cuda")
    ops.bn_bwd(dy, x.detach(), bn.weight.data, sm, sr, dx, dg, db)
    assert rel(dx, x.grad) < 1e-4 and rel(dg, bn.weight.grad) < 1e-5 and rel(db, bn.bias.grad) < 1e-5
    bn.eval()
    ops.bn_fwd(x.detach(), bn.weight.data, bn.bias.data, y, sm, sr, rm, rv, False)
    assert rel(y, bn(x)) < 1e-5


def test_mse(ops):
    a = torch.randn(40, 64, device="cuda", requires_grad=True)
    b = torch.randn(40, 64, device="cuda", requires_grad=True)
    (0.1 * F.mse_loss(a, b)).backward()
    loss, da, db = torch.zeros(1, device="cuda"), torch.zeros(40, 64, device="cuda"), torch.empty(40, 64, device="cuda")
    ops.mse(a.detach(), b.detach(), 0.1, loss, da, db)
    assert float(loss) == pytest.approx(float(F.mse_loss(a, b)), rel=1e-5)
    assert rel(da, a.grad) < 1e-5 and rel(db, b.grad) < 1e-5


@pytest.mark.parametrize("scale", [0.01, 10.0])
def test_clip_and_adam_match_torch(ops, scale):
    n = 100003
    p0, g0 = torch.randn(n + 1, device="cuda")[:n], torch.randn(n + 1, device="cuda")[:n] * scale
    pr = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([pr], lr=3e-4)
    p, m, v = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    hyper, nsq = torch.zeros(4, device="cuda"), torch.zeros(1, device="cuda")
    for t in range(1, 4):
        pr.grad = g0.clone() * t
        torch.nn.utils.clip_grad_norm_([pr], 5.0)
        opt.step()
        g = g0.clone() * t
        hyper.copy_(torch.tensor([3e-4, 1 - 0.9 ** t, 1 - 0.999 ** t, 1.0]))
        ops.grad_sumsq(g, nsq)
        assert float(nsq) == pytest.approx(float(g.double().pow(2).sum()), rel=1e-5)
        ops.adam_step(p, g, m, v, hyper, nsq)
        assert float(g.abs().max()) == 0.0  # zero_grad fused
        assert rel(p, pr.detach()) < 1e-6
