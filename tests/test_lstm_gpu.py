"""GPU parity of the LSTM variant (through the C ABI): training step vs the golden fixture built from the reference's own
DecoderRNN / EncoderRNN / MLP classes and vs the oracle on a larger ragged batch; staged API; bit-exact greedy ids.
Tolerances as in test_iq_gpu.py: exact path 1e-4 / 2e-4; tensor-core path (fp16 operand planes: 3-term forward products and
3-term TF32 recurrence, single-pass backward) losses 1e-3, gradients: see FULL_TOL / _check."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
FULL_TOL = (1e-3, 8e-2, 2e-2)   # (loss, per tensor, global) of the full-size tensor-core runs


def rel(a, b):
    return float((a.double().cpu() - b.double().cpu()).norm() / (b.double().cpu().norm() + 1e-30))


@pytest.fixture(scope="module")
def env(built_lib):
    from blt_vqg_b200 import _lib
    _lib.require_device()
    return True


def _build(cfg, sd):
    from blt_vqg_b200.lstm_iq import LSTMIQ
    m = LSTMIQ.from_config(cfg).cuda().train()
    m.load_state_dict(sd)
    return m


def _check(m, terms, want, grads, simt, tols=None):
    ltol, ttol, gtol = tols or ((1e-4, 2e-4, 2e-4) if simt else (1e-3, 8e-2, 2e-2))
    t = terms.cpu()
    worst = (0.0, "")
    for i, k in enumerate(("gen", "kl", "rec_i", "rec_a")):
        assert float(t[i]) == pytest.approx(float(want[k]), rel=ltol), k
    assert float(m.total_loss(t)) == pytest.approx(float(want["loss"]), rel=ltol)
    num = den = 0.0
    for k, gref in grads.items():
        g = m.arena.view_of(m.unscaled_grads(), k)
        if k == "encoder_cnn.cnn.fc.bias":  # exactly-zero gradient (bias in front of BatchNorm)
            assert float(g.abs().max()) < 1e-3
            continue
        d = float((g.double().cpu() - gref.double()).norm())
        num += d * d
        den += float(gref.double().norm()) ** 2
        r = d / (float(gref.double().norm()) + 1e-30)
        worst = max(worst, (r, k))
        assert r < ttol or float((g.cpu() - gref.cpu()).abs().max()) < 1e-6, (k, r)
    print("gradient error: global %.3e, worst tensor %.3e (%s)" % ((num / den) ** 0.5, worst[0], worst[1]))
    assert (num / den) ** 0.5 < gtol


@pytest.mark.parametrize("simt", [True, False])
def test_lstm_training_step_matches_reference_blocks_golden(env, golden_dir, simt):
    from blt_vqg_b200 import ops
    gold = torch.load(os.path.join(golden_dir, "iq_lstm_tiny.pt"), weights_only=False)
    b = gold["batch"]
    m = _build(gold["cfg"], gold["state_dict"])
    with ops.gemm_mode(simt=simt):
        terms = m.forward_losses(b["feats"], b["answers"], b["alengths"], b["questions"], eps=gold["eps"].cuda(), eps2=gold["eps2"].cuda())
    _check(m, terms, gold["losses"], gold["grads"], simt)
    assert float(terms[4]) == float((b["questions"][:, 1:] != 0).sum())
    assert rel(m.bn_running_mean, gold["bn_running_mean_after"]) < 1e-3


@pytest.mark.parametrize("size", ["small", "odd"])
@pytest.mark.parametrize("simt", [True, False])
def test_lstm_training_step_matches_oracle_small(env, simt, size):
    from blt_vqg_b200 import ops, synth
    from oracle import lstm_oracle
    cfg = synth.LSTM_CONFIGS[size]
    b = synth.make_batch(cfg)  # unsorted, ragged answer lengths: length masking instead of packing
    sd = lstm_oracle.init_state_dict(cfg, seed=4)
    g = torch.Generator().manual_seed(5)
    eps, eps2 = torch.randn(cfg["B"], cfg["Z"], generator=g), torch.randn(cfg["B"], cfg["Z"], generator=g)
    P = {k: v.clone().requires_grad_("running" not in k) for k, v in sd.items()}
    bn = (P["encoder_cnn.bn.running_mean"].clone(), P["encoder_cnn.bn.running_var"].clone())
    loss, want = lstm_oracle.lstm_iq_losses(P, b["feats"], b["answers"], b["alengths"], b["questions"], eps, eps2, bn_state=bn)
    loss.backward()
    m = _build(cfg, sd)
    with ops.gemm_mode(simt=simt):
        terms = m.forward_losses(b["feats"], b["answers"], b["alengths"], b["questions"], eps=eps.cuda(), eps2=eps2.cuda())
    _check(m, terms, {k: v.detach() for k, v in want.items()}, {k: v.grad for k, v in P.items() if v.requires_grad}, simt)


@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg4"])
def test_lstm_full_size_training_step_matches_fp64_oracle(env, name):
    """BASELINE.json configs[0] (cfg1: H=512, V=5000, B=32), configs[1] (cfg2: GloVe H=300, V=15000, B=256 — the persistent
    recurrence's production shape) and configs[3]'s per-GPU shard (cfg4: H=512, B=512) at their own sizes on the timed path,
    against the oracle executed in fp64 on this GPU.  (The glue's wiring is the builder's restatement of the call-site contract —
    the reference tree pins the blocks, not the wiring: SURVEY 8c.)"""
    from blt_vqg_b200 import synth
    from oracle import lstm_oracle
    cfg = synth.LSTM_CONFIGS[name]
    b = synth.make_batch(cfg, seed=21)
    sd = lstm_oracle.init_state_dict(cfg, seed=4)
    g = torch.Generator().manual_seed(5)
    eps, eps2 = torch.randn(cfg["B"], cfg["Z"], generator=g), torch.randn(cfg["B"], cfg["Z"], generator=g)
    P = {k: v.double().cuda().requires_grad_("running" not in k) for k, v in sd.items()}
    bn = (P["encoder_cnn.bn.running_mean"].detach().clone(), P["encoder_cnn.bn.running_var"].detach().clone())
    bg = {k: v.cuda() for k, v in b.items()}
    loss, want = lstm_oracle.lstm_iq_losses(P, bg["feats"].double(), bg["answers"], bg["alengths"], bg["questions"],
                                            eps.double().cuda(), eps2.double().cuda(), bn_state=bn)
    loss.backward()
    grads = {k: v.grad.detach().cpu() for k, v in P.items() if v.requires_grad}
    want = {k: v.detach().cpu() for k, v in want.items()}
    del loss
    torch.cuda.empty_cache()
    m = _build(cfg, sd)
    terms = m.forward_losses(b["feats"], b["answers"], b["alengths"], b["questions"], eps=eps.cuda(), eps2=eps2.cuda())
    _check(m, terms, want, grads, False, tols=FULL_TOL)


def test_staged_api_and_bit_exact_greedy_ids(env, golden_dir):
    """old_train_iq.py:100-106 call sequence in eval mode; ids of the free-running decoder equal the fp32 reference's"""
    from blt_vqg_b200 import ops, synth
    from oracle import lstm_oracle
    gold = torch.load(os.path.join(golden_dir, "iq_lstm_tiny.pt"), weights_only=False)
    b, cfg = gold["batch"], gold["cfg"]
    sd = dict(gold["state_dict"])
    sd["encoder_cnn.bn.running_mean"], sd["encoder_cnn.bn.running_var"] = gold["bn_running_mean_after"], gold["bn_running_var_after"]
    m = _build(cfg, sd).eval()
    with ops.gemm_mode(simt=True):
        img = m.encode_images(b["feats"])
        ans = m.encode_answers(b["answers"], b["alengths"])
        mus, logvars = m.encode_into_z(img, ans)
        zs = m.reparameterize(mus, logvars)           # eval: zs = mus
        outputs, hidden, ret = m.decode_questions(img, zs, questions=None, teacher_forcing_ratio=0)
        tf_out, _, tf_ret = m.decode_questions(img, zs, questions=b["questions"], teacher_forcing_ratio=1.0)
        ri, ra = m.reconstruct_inputs(img, ans)
    P = {k: v for k, v in sd.items()}
    with torch.no_grad():
        img_r = lstm_oracle.encode_images(P, b["feats"], training=False)
        ans_r = lstm_oracle.encode_answers(P, b["answers"], b["alengths"])
        mu_r, lv_r = lstm_oracle.encode_into_z(P, img_r, ans_r)
        logp_r = lstm_oracle.decode_questions(P, img_r, mu_r, b["questions"])
    assert rel(img, img_r) < 1e-5 and rel(ans, ans_r) < 1e-5 and rel(mus, mu_r) < 1e-5 and rel(logvars, lv_r) < 1e-5
    ids = torch.cat(ret[m.KEY_SEQUENCE], dim=1).cpu()
    assert ids.shape == (cfg["B"], cfg["T"]) and torch.equal(ids, gold["greedy_ids"])
    assert len(outputs) == cfg["T"] and outputs[0].shape == (cfg["B"], cfg["V"])
    assert len(tf_out) == cfg["T"] - 1
    assert rel(torch.stack(tf_out, 1), logp_r) < 1e-5
    assert hidden[0].shape == (1, cfg["B"], cfg["H"]) and len(ret[m.KEY_LENGTH]) == cfg["B"]
    assert ri.shape == ra.shape == (cfg["B"], cfg["H"])


def test_lstm_train_step_runs_with_dropout_and_adam(env):
    from blt_vqg_b200 import synth
    from blt_vqg_b200.lstm_iq import LSTMIQ
    from blt_vqg_b200.train import TrainStep
    cfg = synth.LSTM_CONFIGS["small"]
    torch.manual_seed(0)
    m = LSTMIQ.from_config(cfg, input_dropout_p=0.3).cuda().train()
    ts = TrainStep(m, lr_fn=lambda s: 1e-3)
    first = None
    for i in range(8):
        b = synth.make_batch(cfg, seed=7)
        t = ts.step(b["feats"], b["answers"], b["alengths"], b["questions"]).cpu()
        assert torch.isfinite(t).all()
        first = float(t[0]) if first is None else first
    assert float(t[0]) < first  # the generation loss goes down on a repeated batch


@pytest.mark.parametrize("T,B,H,ragged", [(19, 256, 300, False), (5, 256, 300, True), (7, 37, 64, True), (3, 16, 100, False), (4, 50, 304, True)])
def test_persistent_recurrence_matches_the_per_step_kernels(env, T, B, H, ragged):
    """b200iq_lstm_seq_fwd / _bwd (8-CTA clusters, W_hh slice in shared memory, one launch per pass) against the per-timestep
    GEMM + cell kernels on the same buffers: activated gates, h and c sequences, and the BPTT outputs."""
    from blt_vqg_b200 import _lib, ops
    assert ops.lstm_seq_supported(H)
    prev = _lib.lib().b200iq_set_tf32_rounding(1)
    try:
        torch.manual_seed(3)
        rt = lambda x: ops.round_tf32(x, torch.empty_like(x))
        whh = rt(torch.randn(4 * H, H, device="cuda") * H ** -0.5)
        xproj = torch.randn(T, B, 4 * H, device="cuda")
        h0, c0 = rt(torch.randn(B, H, device="cuda") * 0.5), torch.randn(B, H, device="cuda") * 0.5
        lengths = torch.randint(1, T + 1, (B,), device="cuda", dtype=torch.int32) if ragged else None

        def run(persistent):
            gates = xproj.clone()
            hbuf, cbuf = torch.zeros(T + 1, B, H, device="cuda"), torch.zeros(T + 1, B, H, device="cuda")
            hbuf[0].copy_(h0)
            cbuf[0].copy_(c0)
            if persistent:
                ops.lstm_seq_fwd(gates, whh, hbuf, cbuf, lengths)
            else:
                for t in range(T):
                    ops.gemm(hbuf[t], whh, gates[t], residual=gates[t], fp32_out=True)
                    ops.lstm_cell_fwd(gates[t], cbuf[t], hbuf[t], cbuf[t + 1], hbuf[t + 1], lengths, t)
            return gates, hbuf, cbuf
        g1, h1, c1 = run(True)
        g0, h0_, c0_ = run(False)
        assert torch.isfinite(h1).all()
        assert rel(h1, h0_) < 2e-3 and rel(c1, c0_) < 2e-3
        if lengths is None:
            assert rel(g1, g0) < 2e-3
        else:  # rows past their length keep the untouched input projection in both versions
            live = (torch.arange(T, device="cuda")[:, None] < lengths[None, :]).float()[:, :, None]
            assert rel(g1 * live, g0 * live) < 2e-3
        # BPTT: same activated gates / c sequence into both versions
        dh_last, d_out = rt(torch.randn(B, H, device="cuda")), rt(torch.randn(T, B, H, device="cuda"))
        for use_last, use_out in ((True, True), (True, False), (False, True)):
            dgp, dh0p, dc0p = torch.empty(T, B, 4 * H, device="cuda"), torch.empty(B, H, device="cuda"), torch.empty(B, H, device="cuda")
            ops.lstm_seq_bwd(g0, c0_, whh, dh_last if use_last else None, d_out if use_out else None, dgp, dh0p, dc0p, lengths)
            dgs, dh_run, dc = torch.empty(T, B, 4 * H, device="cuda"), torch.zeros(B, H, device="cuda"), torch.zeros(B, H, device="cuda")
            if use_last:
                dh_run.copy_(dh_last)
            for t in reversed(range(T)):
                ops.lstm_cell_bwd(g0[t], c0_[t], c0_[t + 1], dh_run, d_out[t] if use_out else None, dc, dgs[t], lengths, t)
                ops.gemm(dgs[t], whh, dh_run, b_t=True, residual=dh_run)
            assert torch.isfinite(dgp).all()
            assert rel(dgp, dgs) < 3e-3 and rel(dh0p, dh_run) < 3e-3 and rel(dc0p, dc) < 3e-3
    finally:
        _lib.lib().b200iq_set_tf32_rounding(prev)


@pytest.mark.parametrize("T,B,H,ragged", [(19, 256, 300, False), (5, 256, 300, True), (7, 37, 64, True), (3, 16, 100, False), (4, 50, 304, True),
                                           (6, 20, 30, True)])
def test_precise_persistent_forward_matches_an_fp64_recurrence(env, T, B, H, ragged):
    """b200iq_lstm_seq_fwd in the library's precise mode (TF32 rounding off: unrounded fp32 tensors, 3-term products — the fp16
    mma.sync kernel with the W_hh slice split once into fp16 hi / lo half-rows) against the same recurrence in fp64: activated
    gates, h and c sequences.  ~2^-21 per operand of the recurrent product, compounded over T steps."""
    from blt_vqg_b200 import ops
    assert ops.lstm_seq_supported(H)
    torch.manual_seed(4)
    whh = torch.randn(4 * H, H, device="cuda") * H ** -0.5
    xproj = torch.randn(T, B, 4 * H, device="cuda")
    h0, c0 = torch.randn(B, H, device="cuda") * 0.5, torch.randn(B, H, device="cuda") * 0.5
    lengths = torch.randint(1, T + 1, (B,), device="cuda", dtype=torch.int32) if ragged else None
    gates = xproj.clone()
    hbuf, cbuf = torch.zeros(T + 1, B, H, device="cuda"), torch.zeros(T + 1, B, H, device="cuda")
    hbuf[0].copy_(h0)
    cbuf[0].copy_(c0)
    with ops.no_rounding():
        ops.lstm_seq_fwd(gates, whh, hbuf, cbuf, lengths)
    torch.cuda.synchronize()
    h, c, w = h0.double(), c0.double(), whh.double()
    for t in range(T):
        pre = xproj[t].double() + h @ w.t()
        i, f, g, o = pre[:, :H].sigmoid(), pre[:, H:2 * H].sigmoid(), pre[:, 2 * H:3 * H].tanh(), pre[:, 3 * H:].sigmoid()
        cn = f * c + i * g
        hn = o * cn.tanh()
        live = torch.ones(B, 1, device="cuda", dtype=torch.bool) if lengths is None else (lengths > t)[:, None]
        h, c = torch.where(live, hn, h), torch.where(live, cn, c)
        act = torch.cat([i, f, g, o], 1)
        assert rel(hbuf[t + 1], h) < 2e-5 and rel(cbuf[t + 1], c) < 2e-5, t
        assert rel(gates[t] * live, act * live) < 2e-5, t


def test_category_space_matches_oracle_and_bit_exact_ids(env, golden_dir):
    """SURVEY §8f rank 4 — the evaluate() / sampler call sequence of old_train_iq.py:134-142,194-213 and evaluate_iq.py:51-54:
    encode_categories, encode_into_t, the two KL terms, predict_from_category / predict_from_answer (greedy ids bit-exact vs
    the fp32 oracle in exact mode); the answer space's parameters and offsets are untouched by the extra blocks."""
    from blt_vqg_b200 import _lib, ops
    from blt_vqg_b200.lstm_iq import LSTMIQ
    from oracle import lstm_oracle
    gold = torch.load(os.path.join(golden_dir, "iq_lstm_tiny.pt"), weights_only=False)
    b, cfg = gold["batch"], gold["cfg"]
    sd = dict(gold["state_dict"])
    sd["encoder_cnn.bn.running_mean"], sd["encoder_cnn.bn.running_var"] = gold["bn_running_mean_after"], gold["bn_running_var_after"]
    lstm_oracle.add_category_space(sd, cfg, num_categories=16, seed=3)
    m = LSTMIQ.from_config(cfg, no_category_space=False, num_categories=16).cuda()
    missing, unexpected = m.load_state_dict(sd)
    assert not missing and not unexpected
    m.eval()
    plain = _build(cfg, {k: v for k, v in sd.items() if "category" not in k and not k.startswith("t_decoder")})
    for k, off in plain.arena.offsets.items():
        assert m.arena.offsets[k] == off, k
    with pytest.raises(_lib.B200iqError):
        plain.encode_categories(torch.zeros(2, dtype=torch.long))
    B = cfg["B"]
    cats = torch.randint(0, 16, (B,), generator=torch.Generator().manual_seed(5))
    P = sd
    with torch.no_grad():
        img_r = lstm_oracle.encode_images(P, b["feats"], training=False)
        ans_r = lstm_oracle.encode_answers(P, b["answers"], b["alengths"])
        mu_r, lv_r = lstm_oracle.encode_into_z(P, img_r, ans_r)
        catf_r = lstm_oracle.encode_categories(P, cats)
        tmu_r, tlv_r = lstm_oracle.encode_into_t(P, img_r, catf_r)
        tkl_r = lstm_oracle.gaussian_kl_loss(tmu_r, tlv_r)
        ztkl_r = lstm_oracle.compute_two_gaussian_loss(mu_r, lv_r, tmu_r, tlv_r)
        ids_t_r = lstm_oracle.greedy_decode_from_t(P, img_r, tmu_r, 1, cfg["T"])
        ids_z_r = lstm_oracle.greedy_decode(P, img_r, mu_r, 1, cfg["T"])
    with ops.gemm_mode(simt=True):
        img = m.encode_images(b["feats"])
        ans = m.encode_answers(b["answers"], b["alengths"])
        mus, logvars = m.encode_into_z(img, ans)
        catf = m.encode_categories(cats)
        t_mus, t_logvars = m.encode_into_t(img, catf)
        t_kl = m.gaussian_KL_loss(t_mus, t_logvars)
        z_t_kl = m.compute_two_gaussian_loss(mus, logvars, t_mus, t_logvars)
        ids_t = m.predict_from_category(b["feats"], cats)
        ids_z = m.predict_from_answer(b["feats"], b["answers"], b["alengths"])
    assert torch.equal(catf.cpu(), catf_r)
    assert rel(t_mus, tmu_r) < 1e-5 and rel(t_logvars, tlv_r) < 1e-5
    assert float(t_kl) == pytest.approx(float(tkl_r), rel=1e-4)
    assert float(z_t_kl) == pytest.approx(float(ztkl_r), rel=1e-4)
    assert ids_t.shape == (B, cfg["T"]) and torch.equal(ids_t.cpu(), ids_t_r)
    assert torch.equal(ids_z.cpu(), ids_z_r) and torch.equal(ids_z.cpu(), gold["greedy_ids"])
    # tensor-core path: same quantities within the stated TF32 tolerance (losses 1e-3)
    m.train()   # reparameterize samples in train mode: injected eps
    eps = torch.randn(B, cfg["Z"], generator=torch.Generator().manual_seed(9))
    t_mus2, t_logvars2 = m.encode_into_t(img, catf)
    assert rel(t_mus2, tmu_r) < 5e-3 and rel(t_logvars2, tlv_r) < 5e-3
    assert float(m.compute_two_gaussian_loss(mus, logvars, t_mus2, t_logvars2)) == pytest.approx(float(ztkl_r), rel=5e-3)
    ts = m.reparameterize(t_mus, t_logvars, eps.cuda())
    assert rel(ts, lstm_oracle.reparameterize(tmu_r, tlv_r, eps)) < 1e-3
