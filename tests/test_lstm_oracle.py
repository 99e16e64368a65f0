"""CPU: the LSTM-variant oracle (oracle/lstm_oracle.py) against the golden fixture produced from the reference's own
DecoderRNN / EncoderRNN / MLP classes (oracle/gen_golden_lstm.py)."""
import os

import pytest
import torch

from blt_vqg_b200 import synth
from oracle import lstm_oracle, refload


def test_lstm_oracle_matches_reference_blocks(golden_dir):
    gold = torch.load(os.path.join(golden_dir, "iq_lstm_tiny.pt"), weights_only=False)
    b = gold["batch"]
    P = {k: v.clone().requires_grad_("running" not in k) for k, v in gold["state_dict"].items()}
    bn = (P["encoder_cnn.bn.running_mean"].clone(), P["encoder_cnn.bn.running_var"].clone())
    loss, terms = lstm_oracle.lstm_iq_losses(P, b["feats"], b["answers"], b["alengths"], b["questions"], gold["eps"], gold["eps2"], bn_state=bn)
    loss.backward()
    for k in ("loss", "gen", "kl", "rec_i", "rec_a"):
        assert float(terms[k].detach()) == pytest.approx(gold["losses"][k], rel=2e-5), k
    assert torch.allclose(bn[0], gold["bn_running_mean_after"], atol=1e-6)
    n = 0
    for k, gref in gold["grads"].items():
        g = P[k].grad
        assert float((g - gref).norm() / (gref.norm() + 1e-12)) < 1e-4 or float((g - gref).abs().max()) < 1e-7, k
        n += 1
    assert n >= 30


def test_lstm_oracle_greedy_ids_bit_exact(golden_dir):
    gold = torch.load(os.path.join(golden_dir, "iq_lstm_tiny.pt"), weights_only=False)
    b = gold["batch"]
    P = {k: v for k, v in gold["state_dict"].items()}
    P["encoder_cnn.bn.running_mean"], P["encoder_cnn.bn.running_var"] = gold["bn_running_mean_after"], gold["bn_running_var_after"]
    with torch.no_grad():
        img = lstm_oracle.encode_images(P, b["feats"], training=False)
        ans = lstm_oracle.encode_answers(P, b["answers"], b["alengths"])
        mu, _ = lstm_oracle.encode_into_z(P, img, ans)
        ids = lstm_oracle.greedy_decode(P, img, mu, synth.SOQ, gold["cfg"]["T"])
    assert torch.equal(ids, gold["greedy_ids"])


def test_packed_sequence_semantics_of_lstm_seq():
    """length masking == nn.LSTM on a packed sequence (encoder_rnn.py:80-86)"""
    torch.manual_seed(0)
    B, T, H = 5, 4, 8
    rnn = torch.nn.LSTM(H, H, 1, batch_first=True)
    x = torch.randn(B, T, H)
    lengths = torch.tensor([4, 3, 3, 2, 1])
    packed = torch.nn.utils.rnn.pack_padded_sequence(x, lengths, batch_first=True)
    _, (h, c) = rnn(packed)
    z = torch.zeros(B, H)
    _, h2, c2 = lstm_oracle.lstm_seq(x, z, z, rnn.weight_ih_l0, rnn.weight_hh_l0, rnn.bias_ih_l0, rnn.bias_hh_l0, lengths=lengths)
    assert torch.allclose(h[0], h2, atol=1e-6) and torch.allclose(c[0], c2, atol=1e-6)


@pytest.mark.skipif(not refload.available(), reason="reference tree only exists in the build container")
def test_lstm_golden_regenerates_from_reference(golden_dir):
    from oracle import gen_golden_lstm
    gold = torch.load(os.path.join(golden_dir, "iq_lstm_tiny.pt"), weights_only=False)
    fresh = gen_golden_lstm.run_reference("tiny")
    for k in ("loss", "gen", "kl", "rec_i", "rec_a"):
        assert fresh["losses"][k] == pytest.approx(gold["losses"][k], rel=1e-6)
    assert torch.equal(fresh["greedy_ids"], gold["greedy_ids"])


def test_category_space_oracle_kl_identities():
    """compute_two_gaussian_loss against torch.distributions; KL(q || N(0,1)) equals gaussian_kl_loss"""
    import torch.distributions as D
    from oracle import lstm_oracle
    g = torch.Generator().manual_seed(0)
    mu1, lv1, mu2, lv2 = (torch.randn(6, 10, generator=g) for _ in range(4))
    want = D.kl_divergence(D.Normal(mu1, (0.5 * lv1).exp()), D.Normal(mu2, (0.5 * lv2).exp())).sum() / 6
    assert float(lstm_oracle.compute_two_gaussian_loss(mu1, lv1, mu2, lv2)) == pytest.approx(float(want), rel=1e-5)
    z = torch.zeros_like(mu1)
    assert float(lstm_oracle.compute_two_gaussian_loss(mu1, lv1, z, z)) == pytest.approx(float(lstm_oracle.gaussian_kl_loss(mu1, lv1)), rel=1e-5)
    P = lstm_oracle.add_category_space({}, dict(H=32, Z=8, A=16), num_categories=16)
    t_mu, t_lv = lstm_oracle.encode_into_t(P, torch.randn(3, 32, generator=g), lstm_oracle.encode_categories(P, torch.tensor([0, 15, 7])))
    assert t_mu.shape == t_lv.shape == (3, 8)


def test_category_space_arena_is_appended_behind_the_answer_space(golden_dir):
    """host logic (no device): the category-space parameters carry the upstream key names, sit behind the answer space in the flat
    arena (answer-space offsets and checkpoints do not move), and round-trip through state_dict / load_state_dict"""
    from blt_vqg_b200.lstm_iq import LSTMIQ
    gold = torch.load(os.path.join(golden_dir, "iq_lstm_tiny.pt"), weights_only=False)
    cfg = gold["cfg"]
    plain = LSTMIQ.from_config(cfg)
    full = LSTMIQ.from_config(cfg, no_category_space=False, num_categories=16)
    for k, off in plain.arena.offsets.items():
        assert full.arena.offsets[k] == off, k
    extra = [k for k in full.state_dict() if k not in plain.state_dict()]
    assert sorted(extra) == sorted([
        "category_encoder.weight", "category_attention.layers.fc0.weight", "category_attention.layers.fc0.bias",
        "category_attention.layers.fc1.weight", "category_attention.layers.fc1.bias", "mu_category_encoder.weight",
        "mu_category_encoder.bias", "logvar_category_encoder.weight", "logvar_category_encoder.bias", "t_decoder.weight", "t_decoder.bias"])
    off_of = lambda k: full.arena.offsets.get(k + "#pack", full.arena.offsets.get(k))   # packed [mu | logvar] operands carry a #pack mark
    assert min(off_of(k) for k in extra) >= plain.flat.numel() - 64
    sd = lstm_oracle.add_category_space(dict(gold["state_dict"]), cfg, num_categories=16, seed=3)
    missing, unexpected = full.load_state_dict(sd)
    assert not missing and not unexpected
    out = full.state_dict()
    for k in extra:
        assert torch.equal(out[k], sd[k]), k
    # the answer-space golden weights load into the plain model unchanged
    missing, unexpected = plain.load_state_dict(gold["state_dict"])
    assert not missing and not unexpected
    from blt_vqg_b200 import _lib
    with pytest.raises(_lib.B200iqError):
        plain.encode_categories(torch.zeros(2, dtype=torch.long))
