"""CPU: the oracle restatement (oracle/iq_oracle.py) against the committed golden fixtures, which were produced
by running the REAL reference modules (oracle/gen_golden.py).  Bit-for-meaning pin of the checker itself."""
import os

import pytest
import torch

from oracle import iq_oracle, refload

ALIASES = (("answer_encoder.embedding.", "embedding."), ("decoder.embedding.", "embedding."),
           ("answer_encoder.latent_layer.", "latent_layer."))


def _canon(k):
    for a, c in ALIASES:
        if k.startswith(a):
            return c + k[len(a):]
    return k


def _ocfg(gold):
    cfg = gold["cfg"]
    return dict(num_layers=cfg["NL"], num_heads=cfg["NH"], latent_dim=cfg["L"],
                **{k: gold["args"][k] for k in ("image_recon_lambda", "kl_ceiling", "aux_ceiling", "full_kl_step")})


@pytest.mark.parametrize("mode", ["latent", "pretrain"])
def test_oracle_matches_reference_golden(golden_dir, mode):
    gold = torch.load(os.path.join(golden_dir, "iq_transformer_tiny_%s.pt" % mode), weights_only=False)
    lat = gold["latent_mode"]
    b = gold["batch"]
    P = {k: v.clone().requires_grad_(v.dtype.is_floating_point and "running" not in k) for k, v in gold["state_dict"].items()}
    bn = (P["encoder_cnn.bn.running_mean"].clone(), P["encoder_cnn.bn.running_var"].clone())
    out, zl, kld, img, recon = iq_oracle.iq_forward(P, _ocfg(gold), b["feats"], b["answers"], b["posteriors"], b["questions"],
                                                    eps=gold["eps"], latent_mode=lat, training=True, bn_state=bn)
    loss, terms = iq_oracle.iq_losses(_ocfg(gold), out, zl, kld, img, recon, b["questions"], gold["kliter"], lat)
    loss.backward()
    for k in ("loss", "rec", "img", "kld", "aux"):
        assert float(terms[k]) == pytest.approx(gold["losses"][k], rel=2e-5, abs=1e-7), k
    assert torch.allclose(bn[0], gold["bn_running_mean_after"], atol=1e-6)
    assert torch.allclose(bn[1], gold["bn_running_var_after"], atol=1e-6)
    assert torch.allclose(out[:, :, :16], gold["act"]["output_sample"], atol=2e-5)
    n_checked = 0
    for k, gref in gold["grads"].items():
        g = P[_canon(k)].grad
        if gref is None:  # N7: no gradient in pre-training mode
            assert g is None or float(g.abs().max()) == 0.0, k
            continue
        assert float((g - gref).norm() / (gref.norm() + 1e-12)) < 1e-4, k
        n_checked += 1
    assert n_checked > 50


@pytest.mark.parametrize("mode", ["latent", "pretrain"])
def test_oracle_greedy_ids_bit_exact(golden_dir, mode):
    gold = torch.load(os.path.join(golden_dir, "iq_transformer_tiny_%s.pt" % mode), weights_only=False)
    b = gold["batch"]
    P = {k: v for k, v in gold["state_dict"].items()}
    P["encoder_cnn.bn.running_mean"] = gold["bn_running_mean_after"]
    P["encoder_cnn.bn.running_var"] = gold["bn_running_var_after"]
    with torch.no_grad():
        ids, t6, v6 = iq_oracle.greedy_decode(P, _ocfg(gold), b["feats"], b["answers"], eps=gold["eps_greedy"],
                                              latent_mode=gold["latent_mode"], max_decode_length=gold["greedy_len"])
    assert torch.equal(ids, gold["greedy_ids"])
    assert torch.equal(t6, gold["greedy_top6_ids"])
    assert torch.allclose(v6, gold["greedy_top6_vals"], atol=1e-6)


def test_noam_lr():
    """train_iq.py:252-257"""
    assert iq_oracle.noam_lr(0, 512) == 0.0
    assert iq_oracle.noam_lr(4000, 512) == pytest.approx(512 ** -0.5 * 4000 ** -0.5, rel=1e-3)


@pytest.mark.skipif(not refload.available(), reason="reference tree only exists in the build container")
def test_golden_regenerates_from_reference(golden_dir):
    """the committed fixture is what the reference produces today (guards against a stale fixture)"""
    from oracle import gen_golden
    gold = torch.load(os.path.join(golden_dir, "iq_transformer_tiny_latent.pt"), weights_only=False)
    fresh = gen_golden.run_reference("tiny", True)
    for k in ("loss", "rec", "img", "kld", "aux"):
        assert fresh["losses"][k] == pytest.approx(gold["losses"][k], rel=1e-6)
    assert torch.equal(fresh["greedy_ids"], gold["greedy_ids"])
