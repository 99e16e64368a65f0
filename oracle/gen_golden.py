"""ORACLE pin (test infrastructure only): run the REAL reference modules and write golden fixtures.

  python oracle/gen_golden.py          # needs /root/reference; writes tests/golden/iq_transformer_<cfg>_<mode>.pt

For each (config, mode) it builds models.IQ from /root/reference (oracle/refload.py shims), runs
forward + TrainIQ.calculate_losses (verbatim, train_iq.py:81-103) + backward + decode_greedy on the seeded
synthetic batch of blt_vqg_b200/synth.py with dropout off and an injected `eps`, checks the functional
restatement oracle/iq_oracle.py against it (losses, every gradient, greedy ids) and stores inputs, weights,
eps, per-term losses, all gradients, greedy ids and a few activations.  The reference's own tests hold no
vectors for this path, so "reference executed here" IS the pin.
"""
import os
import sys
from types import SimpleNamespace

import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from blt_vqg_b200 import synth  # noqa: E402
from oracle import iq_oracle, refload  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def run_reference(cfg_name, latent_mode, kliter=9000, greedy_len=12):
    cfg = synth.CONFIGS[cfg_name]
    vocab = synth.SynthVocab(cfg["V"])
    args = synth.args_for(cfg)
    batch = synth.make_batch(cfg)
    m = refload.build_reference_iq(vocab, args, latent_mode, seed=0)
    m.train()
    sd = refload.head_state_dict(m)
    g = torch.Generator().manual_seed(4321)
    eps = torch.randn(cfg["B"], cfg["L"], generator=g)
    eps_greedy = torch.randn(cfg["B"], cfg["L"], generator=g)
    calc = refload.import_calculate_losses()
    self_ = SimpleNamespace(criterion=nn.CrossEntropyLoss(ignore_index=synth.PAD), image_recon_criterion=nn.MSELoss(),
                            latent_transformer=latent_mode, kliter=kliter, args=args)
    with refload.inject_eps(eps):
        output, z_logit, kld, (img, recon) = m(batch["feats"], batch["answers"], batch["posteriors"], batch["questions"])
    loss, rec, limg, ppl, kldv, aux, elbo = calc(self_, output, (img, recon), kld, z_logit, batch["questions"])
    loss.backward()
    grads = refload.head_grads(m)
    # BatchNorm running stats after ONE training forward
    sd_after = refload.head_state_dict(m)
    m.eval()
    with torch.no_grad(), refload.inject_eps(eps_greedy):
        _, top_args, top_vals = m.decode_greedy(batch["feats"], batch["answers"], max_decode_length=greedy_len)
    greedy_ids = top_args[:, :, 0].long()
    out = {
        "cfg_name": cfg_name, "cfg": dict(cfg), "latent_mode": latent_mode, "kliter": kliter, "greedy_len": greedy_len,
        "args": dict(vars(args)), "batch": batch, "eps": eps, "eps_greedy": eps_greedy, "state_dict": sd,
        "bn_running_mean_after": sd_after["encoder_cnn.bn.running_mean"], "bn_running_var_after": sd_after["encoder_cnn.bn.running_var"],
        "losses": {"loss": float(loss), "rec": rec, "img": limg, "kld": float(kldv), "aux": float(aux), "elbo": elbo, "ppl": ppl},
        "grads": grads, "greedy_ids": greedy_ids, "greedy_top6_ids": top_args.long(), "greedy_top6_vals": top_vals,
        "act": {"output_sample": output[:, :, :16].detach().clone(), "img": img.detach().clone(), "recon": recon.detach().clone(),
                "z_logit_sample": None if z_logit is None else z_logit[:, :16].detach().clone()},
    }
    return out


def check_restatement(gold, tol=2e-5):
    """oracle/iq_oracle.py against what the reference just produced"""
    cfg, lat = gold["cfg"], gold["latent_mode"]
    b = gold["batch"]
    P = {k: v.clone().requires_grad_(v.dtype.is_floating_point and "running" not in k) for k, v in gold["state_dict"].items()}
    ocfg = dict(num_layers=cfg["NL"], num_heads=cfg["NH"], latent_dim=cfg["L"], **{k: gold["args"][k] for k in
                ("image_recon_lambda", "kl_ceiling", "aux_ceiling", "full_kl_step")})
    bn = (P["encoder_cnn.bn.running_mean"].clone(), P["encoder_cnn.bn.running_var"].clone())
    out, zl, kld, img, recon = iq_oracle.iq_forward(P, ocfg, b["feats"], b["answers"], b["posteriors"], b["questions"],
                                                    eps=gold["eps"], latent_mode=lat, training=True, bn_state=bn)
    loss, terms = iq_oracle.iq_losses(ocfg, out, zl, kld, img, recon, b["questions"], gold["kliter"], lat)
    loss.backward()
    worst = 0.0
    for k in ("loss", "rec", "img", "kld", "aux"):
        r = abs(float(terms[k]) - gold["losses"][k]) / max(1e-6, abs(gold["losses"][k]))
        worst = max(worst, r)
        assert r < tol, (k, float(terms[k]), gold["losses"][k])
    for k, gref in gold["grads"].items():
        canon = k
        for alias, c in (("answer_encoder.embedding.", "embedding."), ("decoder.embedding.", "embedding."),
                         ("answer_encoder.latent_layer.", "latent_layer.")):
            if k.startswith(alias):
                canon = c + k[len(alias):]
        g = P[canon].grad
        if gref is None:
            assert g is None or float(g.abs().max()) == 0.0, k
            continue
        r = float((g - gref).norm() / (gref.norm() + 1e-12))
        worst = max(worst, r)
        assert r < 1e-4, (k, r)
    with torch.no_grad():
        Pe = {k: v.detach() for k, v in gold["state_dict"].items()}
        Pe["encoder_cnn.bn.running_mean"] = gold["bn_running_mean_after"]
        Pe["encoder_cnn.bn.running_var"] = gold["bn_running_var_after"]
        ids, t6, v6 = iq_oracle.greedy_decode(Pe, ocfg, b["feats"], b["answers"], eps=gold["eps_greedy"], latent_mode=lat,
                                              max_decode_length=gold["greedy_len"])
    assert torch.equal(ids, gold["greedy_ids"]), "greedy ids differ from the reference"
    return worst


def main():
    if not refload.available():
        print("reference tree not found at %s" % refload.REF_ROOT)
        return 1
    os.makedirs(GOLDEN, exist_ok=True)
    for cfg_name in ("tiny",):
        for lat in (True, False):
            gold = run_reference(cfg_name, lat)
            worst = check_restatement(gold)
            path = os.path.join(GOLDEN, "iq_transformer_%s_%s.pt" % (cfg_name, "latent" if lat else "pretrain"))
            torch.save(gold, path)
            print("%s  losses=%s  restatement worst rel err %.2e  (%d KB)" % (path, {k: round(v, 5) for k, v in gold["losses"].items()},
                                                                              worst, os.path.getsize(path) // 1024))
    return 0


if __name__ == "__main__":
    sys.exit(main())
