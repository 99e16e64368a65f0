"""CPU: reference checkpoint formats <-> the flat arenas (blt_vqg_b200/checkpoint.py)."""
import os

import torch

from blt_vqg_b200 import checkpoint, synth
from blt_vqg_b200.iq import IQ
from blt_vqg_b200.train import TrainStep

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "iq_transformer_tiny_latent.pt")


def _model(cfg, seed):
    torch.manual_seed(seed)
    return IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg))


def test_reference_state_dict_loads_from_a_lightning_checkpoint_and_a_legacy_file(tmp_path):
    """the golden fixture holds the state_dict of the REAL reference IQ (oracle/gen_golden.py): wrapped the way
    trainer.save_checkpoint writes it (`model.` prefix, trunk tensors beside the head) it fills the arena completely"""
    gold = torch.load(GOLD, weights_only=False)
    ref_sd = gold["state_dict"]
    assert "answer_encoder.embedding.0.weight" in ref_sd and "decoder.embedding.0.weight" in ref_sd   # shared modules, every path
    ckpt = {"epoch": 3, "global_step": 1200, "pytorch-lightning_version": "1.1.0",
            "state_dict": {"model." + k: v for k, v in ref_sd.items()}, "optimizer_states": [{"state": {}, "param_groups": []}]}
    ckpt["state_dict"]["model.encoder_cnn.cnn.conv1.weight"] = torch.zeros(64, 3, 7, 7)     # frozen trunk: not ours
    ckpt["state_dict"]["model.encoder_cnn.cnn.layer1.0.bn1.running_mean"] = torch.zeros(64)
    m = _model(gold["cfg"], seed=7)
    ts = TrainStep(m)
    missing, unexpected = checkpoint.load_checkpoint(ckpt, m, ts)
    assert missing == [] and unexpected == []
    assert ts.iter == 1200
    for k in ("embedding.0.weight", "encoder_cnn.cnn.fc.weight", "decoder.decoder.dec.1.multi_head_attention_dec.key_linear.weight",
              "latent_layer.mean_logvar_posterior.6.bias", "decoder.output.weight"):
        assert torch.equal(m.state_dict()[k], ref_sd[k]), k
    # legacy torch.save(model.state_dict()) (old_train_iq.py:461-469): no prefix, through a file
    path = str(tmp_path / "legacy.pt")
    torch.save(ref_sd, path)
    m2 = _model(gold["cfg"], seed=8)
    assert checkpoint.load_checkpoint(path, m2) == ([], [])
    assert torch.equal(m2.flat.data, m.flat.data)


def test_checkpoint_round_trip_with_trainer_state(tmp_path):
    cfg = torch.load(GOLD, weights_only=False)["cfg"]
    m = _model(cfg, seed=1)
    ts = TrainStep(m)
    ts.exp_avg.normal_()
    ts.exp_avg_sq.uniform_()
    ts.t, ts.iter, ts.kliter, m.step_idx = 17, 4017, 17, 4017
    m.switch_GVT_train_mode(True)
    path = str(tmp_path / "step.ckpt")
    checkpoint.save_checkpoint(path, m, ts, epoch=2, hyper_parameters={"hidden_dim": cfg["H"]})
    ck = torch.load(path, weights_only=False)
    assert ck["global_step"] == 4017 and ck["epoch"] == 2 and ck["hyper_parameters"]["hidden_dim"] == cfg["H"]
    sd = ck["state_dict"]
    # what TrainIQ.load_state_dict would look for: the prefix, and every path of the shared sub-modules
    for k in ("model.embedding.0.weight", "model.answer_encoder.embedding.0.weight", "model.decoder.embedding.0.weight",
              "model.latent_layer.mean_logvar_prior.0.weight", "model.answer_encoder.latent_layer.mean_logvar_prior.0.weight",
              "model.encoder_cnn.bn.running_mean"):
        assert k in sd, k
    assert sd["model.embedding.0.weight"].data_ptr() != m.state_dict()["embedding.0.weight"].data_ptr()   # a copy, not a view
    m2 = _model(cfg, seed=2)
    ts2 = TrainStep(m2)
    assert checkpoint.load_checkpoint(path, m2, ts2) == ([], [])
    assert torch.equal(m2.flat.data, m.flat.data)
    assert torch.equal(ts2.exp_avg, ts.exp_avg) and torch.equal(ts2.exp_avg_sq, ts.exp_avg_sq)
    assert (ts2.t, ts2.iter, ts2.kliter, m2.step_idx) == (17, 4017, 17, 4017)
    assert m2.latent_transformer is True
