"""ORACLE pin for the LSTM variant (test infrastructure only; needs /root/reference):

  python oracle/gen_golden_lstm.py      # writes tests/golden/iq_lstm_tiny.pt

Builds the LSTM-IQ glue out of the REAL reference classes — models.decoder_rnn.DecoderRNN, models.encoder_rnn.EncoderRNN
(variable_lengths=True, packed sequences), models.mlp.MLP and the encoder_cnn head — following the call-site contract of
old_train_iq.py:100-155, runs forward + losses + backward and greedy decoding on a seeded synthetic batch, checks the
functional restatement oracle/lstm_oracle.py against it and stores the vectors.
"""
import os
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from blt_vqg_b200 import synth  # noqa: E402
from oracle import lstm_oracle, refload  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


class RefLSTMIQ(nn.Module):
    """upstream-style glue over the reference's own blocks (SURVEY.md §8c 'LSTM glue contract')"""

    def __init__(self, cfg, sos_id, eos_id):
        super().__init__()
        refload.import_models()
        from models.decoder_rnn import DecoderRNN
        from models.encoder_rnn import EncoderRNN
        from models.mlp import MLP
        V, H, Z, A, T = cfg["V"], cfg["H"], cfg["Z"], cfg["A"], cfg["T"]

        class Head(nn.Module):  # models/encoder_cnn.py:20-21 with the trunk cut
            def __init__(self):
                super().__init__()
                self.cnn = nn.Module()
                self.cnn.fc = nn.Linear(cfg["FEAT"], H)
                self.bn = nn.BatchNorm1d(H, momentum=0.01)

            def forward(self, x):
                return self.bn(self.cnn.fc(x))

        self.encoder_cnn = Head()
        self.answer_encoder = EncoderRNN(V, cfg["S"], H, input_dropout_p=0, dropout_p=0, n_layers=1, rnn_cell="lstm", variable_lengths=True)
        self.answer_attention = MLP(2 * H, A, H, num_layers=2)
        self.mu_answer_encoder = nn.Linear(H, Z)
        self.logvar_answer_encoder = nn.Linear(H, Z)
        self.z_decoder = nn.Linear(Z, H)
        self.gen_decoder = MLP(H, A, H, num_layers=2)
        self.decoder = DecoderRNN(V, T, H, sos_id=sos_id, eos_id=eos_id, n_layers=1, rnn_cell="lstm", input_dropout_p=0, dropout_p=0)
        self.image_reconstructor = MLP(Z, A, H, num_layers=2)
        self.answer_reconstructor = MLP(Z, A, H, num_layers=2)

    def encode_answers(self, answers, alengths):
        _, (h, _) = self.answer_encoder(answers, alengths, None)
        return h[-1]

    def encode_into_z(self, img, ans):
        att = self.answer_attention(torch.cat((img, ans), dim=1))
        return self.mu_answer_encoder(att), self.logvar_answer_encoder(att)

    def decode_questions(self, img, zs, questions=None, teacher_forcing_ratio=0):
        hid = self.gen_decoder(img + self.z_decoder(zs)).unsqueeze(0).contiguous()
        return self.decoder(inputs=questions, encoder_hidden=(hid, hid), function=F.log_softmax, teacher_forcing_ratio=teacher_forcing_ratio)


def run_reference(cfg_name="tiny"):
    cfg = synth.LSTM_CONFIGS[cfg_name]
    batch = synth.make_batch(cfg)
    order = torch.argsort(batch["alengths"], descending=True, stable=True)  # pack_padded_sequence wants sorted lengths
    batch = {k: v[order].contiguous() for k, v in batch.items()}
    sd = lstm_oracle.init_state_dict(cfg, seed=0)
    m = RefLSTMIQ(cfg, synth.SOQ, synth.EOS)
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all("num_batches_tracked" in k for k in missing), (missing, unexpected)
    m.train()
    g = torch.Generator().manual_seed(99)
    eps, eps2 = torch.randn(cfg["B"], cfg["Z"], generator=g), torch.randn(cfg["B"], cfg["Z"], generator=g)
    lam = lstm_oracle.LAMBDAS
    img = m.encoder_cnn(batch["feats"])
    ans = m.encode_answers(batch["answers"], batch["alengths"].tolist())
    mus, logvars = m.encode_into_z(img, ans)
    zs = eps * torch.exp(0.5 * logvars) + mus
    outputs, _, _ = m.decode_questions(img, zs, questions=batch["questions"], teacher_forcing_ratio=1.0)
    logp = torch.cat([o.unsqueeze(1) for o in outputs], dim=1)          # old_train_iq.py:117-118
    gen = F.nll_loss(logp.reshape(-1, cfg["V"]), batch["questions"][:, 1:].reshape(-1), ignore_index=synth.PAD)
    kl = lstm_oracle.gaussian_kl_loss(mus, logvars)                     # utils/train_utils.py:148 (2 lines, torchtext-free)
    mus2, logvars2 = m.encode_into_z(img.detach(), ans.detach())        # old_train_iq.py:145-149
    z2 = eps2 * torch.exp(0.5 * logvars2) + mus2
    rec_i = F.mse_loss(m.image_reconstructor(z2), img.detach())
    rec_a = F.mse_loss(m.answer_reconstructor(z2), ans.detach())
    loss = lam["lambda_gen"] * gen + lam["lambda_z"] * kl + lam["lambda_i"] * rec_i + lam["lambda_a"] * rec_a
    loss.backward()
    grads = {k: p.grad.detach().clone() for k, p in m.named_parameters()}
    sd_after = {k: v.detach().clone() for k, v in m.state_dict().items()}
    m.eval()
    with torch.no_grad():
        img_e = m.encoder_cnn(batch["feats"])
        ans_e = m.encode_answers(batch["answers"], batch["alengths"].tolist())
        mu_e, _ = m.encode_into_z(img_e, ans_e)
        _, _, ret = m.decode_questions(img_e, mu_e, questions=None, teacher_forcing_ratio=0)   # eval: zs = mus
        greedy = torch.cat(ret["sequence"], dim=1)
    return {"cfg_name": cfg_name, "cfg": dict(cfg), "batch": batch, "eps": eps, "eps2": eps2, "state_dict": sd, "lambdas": dict(lam),
            "losses": {"loss": float(loss), "gen": float(gen), "kl": float(kl), "rec_i": float(rec_i), "rec_a": float(rec_a)},
            "grads": grads, "greedy_ids": greedy, "bn_running_mean_after": sd_after["encoder_cnn.bn.running_mean"],
            "bn_running_var_after": sd_after["encoder_cnn.bn.running_var"],
            "act": {"img": img.detach().clone(), "ans": ans.detach().clone(), "mus": mus.detach().clone(), "logp_sample": logp[:, :, :12].detach().clone()}}


def check_restatement(gold):
    b = gold["batch"]
    P = {k: v.clone().requires_grad_("running" not in k) for k, v in gold["state_dict"].items()}
    bn = (P["encoder_cnn.bn.running_mean"].clone(), P["encoder_cnn.bn.running_var"].clone())
    loss, terms = lstm_oracle.lstm_iq_losses(P, b["feats"], b["answers"], b["alengths"], b["questions"], gold["eps"], gold["eps2"], bn_state=bn)
    loss.backward()
    worst = 0.0
    for k in ("loss", "gen", "kl", "rec_i", "rec_a"):
        r = abs(float(terms[k]) - gold["losses"][k]) / max(abs(gold["losses"][k]), 1e-9)
        worst = max(worst, r)
        assert r < 2e-5, (k, float(terms[k]), gold["losses"][k])
    for k, gref in gold["grads"].items():
        r = float((P[k].grad - gref).norm() / (gref.norm() + 1e-12))
        worst = max(worst, r)
        assert r < 1e-4 or float((P[k].grad - gref).abs().max()) < 1e-7, (k, r)
    with torch.no_grad():
        Pe = {k: v.detach() for k, v in gold["state_dict"].items()}
        Pe["encoder_cnn.bn.running_mean"], Pe["encoder_cnn.bn.running_var"] = gold["bn_running_mean_after"], gold["bn_running_var_after"]
        img = lstm_oracle.encode_images(Pe, b["feats"], training=False)
        ans = lstm_oracle.encode_answers(Pe, b["answers"], b["alengths"])
        mu, _ = lstm_oracle.encode_into_z(Pe, img, ans)
        ids = lstm_oracle.greedy_decode(Pe, img, mu, synth.SOQ, gold["cfg"]["T"])
    assert torch.equal(ids, gold["greedy_ids"]), "greedy ids differ from the reference blocks"
    return worst


def main():
    if not refload.available():
        print("reference tree not found")
        return 1
    os.makedirs(GOLDEN, exist_ok=True)
    gold = run_reference("tiny")
    worst = check_restatement(gold)
    path = os.path.join(GOLDEN, "iq_lstm_tiny.pt")
    torch.save(gold, path)
    print("%s losses=%s restatement worst rel err %.2e (%d KB)" % (path, {k: round(v, 5) for k, v in gold["losses"].items()}, worst, os.path.getsize(path) // 1024))
    return 0


if __name__ == "__main__":
    sys.exit(main())
