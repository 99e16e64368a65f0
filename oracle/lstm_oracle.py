"""ORACLE (test infrastructure only): fp32 PyTorch restatement of the LSTM variant of IQ (upstream IQ, CVPR'19 lineage).

The reference tree holds the BLOCKS (`models/decoder_rnn.py:12` DecoderRNN, `models/encoder_rnn.py:6` EncoderRNN,
`models/mlp.py:10` MLP, `models/encoder_cnn.py` head) but not the glue that wires them; the glue contract is pinned only
by the call sites `old_train_iq.py:46-59,100-106,130,145-155` and `evaluate_iq.py:52-54,122-136`.
This file restates blocks + glue functionally over a state_dict whose names follow the upstream module tree:
  encoder_cnn.cnn.fc / encoder_cnn.bn, answer_encoder.{embedding,rnn}, answer_attention (MLP 2H->A->H),
  mu_answer_encoder / logvar_answer_encoder (H->Z), z_decoder (Z->H), gen_decoder (MLP H->A->H),
  decoder.{embedding,rnn,out}, image_reconstructor / answer_reconstructor (MLP Z->A->H).
Pin: oracle/gen_golden_lstm.py builds the same model from the REAL reference classes, checks this restatement against it
and writes tests/golden/iq_lstm_tiny.pt.  The glue itself is ours (parity of the glue is against this restatement).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
"""
import math

import torch
import torch.nn.functional as F

PAD = 0
LAMBDAS = dict(lambda_gen=1.0, lambda_z=0.001, lambda_i=0.001, lambda_a=0.001)


def lstm_seq(x, h0, c0, w_ih, w_hh, b_ih, b_hh, lengths=None):
    """nn.LSTM(batch_first=True, 1 layer) as used at decoder_rnn.py:105 / encoder_rnn.py:83; gate order i,f,g,o.
    With `lengths` (packed-sequence semantics, encoder_rnn.py:80-82) state stops updating after a row's last token.
    Returns (outputs (B,T,H), h_T, c_T)."""
    B, T, _ = x.shape
    h, c = h0, c0
    outs = []
    xp = F.linear(x, w_ih, b_ih + b_hh)
    for t in range(T):
        g = xp[:, t] + F.linear(h, w_hh)
        i, f, gg, o = g.chunk(4, dim=1)
        c_new = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
        h_new = torch.sigmoid(o) * torch.tanh(c_new)
        if lengths is not None:
            m = (lengths > t).to(x.dtype).unsqueeze(1)
            c_new = m * c_new + (1 - m) * c
            h_new = m * h_new + (1 - m) * h
        h, c = h_new, c_new
        outs.append(h)
    return torch.stack(outs, 1), h, c


def mlp2(P, pre, x):
    """models/mlp.py:49-56 with num_layers=2: fc0 -> ReLU -> (dropout 0) -> fc1"""
    return F.linear(torch.relu(F.linear(x, P[pre + "layers.fc0.weight"], P[pre + "layers.fc0.bias"])),
                    P[pre + "layers.fc1.weight"], P[pre + "layers.fc1.bias"])


def encode_images(P, feats, training=True, bn_state=None):
    """models/encoder_cnn.py:33-34 (trunk cut)"""
    y = F.linear(feats, P["encoder_cnn.cnn.fc.weight"], P["encoder_cnn.cnn.fc.bias"])
    rm, rv = bn_state if bn_state is not None else (P["encoder_cnn.bn.running_mean"], P["encoder_cnn.bn.running_var"])
    return F.batch_norm(y, rm, rv, P["encoder_cnn.bn.weight"], P["encoder_cnn.bn.bias"], training, 0.01, 1e-5)


def encode_answers(P, answers, alengths, drop_mask=None):
    """EncoderRNN(variable_lengths=True) final hidden of the top layer (encoder_rnn.py:78-86)"""
    e = F.embedding(answers, P["answer_encoder.embedding.weight"])
    if drop_mask is not None:
        e = e * drop_mask
    B, H = e.shape[0], e.shape[2]
    z = torch.zeros(B, H, dtype=e.dtype, device=e.device)
    _, h, _ = lstm_seq(e, z, z, P["answer_encoder.rnn.weight_ih_l0"], P["answer_encoder.rnn.weight_hh_l0"],
                       P["answer_encoder.rnn.bias_ih_l0"], P["answer_encoder.rnn.bias_hh_l0"], lengths=alengths)
    return h


def encode_into_z(P, img, ans):
    att = mlp2(P, "answer_attention.", torch.cat((img, ans), dim=1))
    return (F.linear(att, P["mu_answer_encoder.weight"], P["mu_answer_encoder.bias"]),
            F.linear(att, P["logvar_answer_encoder.weight"], P["logvar_answer_encoder.bias"]))


def reparameterize(mus, logvars, eps):
    return eps * torch.exp(0.5 * logvars) + mus


def decode_questions(P, img, zs, questions, drop_mask=None):
    """teacher-forced DecoderRNN (decoder_rnn.py:141-144,99-108): returns log-probs (B, T-1, V)"""
    zh = F.linear(zs, P["z_decoder.weight"], P["z_decoder.bias"])
    hid = mlp2(P, "gen_decoder.", img + zh)
    e = F.embedding(questions[:, :-1], P["decoder.embedding.weight"])
    if drop_mask is not None:
        e = e * drop_mask
    out, _, _ = lstm_seq(e, hid, hid, P["decoder.rnn.weight_ih_l0"], P["decoder.rnn.weight_hh_l0"],
                         P["decoder.rnn.bias_ih_l0"], P["decoder.rnn.bias_hh_l0"])
    logits = F.linear(out, P["decoder.out.weight"], P["decoder.out.bias"])
    return F.log_softmax(logits, dim=-1)


def gaussian_kl_loss(mus, logvars, eps=1e-8):
    """utils/train_utils.py:148-159"""
    return -0.5 * torch.sum(1 + logvars - mus.pow(2) - logvars.exp()) / (mus.size(0) + eps)


# ---- category space (old_train_iq.py:134-142,194-213; evaluate_iq.py:51-54).  Composition from the call sites: categories ->
# nn.Embedding(num_categories, H); encode_into_t mirrors encode_into_z with its own MLP + two Linear(H, Z); a sample of the
# category space reaches the decoder through its own Linear(Z, H) (t_decoder).
def encode_categories(P, categories):
    return F.embedding(categories, P["category_encoder.weight"])


def encode_into_t(P, img, catf):
    att = mlp2(P, "category_attention.", torch.cat((img, catf), dim=1))
    return (F.linear(att, P["mu_category_encoder.weight"], P["mu_category_encoder.bias"]),
            F.linear(att, P["logvar_category_encoder.weight"], P["logvar_category_encoder.bias"]))


def compute_two_gaussian_loss(mu1, logvar1, mu2, logvar2):
    """old_train_iq.py:262-278 (statement by statement)"""
    numerator = logvar1.exp() + torch.pow(mu1 - mu2, 2)
    fraction = torch.div(numerator, (logvar2.exp() + 1e-8))
    kl = 0.5 * torch.sum(logvar2 - logvar1 + fraction - 1)
    return kl / (mu1.size(0) + 1e-8)


def greedy_decode_from_t(P, img, ts, sos_id, max_len):
    Q = dict(P)
    Q["z_decoder.weight"], Q["z_decoder.bias"] = P["t_decoder.weight"], P["t_decoder.bias"]
    return greedy_decode(Q, img, ts, sos_id, max_len)


def add_category_space(P, cfg, num_categories=16, seed=1):
    """reference-style initialisation of the category-space blocks, appended to an init_state_dict() result"""
    g = torch.Generator().manual_seed(seed)
    H, Z, A = cfg["H"], cfg["Z"], cfg["A"]

    def uni(shape, b):
        return (torch.rand(shape, generator=g) * 2 - 1) * b

    P["category_encoder.weight"] = torch.randn(num_categories, H, generator=g)
    P["category_attention.layers.fc0.weight"] = torch.randn(A, 2 * H, generator=g) * math.sqrt(2.0 / (2 * H))
    P["category_attention.layers.fc0.bias"] = torch.zeros(A)
    P["category_attention.layers.fc1.weight"] = torch.randn(H, A, generator=g) * math.sqrt(2.0 / A)
    P["category_attention.layers.fc1.bias"] = torch.zeros(H)
    for nm in ("mu_category_encoder", "logvar_category_encoder"):
        P[nm + ".weight"], P[nm + ".bias"] = uni((Z, H), 1 / math.sqrt(H)), uni((Z,), 1 / math.sqrt(H))
    P["t_decoder.weight"], P["t_decoder.bias"] = uni((H, Z), 1 / math.sqrt(Z)), uni((H,), 1 / math.sqrt(Z))
    return P


def lstm_iq_losses(P, feats, answers, alengths, questions, eps, eps2, lambdas=LAMBDAS, training=True, bn_state=None):
    """one combined training loss: lambda_gen*CE + lambda_z*KL + lambda_i*MSE_img + lambda_a*MSE_ans
    (old_train_iq.py:100-155: CE over non-pad targets of questions[:,1:], KL to the unit normal, reconstructions from the
    DETACHED features through a second reparameterisation)."""
    img = encode_images(P, feats, training, bn_state)
    ans = encode_answers(P, answers, alengths)
    mus, logvars = encode_into_z(P, img, ans)
    zs = reparameterize(mus, logvars, eps)
    logp = decode_questions(P, img, zs, questions)
    V = logp.shape[-1]
    gen = F.nll_loss(logp.reshape(-1, V), questions[:, 1:].reshape(-1), ignore_index=PAD)
    kl = gaussian_kl_loss(mus, logvars)
    mus2, logvars2 = encode_into_z(P, img.detach(), ans.detach())
    z2 = reparameterize(mus2, logvars2, eps2)
    rec_i = F.mse_loss(mlp2(P, "image_reconstructor.", z2), img.detach())
    rec_a = F.mse_loss(mlp2(P, "answer_reconstructor.", z2), ans.detach())
    loss = lambdas["lambda_gen"] * gen + lambdas["lambda_z"] * kl + lambdas["lambda_i"] * rec_i + lambdas["lambda_a"] * rec_a
    return loss, {"gen": gen, "kl": kl, "rec_i": rec_i, "rec_a": rec_a, "loss": loss}


def greedy_decode(P, img, zs, sos_id, max_len):
    """free-running DecoderRNN (decoder_rnn.py:150-158): argmax ids (B, max_len)"""
    zh = F.linear(zs, P["z_decoder.weight"], P["z_decoder.bias"])
    h = c = mlp2(P, "gen_decoder.", img + zh)
    tok = torch.full((img.shape[0],), sos_id, dtype=torch.long, device=img.device)
    ids = []
    for _ in range(max_len):
        e = F.embedding(tok, P["decoder.embedding.weight"]).unsqueeze(1)
        out, h, c = lstm_seq(e, h, c, P["decoder.rnn.weight_ih_l0"], P["decoder.rnn.weight_hh_l0"],
                             P["decoder.rnn.bias_ih_l0"], P["decoder.rnn.bias_hh_l0"])
        tok = F.linear(out[:, 0], P["decoder.out.weight"], P["decoder.out.bias"]).argmax(dim=1)
        ids.append(tok)
    return torch.stack(ids, 1)


def init_state_dict(cfg, seed=0):
    """reference-style initialisation of every block (decoder_rnn.py:95-97, encoder_rnn.py:56, mlp.py:38-39,
    encoder_cnn.py:27-28, nn.LSTM / nn.Linear defaults)"""
    g = torch.Generator().manual_seed(seed)
    V, H, Z, A, FEAT = cfg["V"], cfg["H"], cfg["Z"], cfg["A"], cfg["FEAT"]
    P = {}

    def uni(shape, b):
        return (torch.rand(shape, generator=g) * 2 - 1) * b

    def linear(name, o, i):
        P[name + ".weight"], P[name + ".bias"] = uni((o, i), 1 / math.sqrt(i)), uni((o,), 1 / math.sqrt(i))

    def mlp(name, i, o):
        P[name + ".layers.fc0.weight"] = torch.randn(A, i, generator=g) * math.sqrt(2.0 / i)
        P[name + ".layers.fc0.bias"] = torch.zeros(A)
        P[name + ".layers.fc1.weight"] = torch.randn(o, A, generator=g) * math.sqrt(2.0 / A)
        P[name + ".layers.fc1.bias"] = torch.zeros(o)

    def lstm(name):
        k = 1 / math.sqrt(H)
        P[name + ".weight_ih_l0"], P[name + ".weight_hh_l0"] = uni((4 * H, H), k), uni((4 * H, H), k)
        P[name + ".bias_ih_l0"], P[name + ".bias_hh_l0"] = uni((4 * H,), k), uni((4 * H,), k)

    P["encoder_cnn.cnn.fc.weight"] = torch.randn(H, FEAT, generator=g) * 0.02
    P["encoder_cnn.cnn.fc.bias"] = torch.zeros(H)
    P["encoder_cnn.bn.weight"], P["encoder_cnn.bn.bias"] = torch.ones(H), torch.zeros(H)
    P["encoder_cnn.bn.running_mean"], P["encoder_cnn.bn.running_var"] = torch.zeros(H), torch.ones(H)
    P["answer_encoder.embedding.weight"] = uni((V, H), 0.1)
    lstm("answer_encoder.rnn")
    mlp("answer_attention", 2 * H, H)
    linear("mu_answer_encoder", Z, H)
    linear("logvar_answer_encoder", Z, H)
    linear("z_decoder", H, Z)
    mlp("gen_decoder", H, H)
    P["decoder.embedding.weight"] = uni((V, H), 0.1)
    lstm("decoder.rnn")
    P["decoder.out.weight"], P["decoder.out.bias"] = uni((V, H), 0.1), torch.zeros(V)
    mlp("image_reconstructor", Z, H)
    mlp("answer_reconstructor", Z, H)
    return P


def make_training_closure(cfg, seed=0):
    """(params, loss_fn(batch, step)) for bench.py's CPU arm: reference behaviour (CPU randn, input dropout 0.3)"""
    P = init_state_dict(cfg, seed)
    params = [v.requires_grad_(True) for k, v in P.items() if "running" not in k]
    bn = (P["encoder_cnn.bn.running_mean"], P["encoder_cnn.bn.running_var"])

    def loss_fn(b, step):
        Bn = b["feats"].shape[0]
        eps, eps2 = torch.randn(Bn, cfg["Z"]), torch.randn(Bn, cfg["Z"])
        return lstm_iq_losses(P, b["feats"], b["answers"], b["alengths"], b["questions"], eps, eps2, bn_state=bn)[0]

    return params, loss_fn
