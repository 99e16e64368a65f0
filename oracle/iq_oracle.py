"""ORACLE (test infrastructure only — never imported by the product path).

Plain-PyTorch fp32 CPU restatement of the reference's transformer IQ training step
(nihirv/blt-vqg).  It is written functionally over a `state_dict` that uses the
reference's own parameter names, so weights copy both ways.  Each function cites the
reference file:line it follows.  Parity pin: `oracle/gen_golden.py` runs the REAL
reference modules (imported from /root/reference) on seeded inputs, checks this
restatement against them and writes `tests/golden/*.pt`; `tests/test_oracle.py`
re-checks the restatement against those fixtures.  The reference's own tests hold no
golden vectors for this path, so the pin is "reference executed here".

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.
"""
import math

import torch
import torch.nn.functional as F

PAD, SOQ = 0, 1  # utils/train_utils.py:18-37 reserved ids (<pad>=0, <start>=1)


def timing_signal(length, channels, min_timescale=1.0, max_timescale=1.0e4):
    """models/transformer_layers.py:542-558 — [sin | cos] concatenated (not interleaved)."""
    pos = torch.arange(length, dtype=torch.float64)
    nts = channels // 2
    inc = math.log(float(max_timescale) / float(min_timescale)) / (float(nts) - 1)
    inv = min_timescale * torch.exp(torch.arange(nts, dtype=torch.float64) * -inc)
    st = pos[:, None] * inv[None, :]
    sig = torch.cat([torch.sin(st), torch.cos(st)], dim=1)
    if channels % 2:
        sig = F.pad(sig, (0, 1))
    return sig.float()


def embed(P, ids):
    """models/iq.py:57-79 — Embedding(V,E,padding_idx=PAD) -> Linear(E,H).  The pad row is NOT zero
    (overwritten by copy_ at iq.py:73); padding_idx only blocks its gradient."""
    e = F.embedding(ids, P["embedding.0.weight"], padding_idx=PAD)
    return F.linear(e, P["embedding.1.weight"], P["embedding.1.bias"])


def mha(P, pre, q_in, kv_in, mask, nh, drop_p=0.0, training=False):
    """models/transformer_layers.py:486-532.  mask: bool, True = masked, [B,1,nk] or [B,nq,nk]."""
    q = F.linear(q_in, P[pre + "query_linear.weight"])
    k = F.linear(kv_in, P[pre + "key_linear.weight"])
    v = F.linear(kv_in, P[pre + "value_linear.weight"])
    B, nq, H = q.shape
    nk, d = k.shape[1], H // nh
    q = q.view(B, nq, nh, d).permute(0, 2, 1, 3) * (d ** -0.5)
    k = k.view(B, nk, nh, d).permute(0, 2, 1, 3)
    v = v.view(B, nk, nh, d).permute(0, 2, 1, 3)
    logits = q @ k.transpose(-1, -2)
    logits = logits.masked_fill(mask.unsqueeze(1), -1e18)  # finite fill: fully masked rows -> uniform
    w = F.dropout(torch.softmax(logits, dim=-1), drop_p, training)
    ctx = (w @ v).permute(0, 2, 1, 3).reshape(B, nq, H)
    return F.linear(ctx, P[pre + "output_linear.weight"])


def ffn(P, pre, x, drop_p=0.0, training=False):
    """models/transformer_layers.py:400-408 — ReLU (+dropout) after BOTH linears (N1)."""
    x = F.dropout(torch.relu(F.linear(x, P[pre + "layers.0.weight"], P[pre + "layers.0.bias"])), drop_p, training)
    x = F.dropout(torch.relu(F.linear(x, P[pre + "layers.1.weight"], P[pre + "layers.1.bias"])), drop_p, training)
    return x


def ln(P, pre, x):
    return F.layer_norm(x, (x.shape[-1],), P[pre + "weight"], P[pre + "bias"], 1e-5)


def encoder(P, pre, x, pad_mask, nl, nh, drop=(0.0, 0.0), training=False):
    """models/transformer_layers.py:138-152 + EncoderLayer :260-282 (pre-LN)."""
    x = x + timing_signal(x.shape[1], x.shape[2]).to(x)
    for i in range(nl):
        lp = "%senc.%d." % (pre, i)
        x = x + mha(P, lp + "multi_head_attention.", ln(P, lp + "layer_norm_mha.", x), ln(P, lp + "layer_norm_mha.", x),
                    pad_mask, nh, drop[0], training)
        x = x + ffn(P, lp + "positionwise_feed_forward.", ln(P, lp + "layer_norm_ffn.", x), drop[1], training)
    return ln(P, pre + "layer_norm.", x)


def decoder(P, pre, x, enc_out, src_mask, trg_pad_mask, nl, nh, drop=(0.0, 0.0), training=False):
    """models/transformer_layers.py:205-221 + DecoderLayer :326-364."""
    T = x.shape[1]
    causal = torch.triu(torch.ones(T, T, dtype=torch.bool, device=x.device), 1)[None]
    dec_mask = trg_pad_mask | causal  # :207  (B,T,T)
    x = x + timing_signal(T, x.shape[2]).to(x)
    for i in range(nl):
        lp = "%sdec.%d." % (pre, i)
        xn = ln(P, lp + "layer_norm_mha_dec.", x)
        x = x + mha(P, lp + "multi_head_attention_dec.", xn, xn, dec_mask, nh, drop[0], training)
        xn = ln(P, lp + "layer_norm_mha_enc.", x)
        x = x + mha(P, lp + "multi_head_attention_enc_dec.", xn, enc_out, src_mask, nh, drop[0], training)
        x = x + ffn(P, lp + "positionwise_feed_forward.", ln(P, lp + "layer_norm_ffn.", x), drop[1], training)
    return ln(P, pre + "layer_norm.", x)


def seq3(P, pre, x):
    """Latent MLP: Linear, ReLU, Linear, ReLU, Linear (models/transformer_layers.py:21-39; dropout p=0)."""
    x = F.linear(x, P[pre + "0.weight"], P[pre + "0.bias"])
    x = F.linear(torch.relu(x), P[pre + "3.weight"], P[pre + "3.bias"])
    return F.linear(torch.relu(x), P[pre + "6.weight"], P[pre + "6.bias"])


def gaussian_kld(mq, lq, mp, lp):
    """models/transformer_layers.py:536-540."""
    return -0.5 * torch.sum(1 + (lq - lp) - (mp - mq) ** 2 / torch.exp(lp) - torch.exp(lq) / torch.exp(lp), dim=-1)


def latent(P, x, x_p, eps, L):
    """models/transformer_layers.py:41-59 with `eps` injected (reference draws it on the CPU, :45)."""
    pr = seq3(P, "latent_layer.mean_logvar_prior.", x)
    mp, lp = pr[:, :L], pr[:, L:]
    if x_p is None:
        return None, eps * torch.exp(0.5 * lp) + mp
    po = seq3(P, "latent_layer.mean_logvar_posterior.", torch.cat((x_p, x), dim=-1))
    mq, lq = po[:, :L], po[:, L:]
    kld = gaussian_kld(mq, lq, mp, lp).mean()
    return kld, eps * torch.exp(0.5 * lq) + mq


def image_head(P, feats, training=True, bn_state=None):
    """models/encoder_cnn.py:20-21,33-34 — fc + BatchNorm1d(momentum=0.01); trunk excluded."""
    y = F.linear(feats, P["encoder_cnn.cnn.fc.weight"], P["encoder_cnn.cnn.fc.bias"])
    rm = P["encoder_cnn.bn.running_mean"] if bn_state is None else bn_state[0]
    rv = P["encoder_cnn.bn.running_var"] if bn_state is None else bn_state[1]
    return F.batch_norm(y, rm, rv, P["encoder_cnn.bn.weight"], P["encoder_cnn.bn.bias"], training, 0.01, 1e-5)


def iq_forward(P, cfg, feats, context, response, target, eps=None, latent_mode=True, training=True,
               drop=(0.0, 0.0), bn_state=None):
    """models/iq.py:82-114 (+ encoder_transformer.py:22-37, decoder_transformer.py:22-41).
    Returns (output (B,T,V), z_logit (B,V)|None, kld|None, image_features, recon)."""
    nl, nh, L = cfg["num_layers"], cfg["num_heads"], cfg["latent_dim"]
    img = image_head(P, feats, training, bn_state)
    src_mask = context.eq(PAD).unsqueeze(1)
    enc_out = encoder(P, "answer_encoder.encoder.", embed(P, context), src_mask, nl, nh, drop, training)
    enc_out = torch.cat([(enc_out[:, 0] + img).unsqueeze(1), enc_out[:, 1:]], dim=1)  # encoder_transformer.py:32
    kld, z = None, None
    if latent_mode:
        res_mask = response.eq(PAD).unsqueeze(1)
        r_out = encoder(P, "answer_encoder.r_encoder.", embed(P, response), res_mask, nl, nh, drop, training)
        kld, z = latent(P, enc_out[:, 0], r_out[:, 0], eps, L)
        z = F.linear(z, P["latent_projection.weight"], P["latent_projection.bias"])  # iq.py:106
    sos = torch.full_like(target[:, :1], SOQ)
    shifted = torch.cat((sos, target[:, :-1]), 1)  # decoder_transformer.py:24-27
    trg_mask = shifted.eq(PAD).unsqueeze(1)
    temb = embed(P, shifted)
    tok0 = temb[:, 0] + img
    z_logit = None
    if latent_mode:
        tok0 = tok0 + z
        z_logit = F.linear(z + img, P["decoder.z_classifier.weight"], P["decoder.z_classifier.bias"])
    temb = torch.cat([tok0.unsqueeze(1), temb[:, 1:]], dim=1)
    dec = decoder(P, "decoder.decoder.", temb, enc_out, src_mask, trg_mask, nl, nh, drop, training)
    output = F.linear(dec, P["decoder.output.weight"], P["decoder.output.bias"])
    rin = enc_out[:, 0] + z if latent_mode else enc_out[:, 0]
    h = torch.relu(F.linear(rin, P["image_reconstructor.layers.fc0.weight"], P["image_reconstructor.layers.fc0.bias"]))
    recon = F.linear(h, P["image_reconstructor.layers.fc1.weight"], P["image_reconstructor.layers.fc1.bias"])
    return output, z_logit, kld, img, recon


def iq_losses(cfg, output, z_logit, kld, img, recon, target, kliter=0, latent_mode=True):
    """train_iq.py:81-103.  Returns (loss, dict of terms)."""
    V = output.shape[-1]
    rec = F.cross_entropy(output.reshape(-1, V), target.reshape(-1), ignore_index=PAD)
    l_img = F.mse_loss(img, recon)
    if not latent_mode:
        loss = rec + cfg["image_recon_lambda"] * l_img
        return loss, {"rec": rec, "img": l_img, "kld": torch.zeros(()), "aux": torch.zeros(()), "loss": loss}
    zl = z_logit.unsqueeze(1).repeat(1, output.shape[1], 1)
    aux = F.cross_entropy(zl.reshape(-1, V), target.reshape(-1), ignore_index=PAD)
    kl_w = min(math.tanh(6 * kliter / cfg["full_kl_step"] - 3) + 1, 1)
    loss = rec + cfg["kl_ceiling"] * kl_w * kld + cfg["aux_ceiling"] * aux + cfg["image_recon_lambda"] * l_img
    return loss, {"rec": rec, "img": l_img, "kld": kld, "aux": aux, "loss": loss}


def noam_lr(step, hidden_dim, warmup=4000):
    """train_iq.py:252-257."""
    return math.sqrt(1 / hidden_dim) * min(math.sqrt(1 / (step + 1)), step * warmup ** -1.5)


def greedy_decode(P, cfg, feats, context, eps=None, latent_mode=True, max_decode_length=50):
    """models/iq.py:117-152 — eval mode; prefix starts with PAD (iq.py:129); no KV cache.
    Returns ids (B, max_decode_length+1) and top-6 ids/probs per step."""
    nl, nh, L = cfg["num_layers"], cfg["num_heads"], cfg["latent_dim"]
    img = image_head(P, feats, training=False)
    src_mask = context.eq(PAD).unsqueeze(1)
    enc_out = encoder(P, "answer_encoder.encoder.", embed(P, context), src_mask, nl, nh)
    enc_out = torch.cat([(enc_out[:, 0] + img).unsqueeze(1), enc_out[:, 1:]], dim=1)
    z = 0
    if latent_mode:
        _, z = latent(P, enc_out[:, 0], None, eps, L)
        z = F.linear(z, P["latent_projection.weight"], P["latent_projection.bias"])
    ys = torch.full((context.shape[0], 1), PAD, dtype=torch.long)
    top_i, top_v = [], []
    for _ in range(max_decode_length + 1):
        emb = embed(P, ys)
        emb = torch.cat([(emb[:, 0] + z + img).unsqueeze(1), emb[:, 1:]], dim=1)  # decoder_transformer.py:46
        dec = decoder(P, "decoder.decoder.", emb, enc_out, src_mask, ys.eq(PAD).unsqueeze(1), nl, nh)
        logit = F.linear(dec[:, -1], P["decoder.output.weight"], P["decoder.output.bias"])
        nxt = logit.argmax(dim=1)
        v6, i6 = torch.topk(torch.softmax(logit, -1), 6, dim=1)
        top_i.append(i6); top_v.append(v6)
        ys = torch.cat([ys, nxt[:, None]], dim=1)
    return ys[:, 1:], torch.stack(top_i, 1), torch.stack(top_v, 1)
