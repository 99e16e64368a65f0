"""`LSTMIQ` — the LSTM variant of IQ (upstream IQ lineage) behind the staged API that `old_train_iq.py:100-155` and
`evaluate_iq.py:52-54` call: encode_images / encode_answers / encode_into_z / reparameterize / decode_questions /
reconstruct_inputs, plus the fused training entry `forward_losses` (forward + losses + complete hand-written backward).

Blocks and the reference code each replaces:
  image head      fc + BatchNorm1d                           models/encoder_cnn.py:20-21,33-34
  answer encoder  Embedding -> LSTM (packed by length)       models/encoder_rnn.py:62-87
  MLPs            Linear -> ReLU -> Linear                   models/mlp.py:14-56
  decoder         Embedding -> LSTM -> Linear(H,V) -> log_softmax, teacher-forced or free-running   models/decoder_rnn.py:99-162
  KL to N(0,1)    utils/train_utils.py:148-159
The wiring (z from MLP(cat(img, ans)); h0 = c0 = MLP(img + W z); reconstructions from a second reparameterisation of the
DETACHED features) is the call-site contract restated in oracle/lstm_oracle.py (the glue is not in the reference tree).

All arithmetic runs in libb200iq.so.  The recurrence is time-major: one GEMM projects the inputs of all T steps; the T
dependent steps then run inside ONE persistent cluster kernel (lstm_seq.cu: W_hh resident in shared memory, forward and
BPTT each one launch) wherever b200iq_lstm_seq_supported(H) — the per-step pair "tcgen05 GEMM (h.W_hh^T + x-projection,
residual epilogue) + cell kernel" is the fallback for other widths and for exact mode; every weight gradient is one GEMM
over all T*B rows.  State-dict names follow the upstream module tree (see oracle/lstm_oracle.py).
"""
import math
from types import SimpleNamespace

import os

import torch
import torch.nn as nn

from . import _lib, ops
from . import blocks as tf
from .params import Arena, export_state_dict, import_state_dict

PAD = 0


def _planes_mode(fn):
    """public entry points run with the library's TF32 write-rounding off (fp32 tensors stay exact; precision is decided per GEMM
    operand by its fp16 planes, and the persistent recurrence takes its 3-term form)"""
    import functools

    @functools.wraps(fn)
    def wrapped(*a, **k):
        with ops.no_rounding():
            return fn(*a, **k)
    return wrapped


TERMS = ("gen", "kl", "rec_i", "rec_a", "count", "unused0", "unused1", "unused2")


def lstm_iq_specs(a, V, H, Z, A, FEAT, num_categories=0):
    """parameter inventory in forward order (reverse = gradient-ready order); num_categories > 0 appends the category space
    (encode_categories / encode_into_t / predict_from_category, old_train_iq.py:134-142,194-213) AFTER everything else, so
    the offsets of the answer space do not depend on it"""
    k = 1.0 / math.sqrt(H)
    a.add("encoder_cnn.cnn.fc.weight", (H, FEAT), ("normal", 0.02))      # encoder_cnn.py:27-28
    a.add("encoder_cnn.cnn.fc.bias", (H,), ("zeros", 0))
    a.layernorm("encoder_cnn.bn", H)
    for rnn, emb in (("answer_encoder", True),):
        a.add("answer_encoder.embedding.weight", (V, H), ("uniform", 0.1))   # encoder_rnn.py:56
        for nm, shp in (("weight_ih_l0", (4 * H, H)), ("weight_hh_l0", (4 * H, H)), ("bias_ih_l0", (4 * H,)), ("bias_hh_l0", (4 * H,))):
            a.add("answer_encoder.rnn." + nm, shp, ("uniform", k))
    mlp_specs(a, "answer_attention", 2 * H, A, H)
    a.add("mu_answer_encoder.weight#pack", (Z, H), ("linear", H))        # mu | logvar: one [2Z, H] GEMM operand
    a.add("logvar_answer_encoder.weight", (Z, H), ("linear", H))
    a.add("mu_answer_encoder.bias#pack", (Z,), ("linear_bias", H))
    a.add("logvar_answer_encoder.bias", (Z,), ("linear_bias", H))
    a.linear("z_decoder", H, Z)
    mlp_specs(a, "gen_decoder", H, A, H)
    a.add("decoder.embedding.weight", (V, H), ("uniform", 0.1))          # decoder_rnn.py:95
    for nm, shp in (("weight_ih_l0", (4 * H, H)), ("weight_hh_l0", (4 * H, H)), ("bias_ih_l0", (4 * H,)), ("bias_hh_l0", (4 * H,))):
        a.add("decoder.rnn." + nm, shp, ("uniform", k))
    a.add("decoder.out.weight", (V, H), ("uniform", 0.1))                # decoder_rnn.py:96-97
    a.add("decoder.out.bias", (V,), ("zeros", 0))
    mlp_specs(a, "image_reconstructor", Z, A, H)
    mlp_specs(a, "answer_reconstructor", Z, A, H)
    if num_categories > 0:
        a.add("category_encoder.weight", (num_categories, H), ("normal", 1.0))   # nn.Embedding default init
        mlp_specs(a, "category_attention", 2 * H, A, H)
        a.add("mu_category_encoder.weight#pack", (Z, H), ("linear", H))       # mu | logvar: one [2Z, H] GEMM operand
        a.add("logvar_category_encoder.weight", (Z, H), ("linear", H))
        a.add("mu_category_encoder.bias#pack", (Z,), ("linear_bias", H))
        a.add("logvar_category_encoder.bias", (Z,), ("linear_bias", H))
        a.linear("t_decoder", H, Z)


def mlp_specs(a, name, i, hdim, o):
    """models/mlp.py:37-39 (2 layers)"""
    a.add(name + ".layers.fc0.weight", (hdim, i), ("he", i))
    a.add(name + ".layers.fc0.bias", (hdim,), ("zeros", 0))
    a.add(name + ".layers.fc1.weight", (o, hdim), ("he", hdim))
    a.add(name + ".layers.fc1.bias", (o,), ("zeros", 0))


# persistent recurrence kernels (one launch per LSTM pass) when the width fits; B200IQ_LSTM_PERSISTENT=0 = per-step launches
PERSISTENT = os.environ.get("B200IQ_LSTM_PERSISTENT", "1") == "1"


class LSTMIQ(nn.Module):
    KEY_LENGTH, KEY_SEQUENCE = "length", "sequence"  # models/decoder_rnn.py:51-52

    def __init__(self, vocab_size, max_len, hidden_size, num_categories=16, sos_id=1, eos_id=3, num_layers=1, rnn_cell="LSTM",
                 dropout_p=0.0, input_dropout_p=0.0, encoder_max_len=4, embedding=None, num_att_layers=2, att_ff_size=512,
                 z_size=100, feat_dim=2048, no_answer_recon=False, no_image_recon=False, no_category_space=True,
                 lambda_gen=1.0, lambda_z=0.001, lambda_i=0.001, lambda_a=0.001):
        super().__init__()
        if rnn_cell.upper() != "LSTM" or num_layers != 1 or num_att_layers != 2:
            raise NotImplementedError("LSTMIQ implements the default upstream configuration: 1-layer LSTM, 2-layer MLPs")
        # category space: forward / inference only — the in-tree training loop never puts a t-term into the
        # loss it backpropagates (old_train_iq.py:399-444), only `evaluate` and the samplers use it (:134-142,194-213)
        self.no_category_space, self.num_categories = no_category_space, num_categories
        self.V, self.H, self.Z, self.A, self.FEAT = vocab_size, hidden_size, z_size, att_ff_size, feat_dim
        self.max_len, self.sos_id, self.eos_id = max_len, sos_id, eos_id
        # model widths with 16-byte rows; the vocabulary size and the feature width are arbitrary (padded row pitches)
        assert self.H % 4 == 0 and self.Z % 4 == 0 and self.A % 4 == 0, "hidden_size, z_size and att_ff_size must be multiples of 4"
        self.input_dropout_p = input_dropout_p
        self.no_answer_recon, self.no_image_recon = no_answer_recon, no_image_recon
        self.lambdas = dict(lambda_gen=lambda_gen, lambda_z=lambda_z, lambda_i=lambda_i, lambda_a=lambda_a)
        self.arena = Arena()
        lstm_iq_specs(self.arena, self.V, self.H, self.Z, self.A, self.FEAT, 0 if no_category_space else num_categories)
        self.flat = nn.Parameter(self.arena.finalize())
        self.arena.bind(self.flat.data)
        if embedding is not None:  # GloVe (V, H) tensor: old_train_iq.py:38-40 (copied, see SURVEY §8c on decoder_rnn.py:87-88)
            with torch.no_grad():
                self.arena.views["decoder.embedding.weight"].copy_(embedding)
        self.register_buffer("bn_running_mean", torch.zeros(self.H), persistent=False)
        self.register_buffer("bn_running_var", torch.ones(self.H), persistent=False)
        self.register_buffer("bn_num_batches_tracked", torch.zeros((), dtype=torch.long), persistent=False)
        # what train.TrainStep reads
        self.args = SimpleNamespace(num_pretraining_steps=-1)
        self.latent_transformer = True
        # step_idx counts this module's stochastic passes: every context (a training step, or a stage call of the staged API)
        # draws a fresh Philox key, like the reference's nn.Dropout / torch.randn draw fresh numbers on every call
        self.seed, self.seed_rank, self.step_idx = 1234, 0, 0
        self.seed_dev = None       # device int64 word added to every Philox seed (train.TrainStep's CUDA-graph path, see IQ.seed_dev)
        self.kld_coef_dev = None   # unused here (the LSTM variant's KL weight is constant); TrainStep sets it on both models
        self.feature_cache = None
        self.on_grad_ready = None
        self._ws, self._gflat, self._terms, self._wops, self._wops_exact = {}, None, None, None, None

    @classmethod
    def from_config(cls, cfg, **kw):
        return cls(cfg["V"], cfg["T"], cfg["H"], z_size=cfg["Z"], att_ff_size=cfg["A"], feat_dim=cfg["FEAT"], **kw)

    # ------------------------------------------------------------------ module plumbing
    def _apply(self, fn, *a, **k):
        super()._apply(fn, *a, **k)
        self.arena.bind(self.flat.data)
        self._ws, self._gflat, self._terms, self._wops, self._wops_exact = {}, None, None, None, None
        return self

    def _buffers_map(self):
        return {"encoder_cnn.bn.running_mean": self.bn_running_mean, "encoder_cnn.bn.running_var": self.bn_running_var,
                "encoder_cnn.bn.num_batches_tracked": self.bn_num_batches_tracked}

    def state_dict(self, *args, destination=None, prefix="", keep_vars=False):
        sd = export_state_dict(self.arena, self._buffers_map(), (), prefix)
        out = destination if destination is not None else type(sd)()
        for k, v in sd.items():
            out[k] = v if keep_vars else v.detach()
        return out

    def load_state_dict(self, state_dict, strict=True):
        return import_state_dict(self.arena, self._buffers_map(), (), state_dict, strict)

    @property
    def grad_arena(self):
        if self._gflat is None:
            self._gflat = torch.zeros_like(self.flat.data)
        return self._gflat

    @property
    def terms(self):
        if self._terms is None:
            self._terms = torch.zeros(len(TERMS), device=self.flat.device)
        return self._terms

    def total_loss(self, terms, kliter=0):
        lam = self.lambdas
        return lam["lambda_gen"] * terms[0] + lam["lambda_z"] * terms[1] + lam["lambda_i"] * terms[2] + lam["lambda_a"] * terms[3]

    # ------------------------------------------------------------------ internals
    def _workspace(self, key):
        ws = self._ws.get(key)
        if ws is None:
            ws = tf.Workspace(self.flat.device)
            self._ws[key] = ws
        return ws

    def _make_ctx(self, key, training):
        _lib.require_device()
        ws = self._workspace(key + (ops.exact_mode(),))
        if ops.exact_mode():
            if self._wops_exact is None:
                self._wops_exact = tf.WeightOperands(self.arena)
            weights = self._wops_exact
        else:
            if self._wops is None:   # fp16 planes (hi, lo) of the weights; the embedding tables are gathered, never multiplied
                self._wops = tf.build_weight_operands(self.arena, self.flat.device, skip=tuple(n for n in self.arena.views if n.endswith("embedding.weight") or n == "category_encoder.weight"))
            tf.refresh_weight_operands(self._wops, self.arena, self.flat.data)
            weights = self._wops
        if self.seed_dev is None:
            self.step_idx += 1
        c = tf.Ctx(self.arena, self.grad_arena, ws, self.H, 1, self.A, self.H, 0.0, 0.0, training,
                   self.seed + (0 if self.seed_dev is not None else self.step_idx) + 1000003 * self.seed_rank, weights)
        c.on_ready = self.on_grad_ready
        c.in_drop = self.input_dropout_p if training else 0.0
        return c

    def _feats(self, c, images):
        if images.dim() == 1 and not images.is_floating_point():
            cache = getattr(self, "feature_cache", None)
            if cache is None:
                raise _lib.B200iqError("image rows given but no feature cache attached (set .feature_cache = batch.FeatureCache)")
            rows = self._ids(images, self.flat.device)
            return cache.gather(rows, c.ws.buf("feats", rows.numel(), cache.rows.shape[1]))
        f = images.to(self.flat.device, torch.float32, non_blocking=True).contiguous()
        if f.dim() != 2:
            raise _lib.B200iqError("LSTMIQ takes pooled trunk features (B, FEAT); run the frozen CNN trunk first")
        return f

    @staticmethod
    def _ids(t, device):
        if t.dtype != torch.int32:
            t = t.to(torch.int32)
        return t.to(device=device, non_blocking=True).contiguous()

    def _head_fwd(self, c, feats):
        """models/encoder_cnn.py:33-34"""
        ws, P, B, H = c.ws, c.P, feats.shape[0], self.H
        img_pre = ws.buf("img_pre", B, H)
        # split-K: K = FEAT = 2048 is the longest forward reduction and the tensor core's fp32 accumulation error grows with it
        tf.mm(c, feats, "encoder_cnn.cnn.fc.weight", img_pre, bias=P["encoder_cnn.cnn.fc.bias"], auto_split=True)
        img = ws.buf("img", B, H)
        ops.bn_fwd(img_pre, P["encoder_cnn.bn.weight"], P["encoder_cnn.bn.bias"], img, ws.buf("bn.mean", H), ws.buf("bn.rstd", H),
                   self.bn_running_mean, self.bn_running_var, c.training)
        return img

    def _lstm_fwd(self, c, tag, pfx, emb_name, ids_tm, T, B, h0, c0, lengths):
        """Embedding -> input dropout -> LSTM over T steps (time-major).  Returns hbuf [T+1,B,H] (slot 0 = h0)."""
        ws, P, H = c.ws, c.P, self.H
        e = ws.buf(tag + ".e", T * B, H)
        ops.embed_gather(ids_tm, P[emb_name], e)
        ops.dropout_(e, c.in_drop, *c.key(tag + ".indrop"))
        bsum = ws.buf(tag + ".bsum", 4 * H)
        ops.axpby(P[pfx + "bias_ih_l0"], P[pfx + "bias_hh_l0"], bsum)
        gates = ws.buf(tag + ".gates", T, B, 4 * H)
        tf.mm(c, e, pfx + "weight_ih_l0", gates.view(T * B, 4 * H), bias=bsum)  # x-projection of all steps
        hbuf, cbuf = ws.buf(tag + ".h", T + 1, B, H), ws.buf(tag + ".c", T + 1, B, H)
        if h0 is None:
            hbuf[0].zero_()
            cbuf[0].zero_()
        else:
            hbuf[0].copy_(h0)
            cbuf[0].copy_(c0)
        if PERSISTENT and ops.lstm_seq_supported(H):
            # the whole recurrence in one launch; the kernel splits its slice of the fp32 master W_hh (once) and h (per step) into fp16 hi / lo itself (3-term)
            ops.lstm_seq_fwd(gates, P[pfx + "weight_hh_l0"], hbuf, cbuf, lengths)
            return hbuf
        for t in range(T):
            tf.mm(c, hbuf[t], pfx + "weight_hh_l0", gates[t], residual=gates[t])
            ops.lstm_cell_fwd(gates[t], cbuf[t], hbuf[t], cbuf[t + 1], hbuf[t + 1], lengths, t)
        return hbuf

    def _lstm_bwd(self, c, tag, pfx, emb_name, ids_tm, T, B, dh_last, d_out, lengths):
        """BPTT.  dh_last: gradient wrt the final hidden state (or None); d_out: [T*B,H] gradient wrt every output (or None).
        Returns (d h0, d c0) in workspace buffers; accumulates weight / embedding gradients."""
        ws, G, H, P = c.ws, c.G, self.H, c.P
        gates, hbuf, cbuf, e = ws[tag + ".gates"], ws[tag + ".h"], ws[tag + ".c"], ws[tag + ".e"]
        dgates = ws.buf(tag + ".dgates", T, B, 4 * H)
        dh_run, dc = ws.buf(tag + ".dh_run", B, H), ws.buf(tag + ".dc", B, H)
        if dh_last is None:
            dh_run.zero_()
        else:
            dh_run.copy_(dh_last)
        dc.zero_()
        dov = d_out.view(T, B, H) if d_out is not None else None
        if PERSISTENT and ops.lstm_seq_supported(H):
            ops.lstm_seq_bwd(gates, cbuf, P[pfx + "weight_hh_l0"], dh_run, dov, dgates, dh_run, dc, lengths)   # BPTT in one launch
        else:
            for t in reversed(range(T)):
                ops.lstm_cell_bwd(gates[t], cbuf[t], cbuf[t + 1], dh_run, dov[t] if dov is not None else None, dc, dgates[t], lengths, t)
                tf.mm(c, dgates[t], pfx + "weight_hh_l0", dh_run, b_t=True, residual=dh_run)
        dg = dgates.view(T * B, 4 * H)
        dga = tf.act(c, dg, lo=False)   # one split of the gate gradients for the three products that read them
        def weight_grads():   # off the dX chain: side stream, under the input-gradient GEMM and the embedding scatter
            tf.grad_w(c, dg, hbuf[:T].view(T * B, H), G[pfx + "weight_hh_l0"], G[pfx + "bias_hh_l0"], dya=dga)
            tf.grad_w(c, dg, e, G[pfx + "weight_ih_l0"], G[pfx + "bias_ih_l0"], dya=dga, x_cached=True)
        ev = c.mark()
        de = ws.buf(tag + ".de", T * B, H)
        tf.mm(c, dga, pfx + "weight_ih_l0", de, b_t=True)
        c.on_side(weight_grads, ev)
        ops.dropout_(de, c.in_drop, *c.key(tag + ".indrop"))
        ops.embed_scatter_add(ids_tm, de, G[emb_name], -1)
        c.join_side()
        return dh_run, dc

    def _mlp_fwd(self, c, tag, name, x, fp32_out=False):
        h = tf.linear_fwd(c, tag + ".h", name + ".layers.fc0", x, relu=True)
        return tf.linear_fwd(c, tag + ".y", name + ".layers.fc1", h, fp32_out=fp32_out)

    def _mlp_bwd(self, c, tag, name, dy, x, need_dx=True, dx=None):
        h = c.ws[tag + ".h"]
        dh = tf.linear_bwd(c, tag + ".dh", name + ".layers.fc1", dy, h, relu_mask=h)
        return tf.linear_bwd(c, tag + ".dx", name + ".layers.fc0", dh, x, need_dx=need_dx, dx=dx)

    def _encode_answers_fwd(self, c, aids, alens, B, S):
        ids_tm = ops.transpose_ids(aids, c.ws.buf("enc.ids", S * B, dtype=torch.int32), 0, S)
        hbuf = self._lstm_fwd(c, "enc", "answer_encoder.rnn.", "answer_encoder.embedding.weight", ids_tm, S, B, None, None, alens)
        return hbuf[S]  # final hidden of every row (state stops updating at its length)

    def _z_params_fwd(self, c, img, ans, B, space="answer"):
        """encode_into_z (space="answer") / encode_into_t (space="category"): MLP(cat(img, other)) -> [mu | logvar] in one GEMM"""
        H = self.H
        cat = c.ws.buf("cat", B, 2 * H)
        ops.tok0_gather(img, cat[:, :H], B, 1, H)
        ops.tok0_gather(ans, cat[:, H:], B, 1, H)
        att = self._mlp_fwd(c, "att", space + "_attention", cat)
        wml = c.pspan("mu_%s_encoder.weight" % space, "logvar_%s_encoder.weight" % space, H)
        o0, o1 = self.arena.span("mu_%s_encoder.bias" % space, "logvar_%s_encoder.bias" % space)
        mulv = c.ws.buf("mulv", B, 2 * self.Z)
        tf.mm(c, att, wml, mulv, bias=self.flat.data[o0:o1])
        return cat, att, mulv

    def _init_hidden_fwd(self, c, img, z, B, proj="z_decoder"):
        """decode_questions prologue: h0 = c0 = gen_decoder(img + z_decoder(z)) (t_decoder for a sample of the category space)"""
        zh = tf.linear_fwd(c, "zh", proj, z)
        s = ops.axpby(img, zh, c.ws.buf("s", B, self.H))
        return s, self._mlp_fwd(c, "gd", "gen_decoder", s)

    # ------------------------------------------------------------------ fused training path
    def forward_losses(self, images, answers, alengths, questions, kliter=0, eps=None, backward=True, eps2=None):
        """forward + losses (old_train_iq.py:100-155 contract) + complete backward; returns the device vector
        (gen, kl, rec_i, rec_a, n_tokens, ...); gradients are ACCUMULATED into `grad_arena`.  `kliter` is unused (signature
        shared with IQ.forward_losses so train.TrainStep drives both)."""
        prev, ops.SEED_DEV = ops.SEED_DEV, self.seed_dev
        try:
            with ops.no_rounding():
                return self._forward_losses(images, answers, alengths, questions, eps, backward, eps2)
        finally:
            ops.SEED_DEV = prev

    def scale(self, tokens=None):
        """loss scale of the fp16-plane path (see IQ.scale): every gradient seed is multiplied by it, train.TrainStep divides it out"""
        if ops.exact_mode():
            return 1.0
        if tokens is None:
            return getattr(self, "_grad_scale", 1.0)
        return 16.0 * 2.0 ** math.floor(math.log2(max(1.0, tokens / 2.0)))

    def unscaled_grads(self):
        return self.grad_arena / getattr(self, "_grad_scale", 1.0)

    def _forward_losses(self, images, answers, alengths, questions, eps, backward, eps2):
        dev = self.flat.device
        aids, qids = self._ids(answers, dev), self._ids(questions, dev)
        alens = self._ids(alengths, dev)
        B, S, T = aids.shape[0], aids.shape[1], qids.shape[1]
        Td = T - 1
        c = self._make_ctx(("train", B, S, T, bool(self.training)), self.training)
        ws, P, G, H, Z, V = c.ws, c.P, c.G, self.H, self.Z, self.V
        SC = self._grad_scale = self.scale(B * Td)
        lam = {k: v * SC for k, v in self.lambdas.items()}   # every gradient seed carries the loss scale (the loss TERMS do not)
        terms = self.terms
        terms.zero_()
        feats = self._feats(c, images)
        # ---- forward
        img = self._head_fwd(c, feats)
        ans = self._encode_answers_fwd(c, aids, alens, B, S)
        cat, att, mulv = self._z_params_fwd(c, img, ans, B)
        zeros2z = ws.buf("zeros2z", B, 2 * Z)
        zeros2z.zero_()  # the unit-normal prior [mu=0 | logvar=0] of utils/train_utils.py:148-159
        e1, e2 = ws.buf("eps", B, Z), ws.buf("eps2", B, Z)
        if eps is not None:
            e1.copy_(eps)
        else:
            ops.randn_(e1, c.seed, 7 << 40)
        if eps2 is not None:
            e2.copy_(eps2)
        else:
            ops.randn_(e2, c.seed, 8 << 40)
        z = ws.buf("z", B, Z)
        ops.latent_fwd(zeros2z, mulv, e1, z, terms[1:2])
        s, hid = self._init_hidden_fwd(c, img, z, B)
        ids_q = ops.transpose_ids(qids, ws.buf("dec.ids", Td * B, dtype=torch.int32), 0, Td)   # inputs[:, :-1]
        tgt = ops.transpose_ids(qids, ws.buf("dec.tgt", Td * B, dtype=torch.int32), 1, Td)     # questions[:, 1:]
        hbuf = self._lstm_fwd(c, "dec", "decoder.rnn.", "decoder.embedding.weight", ids_q, Td, B, hid, hid, None)
        outs = hbuf[1:].view(Td * B, H)
        # ---- vocabulary projection + log-softmax + NLL, chunked so the logits never exist in HBM as a whole
        M = Td * B
        count = terms[4:5]
        ops.count_nonpad(tgt, PAD, count)
        d_out = ws.buf("d_out", M, H)
        rows = max(128, ((48 << 20) // (4 * V)) // 128 * 128)
        chunk = min(M, rows)
        Vp = ops.pad8(V)   # row pitch of the chunk buffers: any vocabulary size keeps 16-byte rows
        logits = ws.buf("logits_chunk", chunk, Vp)[:, :V]
        # the chunk's gradient goes out as an fp16 plane beside the fp32 logits (exact mode: in place); the vocabulary products are
        # smooth (no ReLU behind them): single pass
        lgrad = ops.Act(f32=logits) if ops.exact_mode() else ops.Act(hi=ws.buf("logits_grad", chunk, Vp, dtype=torch.float16)[:, :V])
        outs_a = tf.act(c, outs, lo=False)
        wout, bout = c.W["decoder.out.weight"], P["decoder.out.bias"]
        for r0 in range(0, M, chunk):
            n = min(chunk, M - r0)
            lg, xs, gr = logits[:n], outs_a.rows(r0, r0 + n), lgrad.rows(0, n)
            ops.gemm2(xs, wout, lg, bias=bout, terms=1)                                # decoder_rnn.py:107
            ops.ce_rows_act(lg, tgt[r0:r0 + n], PAD, count, lam["lambda_gen"], terms[0:1], gr if backward else None)
            if backward:
                ev = c.mark()
                ops.gemm2(gr, wout, d_out[r0:r0 + n], b_t=True, auto_split=True, terms=1)
                c.on_side(lambda: tf.grad_w(c, gr, xs, G["decoder.out.weight"], G["decoder.out.bias"]), ev)
                c.join_side()   # before the next chunk's gradient overwrites gr
        # ---- reconstructions from the detached features through a second reparameterisation (old_train_iq.py:145-155)
        z2 = ws.buf("z2", B, Z)
        ops.latent_fwd(zeros2z, mulv, e2, z2, None)
        d_z2 = ws.buf("d_z2", B, Z)
        d_z2_used = False
        for tag, name, target, ti, coef, off in (("ri", "image_reconstructor", img, 2, lam["lambda_i"], self.no_image_recon),
                                                 ("ra", "answer_reconstructor", ans, 3, lam["lambda_a"], self.no_answer_recon)):
            if off:
                continue
            rec = self._mlp_fwd(c, tag, name, z2, fp32_out=True)
            d_rec = ws.buf(tag + ".d", B, H)
            ops.mse(target, rec, coef, terms[ti:ti + 1], None, d_rec if backward else None)
            if backward:
                if d_z2_used:
                    tmp = self._mlp_bwd(c, tag, name, d_rec, z2)
                    ops.axpby(d_z2, tmp, d_z2)
                else:
                    self._mlp_bwd(c, tag, name, d_rec, z2, dx=d_z2)
                    d_z2_used = True
        if not backward:
            return terms
        c.ready("decoder.out.weight")
        # ---- backward: decoder recurrence -> initial hidden -> z
        dh0, dc0 = self._lstm_bwd(c, "dec", "decoder.rnn.", "decoder.embedding.weight", ids_q, Td, B, None, d_out, None)
        d_hid = ops.axpby(dh0, dc0, ws.buf("d_hid", B, H))
        c.ready("decoder.embedding.weight")
        d_s = self._mlp_bwd(c, "gd", "gen_decoder", d_hid, s)
        d_z = tf.linear_bwd(c, "d_z", "z_decoder", d_s, z)
        d_img = ws.buf("d_img", B, H)
        ops.axpby(d_s, d_s, d_img, 1.0, 0.0)
        # ---- [mu | logvar] layer: generation/KL path (flows on to img and ans) and reconstruction path (stops at the features)
        wml = c.pspan("mu_answer_encoder.weight", "logvar_answer_encoder.weight", H)
        gml = c.gspan("mu_answer_encoder.weight", "logvar_answer_encoder.weight", H)
        o0, o1 = self.arena.span("mu_answer_encoder.bias", "logvar_answer_encoder.bias")
        gmlb = c.gflat[o0:o1]
        dprior = ws.buf("d_prior_unused", B, 2 * Z)
        paths = [("gen", e1, d_z, lam["lambda_z"], True)]
        if d_z2_used:
            paths.append(("rec", e2, d_z2, 0.0, False))
        for tag, ee, dzz, kc, flows in paths:
            dpost = ws.buf("d_mulv." + tag, B, 2 * Z)
            ops.latent_bwd(zeros2z, mulv, ee, dzz, kc, dprior, dpost)
            dpa = tf.act(c, dpost, lo=False)
            tf.grad_w(c, dpost, att, gml, gmlb, dya=dpa, x_cached=True)
            d_att = ws.buf("d_att." + tag, B, H)
            tf.mm(c, dpa, wml, d_att, b_t=True)
            if flows:
                d_cat = self._mlp_bwd(c, "att", "answer_attention", d_att, cat)
            else:
                self._mlp_bwd(c, "att", "answer_attention", d_att, cat, need_dx=False)
        # d_cat is [B, 2H]: its halves are column slices with row pitch 2H = "token 0 of a 2-token sequence"
        ops.tok0_gather(d_cat[:, :H], d_img, B, 2, H, accumulate=True)
        d_ans = ws.buf("d_ans", B, H)
        ops.tok0_gather(d_cat[:, H:], d_ans, B, 2, H)
        c.ready("answer_attention.layers.fc0.weight")
        # ---- answer encoder recurrence, image head
        self._lstm_bwd(c, "enc", "answer_encoder.rnn.", "answer_encoder.embedding.weight", ws["enc.ids"], S, B, d_ans, None, alens)
        d_img_pre = ws.buf("d_img_pre", B, H)
        ops.bn_bwd(d_img, ws["img_pre"], P["encoder_cnn.bn.weight"], ws["bn.mean"], ws["bn.rstd"], d_img_pre,
                   G["encoder_cnn.bn.weight"], G["encoder_cnn.bn.bias"])
        tf.grad_w(c, d_img_pre, feats, G["encoder_cnn.cnn.fc.weight"], G["encoder_cnn.cnn.fc.bias"], x_cached=True)
        return terms


    # ------------------------------------------------------------------ staged API (old_train_iq.py:100-106,145-149; evaluate_iq.py:52-54)
    # Forward-only kernels returning plain tensors: evaluation, sampling and the per-stage parity tests use these.  Training
    # goes through forward_losses (the hand-written backward is not exposed through autograd for the staged calls).
    def _stage_ctx(self, tag, B):
        return self._make_ctx(("stage", tag, B, bool(self.training)), self.training)

    @torch.no_grad()
    @_planes_mode
    def encode_images(self, images):
        c = self._stage_ctx("img", images.shape[0])
        return self._head_fwd(c, self._feats(c, images)).clone()

    @torch.no_grad()
    @_planes_mode
    def encode_answers(self, answers, alengths):
        dev = self.flat.device
        aids = self._ids(answers, dev)
        alens = self._ids(torch.as_tensor(alengths), dev)
        B, S = aids.shape
        c = self._stage_ctx("ans", B)
        return self._encode_answers_fwd(c, aids, alens, B, S).clone()

    @torch.no_grad()
    @_planes_mode
    def encode_into_z(self, image_features, answer_features):
        B = image_features.shape[0]
        c = self._stage_ctx("z", B)
        _, _, mulv = self._z_params_fwd(c, image_features.contiguous(), answer_features.contiguous(), B)
        return mulv[:, :self.Z].clone(), mulv[:, self.Z:].clone()

    @torch.no_grad()
    @_planes_mode
    def reparameterize(self, mus, logvars, eps=None):
        """train: mu + eps*exp(.5*logvar) with Philox eps (or the injected one); eval: mu"""
        if not self.training:
            return mus
        B = mus.shape[0]
        c = self._stage_ctx("rep", B)
        e = c.ws.buf("eps", B, self.Z)
        if eps is not None:
            e.copy_(eps)
        else:
            ops.randn_(e, c.seed, 9 << 40)
        post = c.ws.buf("post", B, 2 * self.Z)
        post[:, :self.Z].copy_(mus)
        post[:, self.Z:].copy_(logvars)
        zeros = c.ws.buf("zeros", B, 2 * self.Z)
        zeros.zero_()
        z = c.ws.buf("z", B, self.Z)
        ops.latent_fwd(zeros, post, e, z, None)
        return z.clone()

    @torch.no_grad()
    @_planes_mode
    def reconstruct_inputs(self, image_features, answer_features, eps=None):
        mus, logvars = self.encode_into_z(image_features, answer_features)
        zs = self.reparameterize(mus, logvars, eps)
        c = self._stage_ctx("rec", zs.shape[0])
        ri = None if self.no_image_recon else self._mlp_fwd(c, "ri", "image_reconstructor", zs.contiguous(), fp32_out=True).clone()
        ra = None if self.no_answer_recon else self._mlp_fwd(c, "ra", "answer_reconstructor", zs.contiguous(), fp32_out=True).clone()
        return ri, ra

    # ---- category space (call sites old_train_iq.py:134-142,194-213, evaluate_iq.py:51-54) -- forward only
    def _need_category_space(self):
        if self.no_category_space:
            raise _lib.B200iqError("this LSTMIQ was built with no_category_space=True")

    @torch.no_grad()
    @_planes_mode
    def encode_categories(self, categories):
        """nn.Embedding(num_categories, H) lookup: categories (B,) int -> (B, H)"""
        self._need_category_space()
        ids = self._ids(categories.reshape(-1), self.flat.device)
        c = self._stage_ctx("cat", ids.numel())
        out = c.ws.buf("catf", ids.numel(), self.H)
        return ops.embed_gather(ids, c.P["category_encoder.weight"], out).clone()

    @torch.no_grad()
    @_planes_mode
    def encode_into_t(self, image_features, category_features):
        """(t_mus, t_logvars) = the category space's Gaussian, same shape of computation as encode_into_z"""
        self._need_category_space()
        B = image_features.shape[0]
        c = self._stage_ctx("t", B)
        _, _, mulv = self._z_params_fwd(c, image_features.contiguous(), category_features.contiguous(), B, "category")
        return mulv[:, :self.Z].clone(), mulv[:, self.Z:].clone()

    @torch.no_grad()
    @_planes_mode
    def _kl(self, prior, post):
        """mean_b KL(N(post) || N(prior)) through b200iq_latent_fwd; prior / post are (mu, logvar) pairs, prior None = unit normal"""
        B = post[0].shape[0]
        c = self._stage_ctx("kl", B)
        q, p = c.ws.buf("kl.q", B, 2 * self.Z), c.ws.buf("kl.p", B, 2 * self.Z)
        q[:, :self.Z].copy_(post[0]); q[:, self.Z:].copy_(post[1])
        if prior is None:
            p.zero_()
        else:
            p[:, :self.Z].copy_(prior[0]); p[:, self.Z:].copy_(prior[1])
        e = c.ws.buf("kl.e", B, self.Z)
        e.zero_()
        out = torch.zeros(1, device=q.device)
        ops.latent_fwd(p, q, e, c.ws.buf("kl.z", B, self.Z), out)
        return out[0]

    def gaussian_KL_loss(self, mus, logvars):
        """utils/train_utils.py:148-159: KL to the unit normal, summed over elements, / B"""
        return self._kl(None, (mus, logvars))

    def compute_two_gaussian_loss(self, mu1, logvar1, mu2, logvar2):
        """old_train_iq.py:262-278: KL(N(mu1, e^logvar1) || N(mu2, e^logvar2)) summed over elements, / B (the reference adds 1e-8
        to e^logvar2 and to B)"""
        return self._kl((mu2, logvar2), (mu1, logvar1))

    @torch.no_grad()
    def _predict(self, image_features, zs, from_t):
        _, _, ret = self.decode_questions(image_features, zs, questions=None, teacher_forcing_ratio=0, from_t=from_t)
        return torch.cat(ret[self.KEY_SEQUENCE], dim=1)

    @torch.no_grad()
    def predict_from_answer(self, images, answers, alengths, eps=None):
        """evaluate_iq.py:51-52: (B, max_len) greedy token ids from the answer space (eval mode decodes from the mean)"""
        img = self.encode_images(images)
        mus, logvars = self.encode_into_z(img, self.encode_answers(answers, alengths))
        return self._predict(img, self.reparameterize(mus, logvars, eps), False)

    @torch.no_grad()
    def predict_from_category(self, images, categories, eps=None):
        """evaluate_iq.py:53-54, old_train_iq.py:194-213: (B, max_len) greedy token ids from the category space"""
        img = self.encode_images(images)
        t_mus, t_logvars = self.encode_into_t(img, self.encode_categories(categories))
        return self._predict(img, self.reparameterize(t_mus, t_logvars, eps), True)

    @torch.no_grad()
    @_planes_mode
    def decode_questions(self, image_features, zs, questions=None, teacher_forcing_ratio=0, decode_function=None, from_t=False):
        """models/decoder_rnn.py:110-162 behind the glue: returns (outputs, hidden, ret_dict) with `outputs` a list of (B, V)
        log-probability tensors and ret_dict = {sequence: list of (B,1) argmax ids, length: list of ints}.
        teacher_forcing_ratio >= 1 with `questions`: one pass over questions[:, :-1]; otherwise free-running greedy decoding for
        max_len steps from <start> (the reference draws random.random() < ratio once per call; only 0 and 1 are meaningful)."""
        dev = self.flat.device
        B, H, V = zs.shape[0], self.H, self.V
        c = self._stage_ctx("dec", B)
        ws = c.ws
        if from_t:
            self._need_category_space()
        _, hid = self._init_hidden_fwd(c, image_features.contiguous(), zs.contiguous(), B, "t_decoder" if from_t else "z_decoder")
        outputs, symbols = [], []
        if questions is not None and teacher_forcing_ratio >= 1:
            qids = self._ids(questions, dev)
            Td = qids.shape[1] - 1
            ids_q = ops.transpose_ids(qids, ws.buf("dec.ids", Td * B, dtype=torch.int32), 0, Td)
            hbuf = self._lstm_fwd(c, "dec", "decoder.rnn.", "decoder.embedding.weight", ids_q, Td, B, hid, hid, None)
            logp = ws.buf("logp", Td * B, V)
            tf.mm(c, hbuf[1:].view(Td * B, H), "decoder.out.weight", logp, bias=c.P["decoder.out.bias"], terms=1)
            ops.log_softmax_rows_(logp)
            top = ws.buf("top", Td * B, dtype=torch.int32)
            ops.topk_softmax_rows(logp, 1, top)
            for t in range(Td):
                outputs.append(logp[t * B:(t + 1) * B].clone())
                symbols.append(top[t * B:(t + 1) * B].long().view(B, 1))
            h_last, c_last = hbuf[Td].clone(), ws["dec.c"][Td].clone()
        else:
            steps = self.max_len
            tok = ws.buf("tok", B, dtype=torch.int32)
            tok.fill_(self.sos_id)
            h, cc = hid.clone(), hid.clone()
            for _ in range(steps):
                hb = self._lstm_fwd(c, "dec1", "decoder.rnn.", "decoder.embedding.weight", tok, 1, B, h, cc, None)
                h, cc = hb[1].clone(), ws["dec1.c"][1].clone()
                logp = ws.buf("logp1", B, V)
                tf.mm(c, h, "decoder.out.weight", logp, bias=c.P["decoder.out.bias"], terms=1)
                ops.log_softmax_rows_(logp)
                ops.topk_softmax_rows(logp, 1, tok)
                outputs.append(logp.clone())
                symbols.append(tok.long().view(B, 1))
            h_last, c_last = h, cc
        # lengths bookkeeping of decoder_rnn.py:120-135 (host side, one D2H at the end instead of one .cpu() per step)
        seq = torch.cat(symbols, dim=1).cpu()
        n = seq.shape[1]
        lengths = []
        for b in range(B):
            hit = (seq[b] == self.eos_id).nonzero()
            lengths.append(int(hit[0]) + 1 if len(hit) else n)
        ret = {self.KEY_SEQUENCE: symbols, self.KEY_LENGTH: lengths}
        return outputs, (h_last.unsqueeze(0), c_last.unsqueeze(0)), ret
