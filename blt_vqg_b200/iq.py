"""`IQ` — drop-in for the reference's models/iq.py:22 `IQ` module (transformer / BLT-VQG variant).

Same constructor, attribute surface and return values as the reference:
  IQ(latent_transformer, vocab, args, num_att_layers=2)                      models/iq.py:25
  forward(images, answers, response, target) -> (output, z_logit, kld, (img, recon))   :82-114
  switch_GVT_train_mode(bool)                                                :51-54
  decode_greedy(images, answers, max_decode_length=50)                       :117-152   (see decode.py)
plus the fused entry the reference cannot offer:
  forward_losses(images, answers, response, target, kliter) -> loss-term vector; runs forward, the fused
  vocab-projection + cross-entropy (no (B,T,V) logits in HBM) and the complete hand-written backward,
  leaving gradients in the flat gradient arena.
All arithmetic runs in libb200iq.so (sm_100a); torch is used for memory, streams and autograd glue only.
`images` may be pre-pooled trunk features (B, FEAT) (trunk-cut mode) or raw images (B,3,224,224), in which
case the frozen torchvision trunk of models/encoder_cnn.py:17-19 runs under no_grad first.
"""
import math
import os

import numpy as np
import torch
import torch.nn as nn

from . import _lib, ops
from . import transformer as tf
from .params import Arena, TRANSFORMER_ALIASES, export_state_dict, import_state_dict, transformer_iq_specs

PAD = 0
# B200IQ_PDL=2: programmatic dependent launch for the whole forward and the vocabulary loop (=1: the whole step, train.py;
# the default =3 covers the decoder's forward only, transformer.py).  Measured on cfg3 the wider scopes cost 0.5 ms (=2) /
# 1.2 ms (=1) per step, because a dependent kernel's CTAs hold their SMs while they wait, and those are the SMs the side-stream
# kernels would otherwise fill.
PDL_FORWARD = os.environ.get("B200IQ_PDL", "3") == "2"
# run the 5-token context encoder on a side stream next to the 21-token response encoder (forward and backward): its GEMMs
# have 2560 rows = 20-40 tiles for 148 SMs, so on its own it leaves three quarters of the GPU idle
CONCURRENT_ENCODERS = os.environ.get("B200IQ_CONCURRENT", "1") == "1"
# B200IQ_DEC_SPLIT=n (default 1 = off): the decoder stack as n independent chains over n equal sample ranges of the batch, each on
# its own stream, forward and backward (WorkspacePart: every chain works on its row range of the same full-size buffers).
# The decoder is the one serial phase of the step (timeline profiles/r02m_timeline_summary.txt: one kernel resident for the
# 2.25 ms of its forward), but cutting it does not pay on cfg3: half-size products run one tile round on 40 of the 74 SM pairs
# in almost the time the full-size ones need for two, and two 80-CTA persistent grids do not fit the 148 SMs side by side —
# 11.15 ms per step with n = 2, 11.85 with n = 4, against 10.92 (profiles/r02m_decoder_split.txt).  Kept as a switch.
DEC_SPLIT = int(os.environ.get("B200IQ_DEC_SPLIT", "1"))
DEC_SPLIT_MIN_ROWS = int(os.environ.get("B200IQ_DEC_SPLIT_MIN_ROWS", "2048"))
# B200IQ_NVTX=1: NVTX ranges around the phases of a step (forward / vocabulary CE / backward, and one per stack), for timelines
NVTX = os.environ.get("B200IQ_NVTX", "0") == "1"
CE_CHUNK_MB = int(os.environ.get("B200IQ_CE_CHUNK_MB", "80"))  # size of one logits chunk of the fused vocabulary projection + CE
TERMS = ("rec", "img", "kld", "aux", "count", "gnorm_sq", "unused0", "unused1")


def gen_timing_signal(length, channels, min_timescale=1.0, max_timescale=1.0e4):
    """models/transformer_layers.py:542-558: [sin | cos] halves, float64 math then cast."""
    position = np.arange(length)
    num_timescales = channels // 2
    inc = math.log(float(max_timescale) / float(min_timescale)) / (float(num_timescales) - 1)
    inv = min_timescale * np.exp(np.arange(num_timescales).astype(np.float64) * -inc)
    scaled = position[:, None] * inv[None, :]
    signal = np.concatenate([np.sin(scaled), np.cos(scaled)], axis=1)
    signal = np.pad(signal, [[0, 0], [0, channels % 2]], "constant")
    return torch.from_numpy(signal).float()


class _nvtx:
    """`with _nvtx("name"):` — an NVTX range when B200IQ_NVTX=1, else nothing"""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if NVTX:
            torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *a):
        if NVTX:
            torch.cuda.nvtx.range_pop()


def _ids(t, device):
    if t.dtype != torch.int32:
        t = t.to(torch.int32)
    return t.to(device=device, non_blocking=True).contiguous()


class _IQFunction(torch.autograd.Function):
    """autograd glue for the API-exact forward(): one node whose backward is our hand-written schedule"""

    @staticmethod
    def forward(ctx, flat, model, feats, cids, rids, tids, eps):
        outs = model._forward_api(feats, cids, rids, tids, eps)
        ctx.model = model
        ctx.latent = model.latent_transformer
        return outs

    @staticmethod
    def backward(ctx, d_output, d_z_logit, d_kld, d_img, d_recon):
        g = ctx.model._backward_api(d_output, d_z_logit, d_kld, d_img, d_recon)
        return g, None, None, None, None, None, None


class IQ(nn.Module):
    def __init__(self, latent_transformer, vocab, args, num_att_layers=2):
        super().__init__()
        assert num_att_layers == 2, "image_reconstructor is the 2-layer MLP of models/iq.py:46-48"
        self.vocab = vocab
        self.vocab_size = len(vocab.word2idx)
        self.latent_transformer = latent_transformer
        self.args = args
        self.V, self.E, self.H = self.vocab_size, args.emb_dim, args.hidden_dim
        self.L, self.F, self.NL, self.nh = args.latent_dim, args.pwffn_dim, args.num_layers, args.num_heads
        self.FEAT = getattr(args, "feat_dim", 512)  # resnet18 fc.in_features (encoder_cnn.py:20)
        self.sos_id = vocab.word2idx[vocab.SYM_SOQ]
        assert vocab.word2idx[vocab.SYM_PAD] == PAD
        # widths the kernels' 16-byte row accesses fix: the model dimensions.  The VOCABULARY size (len(vocab.word2idx): whatever the
        # frequency threshold left), the embedding width and the feature width are arbitrary: their buffers get padded row pitches
        assert self.H % self.nh == 0 and self.H % 8 == 0 and self.L % 4 == 0 and self.F % 8 == 0, \
            "hidden_dim and pwffn_dim must be multiples of 8, latent_dim a multiple of 4 (16-byte rows for TMA)"
        self.attn_dropout, self.relu_dropout = 0.1, 0.1  # transformer_layers.py:96-97,163-164 defaults (N4)
        self.arena = Arena()
        transformer_iq_specs(self.arena, self.V, self.E, self.H, self.L, self.F, self.NL, self.FEAT)
        self.flat = nn.Parameter(self.arena.finalize())
        self.arena.bind(self.flat.data)
        self.register_buffer("bn_running_mean", torch.zeros(self.H), persistent=False)
        self.register_buffer("bn_running_var", torch.ones(self.H), persistent=False)
        self.register_buffer("bn_num_batches_tracked", torch.zeros((), dtype=torch.long), persistent=False)
        self.register_buffer("timing_enc", gen_timing_signal(1000, self.H), persistent=False)  # :96,118
        self.register_buffer("timing_dec", gen_timing_signal(200, self.H), persistent=False)   # :163,184
        self.feature_cache = None  # set by attach_feature_cache(): `images` may then be (B,) rows of batch.FeatureCache
        self.trunk = None  # set by attach_trunk(); reference: torchvision resnet18 (encoder_cnn.py:17)
        self.seed = 1234
        self.seed_rank = 0   # data-parallel rank (parallel.GradReducer.begin_step): ranks draw different masks / eps like DDP replicas do
        self.step_idx = 0    # stochastic forward passes so far: every training-mode forward / decode draws fresh dropout masks and eps
        # device-resident per-step scalars (train.TrainStep sets them when it captures the step in a CUDA graph: a captured launch
        # keeps its arguments, so what changes every step must be read from memory): seed_dev = int64 word added to every Philox
        # seed (then it, not step_idx, counts the passes), kld_coef_dev = float holding kl_ceiling * kl_weight * loss_scale
        self.seed_dev = None
        self.kld_coef_dev = None
        self._ws = {}
        self._gflat = None
        self._terms = None
        self._wops = None   # transformer.WeightOperands: fp16 plane arenas (hi, lo) of the weights, or the masters in exact mode
        self._wops_exact = None
        # loss scale of the fp16-plane path: every gradient seed is multiplied by it so that activation gradients (~1e-6) sit in
        # the fp16 normal range as single-plane GEMM operands; folded back out in Adam's grad_scale (train.py) / at the end of the
        # API backward.  A power of two: exact.
        # Default (loss_scale = 0): 16 x 2^floor(log2(B*T/2)) — the largest seed, (softmax - onehot) / N_tokens, then sits near 2^4
        # whatever the batch (cfg3: 65536), with 2^12 of headroom below the fp16 maximum.
        self.loss_scale = float(os.environ.get("B200IQ_LOSS_SCALE", "0"))
        self._ctx = None
        self.on_grad_ready = None  # set by parallel.GradReducer.begin_step
        self.on_grad_range = None  # (lo, hi) pieces that become final out of arena order (the side-stream encoder)

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
        """reference key names (models/iq.py attribute tree), duplicates for shared sub-modules included"""
        sd = export_state_dict(self.arena, self._buffers_map(), TRANSFORMER_ALIASES, prefix)
        out = destination if destination is not None else type(sd)()
        for k, v in sd.items():
            out[k] = v if keep_vars else v.detach()
        return out

    def load_state_dict(self, state_dict, strict=True):
        sd = {k: v for k, v in state_dict.items() if not k.startswith("encoder_cnn.cnn.") or k.startswith("encoder_cnn.cnn.fc.")}
        missing, unexpected = import_state_dict(self.arena, self._buffers_map(), TRANSFORMER_ALIASES, sd, strict)
        return missing, unexpected

    def switch_GVT_train_mode(self, new_mode):
        """models/iq.py:51-54"""
        self.latent_transformer = new_mode

    def attach_trunk(self, trunk):
        """frozen CNN trunk mapping images -> pooled (B, FEAT) features (reference: resnet18 minus fc)"""
        self.trunk = trunk
        for p in trunk.parameters():
            p.requires_grad = False

    def attach_feature_cache(self, cache):
        """batch.FeatureCache of pooled trunk features: `images` may then be a (B,) integer tensor of cache rows"""
        self.feature_cache = cache

    # ------------------------------------------------------------------ internals
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

    def _workspace(self, key):
        ws = self._ws.get(key)
        if ws is None:
            ws = tf.Workspace(self.flat.device)
            self._ws[key] = ws
        return ws

    def _features(self, images):
        if images.dim() == 1 and not images.is_floating_point():
            if self.feature_cache is None:
                raise _lib.B200iqError("image rows given but no feature cache attached (call attach_feature_cache)")
            rows = _ids(images, self.flat.device)
            ws = self._workspace("feats")
            return self.feature_cache.gather(rows, ws.buf("feats", rows.numel(), self.feature_cache.rows.shape[1]))
        if images.dim() == 2:
            return images.to(self.flat.device, torch.float32, non_blocking=True).contiguous()
        if self.trunk is None:
            raise _lib.B200iqError("raw images given but no trunk attached (call attach_trunk, or pass pooled features)")
        with torch.no_grad():
            return self.trunk(images.to(self.flat.device)).flatten(1).float().contiguous()

    def _range_marks(self, prefix):
        """Ctx.on_ready callback for a stack that runs beside the main backward: turns its top-down layer marks into
        (lo, hi) pieces of the stack's own arena region"""
        offs = self.arena.offsets
        ends = [offs[n] + self.arena.views[n.replace("#pack", "")].numel() for n in offs if n.startswith(prefix)]
        state = {"hi": max(ends)}

        def mark(lo):
            if lo < state["hi"]:
                self.on_grad_range(lo, state["hi"])
                state["hi"] = lo
        return mark

    def _fork(self, enable):
        """side stream that starts where the current stream is now (None when disabled)"""
        if not enable:
            return None
        if getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream(device=self.flat.device)
        self._side.wait_stream(torch.cuda.current_stream())
        return self._side

    def _join(self, side):
        if side is not None:
            torch.cuda.current_stream().wait_stream(side)

    def _dec_parts(self, B, T):
        """number of sample ranges the decoder stack is cut into (DEC_SPLIT)"""
        n = DEC_SPLIT
        if n > 1 and B % n == 0 and (B // n) * T >= DEC_SPLIT_MIN_ROWS:
            return n
        return 1

    def _run_parts(self, c, nparts, fn):
        """fn(part, nparts) for every part: part 0 on the current stream, the others on their own streams (same priority) that
        fork here and are joined before this returns; c.ws is the part's row slice of the workspace while fn issues its launches"""
        main = torch.cuda.current_stream()
        ws = c.ws
        streams = [main]
        for i in range(1, nparts):
            key = ("part_stream", main.cuda_stream, i)
            st = ws.meta.get(key)
            if st is None:
                st = ws.meta[key] = torch.cuda.Stream(device=ws.device, priority=main.priority)
            st.wait_stream(main)
            streams.append(st)
        try:
            for i, st in enumerate(streams):
                key = ("part_ws", i, nparts)
                part = ws.meta.get(key)
                if part is None:
                    part = ws.meta[key] = tf.WorkspacePart(ws, i, nparts)
                c.ws = part
                if i == 0:
                    fn(i, nparts)
                else:
                    with torch.cuda.stream(st), _lib.pinned_stream():
                        fn(i, nparts)
        finally:
            c.ws = ws
        for st in streams[1:]:
            main.wait_stream(st)

    def scale(self, tokens=None):
        """the loss scale in force: 1 in exact mode; else self.loss_scale, or (loss_scale = 0) the power of two chosen from the
        number of target tokens of the batch.  Without `tokens`: the scale of the last pass."""
        if ops.exact_mode():
            return 1.0
        if self.loss_scale > 0:
            return self.loss_scale
        if tokens is None:
            return getattr(self, "_grad_scale", 1.0)
        return 16.0 * 2.0 ** math.floor(math.log2(max(1.0, tokens / 2.0)))

    def _weights(self):
        """GEMM weight operands for this step: exact mode = the fp32 masters; else the fp16 planes (hi, lo) of the arena, refreshed
        from the masters (one streaming pass: 4 B read + 4 B written per parameter, like the TF32 copy it replaces)"""
        if ops.exact_mode():
            if self._wops_exact is None:
                self._wops_exact = tf.WeightOperands(self.arena)
            return self._wops_exact
        if self._wops is None:
            self._wops = tf.build_weight_operands(self.arena, self.flat.device, skip=("embedding.0.weight",))
        tf.refresh_weight_operands(self._wops, self.arena, self.flat.data)
        return self._wops

    def _next_seed(self):
        """Philox key of the next stochastic pass.  The reference draws fresh dropout masks and a fresh torch.randn eps on every
        call (transformer_layers.py:45,406,520); a counter-based generator needs a new key per call to do the same, so the
        module counts its own stochastic passes (checkpoint.py saves / restores the counter)."""
        if self.seed_dev is not None:
            return self.seed + 1000003 * self.seed_rank   # + *seed_dev, added by the kernels
        self.step_idx += 1
        return self.seed + self.step_idx + 1000003 * self.seed_rank

    def _make_ctx(self, B, S, R, T, training):
        _lib.require_device()
        key = (B, S, R, T, bool(self.latent_transformer), bool(training), ops.exact_mode())
        ws = self._workspace(key)
        c = tf.Ctx(self.arena, self.grad_arena, ws, self.H, self.nh, self.F, self.E, self.attn_dropout,
                   self.relu_dropout, training, self._next_seed(), self._weights())
        c.on_ready = self.on_grad_ready
        return c

    def _run_forward(self, c, feats, cids, rids, tids, eps):
        """models/iq.py:101-112 down to the decoder's final LayerNorm; everything stays in the workspace.
        GEMM operands travel as ops.Act (fp16 planes hi + lo written by their producers; fp32 in exact mode)."""
        ws, P, H, L = c.ws, c.P, self.H, self.L
        B, S, T = feats.shape[0], cids.shape[1], tids.shape[1]
        latent = self.latent_transformer
        m = ws.meta
        # image head: fc + BatchNorm1d (encoder_cnn.py:33-34)
        feats_a = ops.to_act(ws, "feats.a", feats)
        img_pre = ws.buf("img_pre", B, H)
        # (split-K: K = FEAT = 2048 is the longest forward reduction, and the tensor core's fp32 accumulation error grows with K)
        ops.gemm2(feats_a, c.W["encoder_cnn.cnn.fc.weight"], img_pre, bias=P["encoder_cnn.cnn.fc.bias"], auto_split=True)
        img = ws.buf("img", B, H)
        ops.bn_fwd(img_pre, P["encoder_cnn.bn.weight"], P["encoder_cnn.bn.bias"], img, ws.buf("bn.mean", H),
                   ws.buf("bn.rstd", H), self.bn_running_mean, self.bn_running_var, c.training)
        # context encoder (encoder_transformer.py:27-32)
        def context_encoder():
            x = tf.embed_fwd(c, "enc.emb", cids, B, S, self.timing_enc)
            eo = tf.encoder_fwd(c, "enc", "answer_encoder.encoder.", x, cids, B, S, self.NL, out_kind="f32")
            ops.add_timing(eo, None, img, None, B, S, H)  # enc_out[:,0] += image_features
            return eo, ops.to_act(ws, "enc_out.a", eo)    # the cross-attention K|V projections read it as an operand
        z = None
        # (the two stacks share per-row-count scratch buffers in the backward, so they must differ in length)
        side = self._fork(latent and CONCURRENT_ENCODERS and rids.shape[1] != S)
        if side is None:
            enc_out, enc_a = context_encoder()
        else:
            with torch.cuda.stream(side), _lib.pinned_stream():
                enc_out, enc_a = context_encoder()
        if latent:
            R = rids.shape[1]
            xr = tf.embed_fwd(c, "renc.emb", rids, B, R, self.timing_enc)
            r_out = tf.encoder_fwd(c, "renc", "answer_encoder.r_encoder.", xr, rids, B, R, self.NL, out_kind="f32")
            self._join(side)
            cat = ws.buf("lat.cat", B, 2 * H)  # cat(x_p, x), transformer_layers.py:52
            ops.tok0_gather(r_out, cat[:, :H], B, R, H)
            ops.tok0_gather(enc_out, cat[:, H:], B, S, H)
            cat_a = ops.to_act(ws, "lat.cat.a", cat)
            pp, pq = "latent_layer.mean_logvar_prior.", "latent_layer.mean_logvar_posterior."
            p1 = tf.linear_fwd(c, "lat.p1", pp + "0", cat_a.cols(H, 2 * H), relu=True, planes=True)
            p2 = tf.linear_fwd(c, "lat.p2", pp + "3", p1, relu=True, planes=True)
            prior = tf.linear_fwd(c, "lat.prior", pp + "6", p2)
            q1 = tf.linear_fwd(c, "lat.q1", pq + "0", cat_a, relu=True, planes=True)
            q2 = tf.linear_fwd(c, "lat.q2", pq + "3", q1, relu=True, planes=True)
            post = tf.linear_fwd(c, "lat.post", pq + "6", q2)
            e = ws.buf("lat.eps", B, L)
            if eps is not None:
                e.copy_(eps)
            else:
                ops.randn_(e, c.seed, 7 << 40)
            zl = ws.buf("lat.z", B, L)
            ops.latent_fwd(prior, post, e, zl, self.terms[2:3])
            zl_a = ops.to_act(ws, "lat.z.a", zl)
            z = tf.linear_fwd(c, "z", "latent_projection", zl_a)
            zimg = ops.axpby(z, img, ws.buf("zimg", B, H))
            zimg_a = ops.to_act(ws, "zimg.a", zimg, lo=False)
            m["z_logit"] = tf.linear_fwd(c, "z_logit", "decoder.z_classifier", zimg_a, terms=1)
            m.update(cat_a=cat_a, zl_a=zl_a, zimg_a=zimg_a, p1=p1, p2=p2, q1=q1, q2=q2)
        else:
            self._join(side)
        # decoder (decoder_transformer.py:22-38)
        shifted = ws.buf("dec.ids", B, T, dtype=torch.int32)
        shifted[:, 0] = self.sos_id
        shifted[:, 1:] = tids[:, :-1]
        xd = tf.embed_fwd(c, "dec.emb", shifted, B, T, self.timing_dec, add0=img, add1=z)
        nparts = self._dec_parts(B, T)
        if nparts == 1:
            dec_out = tf.decoder_fwd(c, "dec", "decoder.decoder.", xd, shifted, enc_a, cids, B, T, S, self.NL)
        else:
            def part_fwd(i, n):
                b0, b1 = i * (B // n), (i + 1) * (B // n)
                tf.decoder_fwd(c, "dec", "decoder.decoder.", xd[b0 * T:b1 * T], shifted[b0:b1], enc_a.rows(b0 * S, b1 * S),
                               cids[b0:b1], b1 - b0, T, S, self.NL, pdl=False)
            self._run_parts(c, nparts, part_fwd)
            dec_out = ws.act("dec.lnf.y")   # the parts wrote their rows of the full-size buffers
        # image reconstruction (iq.py:109-112)
        rin = ws.buf("rin", B, H)
        ops.tok0_gather(enc_out, rin, B, S, H)
        if latent:
            ops.axpby(rin, z, rin)
        rin_a = ops.to_act(ws, "rin.a", rin)
        rh = tf.linear_fwd(c, "rec.h", "image_reconstructor.layers.fc0", rin_a, relu=True, planes=True)
        recon = tf.linear_fwd(c, "rec.out", "image_reconstructor.layers.fc1", rh)
        m.update(img=img, img_pre=img_pre, enc_out=enc_out, enc_a=enc_a, z=z, dec_out=dec_out, shifted=shifted, rin_a=rin_a, rh=rh,
                 recon=recon, feats=feats, feats_a=feats_a, cids=cids, rids=rids, tids=tids)
        return dec_out

    def _backward_tail(self, c, d_dec, d_img, d_z, d_recon, kld_coef):
        """everything below the loss seeds: recon MLP, decoder, latent, both encoders, embedder, image head.
        d_img / d_z are fp32 accumulators already holding the seeds from the losses / z_classifier.  Every gradient carries
        the loss scale (self.scale()); backward GEMM operands are single fp16 planes."""
        ws, P, G, H, L = c.ws, c.P, c.G, self.H, self.L
        m = ws.meta
        latent = self.latent_transformer
        B, S, T = m["feats"].shape[0], m["cids"].shape[1], m["tids"].shape[1]
        d_recon_a = ops.to_act(ws, "d_recon.a", d_recon, lo=False)
        d_rh = tf.linear_bwd(c, "d_rh", "image_reconstructor.layers.fc1", d_recon_a, m["rh"], relu_mask=m["rh"], planes=True)
        d_rin = tf.linear_bwd(c, "d_rin", "image_reconstructor.layers.fc0", d_rh, m["rin_a"])
        c.ready("decoder.z_classifier.weight")  # vocab projections + reconstructor: final, largest bucket goes first
        if latent:
            ops.axpby(d_z, d_rin, d_z)
        d_enc_out = ws.buf("d_enc_out", B * S, H)
        nparts = self._dec_parts(B, T)
        if nparts == 1:
            dxd, dxd_a = tf.decoder_bwd(c, "dec", "decoder.decoder.", d_dec, m["shifted"], m["enc_a"], d_enc_out, m["cids"],
                                        B, T, S, self.NL)
        else:
            # the parts add into the same weight gradients (atomic reduce-adds) in no fixed order: a layer's gradients are final
            # when every part has passed it, so the bucket marks of the stack are replaced by one mark after the join
            saved, c.on_ready = c.on_ready, None

            def part_bwd(i, n):
                b0, b1 = i * (B // n), (i + 1) * (B // n)
                tf.decoder_bwd(c, "dec", "decoder.decoder.", d_dec[b0 * T:b1 * T], m["shifted"][b0:b1], m["enc_a"].rows(b0 * S, b1 * S),
                               d_enc_out[b0 * S:b1 * S], m["cids"][b0:b1], b1 - b0, T, S, self.NL)
            try:
                self._run_parts(c, nparts, part_bwd)
            finally:
                c.on_ready = saved
            c.ready("decoder.decoder.dec.0.layer_norm_mha_dec.weight")
            dxd = ws["dec.dx"]
            dxd_a = tf.dx_operand(c, "dec.dx.hi", dxd)
        ops.tok0_gather(dxd, d_img, B, T, H, accumulate=True)
        if latent:
            ops.tok0_gather(dxd, d_z, B, T, H, accumulate=True)
        tf.embed_bwd(c, "dec.emb", m["shifted"], dxd_a)
        ops.tok0_scatter_add(d_rin, d_enc_out, B, S, H)
        if latent:
            R = m["rids"].shape[1]
            pp, pq = "latent_layer.mean_logvar_prior.", "latent_layer.mean_logvar_posterior."
            d_z_a = ops.to_act(ws, "d_z.a", d_z, lo=False)
            d_zl = tf.linear_bwd(c, "d_zl", "latent_projection", d_z_a, m["zl_a"])
            dprior, dpost = ws.buf("d_prior", B, 2 * L), ws.buf("d_post", B, 2 * L)
            ops.latent_bwd(ws["lat.prior"], ws["lat.post"], ws["lat.eps"], d_zl, kld_coef, dprior, dpost)
            cat_a = m["cat_a"]
            d_cat = ws.buf("d_cat", B, 2 * H)
            dpost_a = ops.to_act(ws, "d_post.a", dpost, lo=False)
            d_q2 = tf.linear_bwd(c, "d_q2", pq + "6", dpost_a, m["q2"], relu_mask=m["q2"], planes=True)
            d_q1 = tf.linear_bwd(c, "d_q1", pq + "3", d_q2, m["q1"], relu_mask=m["q1"], planes=True)
            tf.linear_bwd(c, None, pq + "0", d_q1, cat_a, dx=d_cat)
            dprior_a = ops.to_act(ws, "d_prior.a", dprior, lo=False)
            d_p2 = tf.linear_bwd(c, "d_p2", pp + "6", dprior_a, m["p2"], relu_mask=m["p2"], planes=True)
            d_p1 = tf.linear_bwd(c, "d_p1", pp + "3", d_p2, m["p1"], relu_mask=m["p1"], planes=True)
            tf.linear_bwd(c, None, pp + "0", d_p1, cat_a.cols(H, 2 * H), dx=d_cat[:, H:], accumulate_dx=True)
            ops.tok0_scatter_add(d_cat[:, H:], d_enc_out, B, S, H)
            c.ready("latent_layer.mean_logvar_prior.0.weight")

        def context_encoder_bwd():
            ops.tok0_gather(d_enc_out, d_img, B, S, H, accumulate=True)  # enc_out[:,0] = LN(..)[:,0] + image_features
            de_in = tf.encoder_bwd(c, "enc", "answer_encoder.encoder.", d_enc_out, m["cids"], B, S, self.NL)
            tf.embed_bwd(c, "enc.emb", m["cids"], de_in)
        side = self._fork(latent and CONCURRENT_ENCODERS and m["rids"].shape[1] != S)
        if side is not None:
            # both encoder backward passes at once; gradient-bucket marks wait for the join (the two stacks finish in no fixed
            # order, and a mark promises that everything above an arena offset is final)
            # order: the main stream keeps its suffix marks (the response encoder lies ABOVE the context encoder in the arena,
            # so "everything above this offset is final" stays true for it); the context encoder reports its layers as
            # ranges from the side stream
            saved = c.on_ready
            c.on_ready = self._range_marks("answer_encoder.encoder.") if (saved is not None and self.on_grad_range is not None) else None
            with torch.cuda.stream(side), _lib.pinned_stream():
                context_encoder_bwd()
            c.on_ready = saved
        if latent:
            d_r_out = ws.buf("d_r_out", B * R, H)
            d_r_out.zero_()
            ops.tok0_scatter_add(d_cat[:, :H], d_r_out, B, R, H)
            dr_in = tf.encoder_bwd(c, "renc", "answer_encoder.r_encoder.", d_r_out, m["rids"], B, R, self.NL)
            tf.embed_bwd(c, "renc.emb", m["rids"], dr_in)
        if side is not None:
            self._join(side)
            c.ready("answer_encoder.encoder.enc.0.layer_norm_mha.weight")
        else:
            context_encoder_bwd()
        d_img_pre = ws.buf("d_img_pre", B, H)
        ops.bn_bwd(d_img, m["img_pre"], P["encoder_cnn.bn.weight"], ws["bn.mean"], ws["bn.rstd"], d_img_pre,
                   G["encoder_cnn.bn.weight"], G["encoder_cnn.bn.bias"])
        d_img_pre_a = ops.to_act(ws, "d_img_pre.a", d_img_pre, lo=False)
        tf.grad_w(c, d_img_pre_a, m["feats_a"], G["encoder_cnn.cnn.fc.weight"], G["encoder_cnn.cnn.fc.bias"])

    # ------------------------------------------------------------------ fused training path
    def loss_coefs(self, kliter=0):
        """train_iq.py:88-101 weights: (rec, img, kld, aux)"""
        a = self.args
        if not self.latent_transformer:
            return 1.0, a.image_recon_lambda, 0.0, 0.0
        kl_w = min(math.tanh(6 * kliter / a.full_kl_step - 3) + 1, 1)
        return 1.0, a.image_recon_lambda, a.kl_ceiling * kl_w, a.aux_ceiling

    def _ce_chunk_rows(self, M):
        rows = (CE_CHUNK_MB << 20) // (4 * self.V)  # keep a logits chunk (fp32) L2-resident on B200 (126 MB L2)
        rows = max(128, rows // 128 * 128)
        return min(M, rows)

    def forward_losses(self, images, answers, response, target, kliter=0, eps=None, backward=True):
        """Forward + losses (train_iq.py:81-103) + full backward, logits never materialised.
        Returns the device tensor of loss terms (rec, img, kld, aux, count, ...); gradients are ACCUMULATED into
        `self.grad_arena` MULTIPLIED BY self.scale() (the loss scale of the fp16-plane path; train.TrainStep folds it into Adam's
        grad_scale, `unscaled_grads()` gives a plain copy).  total = rec + c_img*img + c_kld*kld + c_aux*aux (loss_coefs / total_loss)."""
        prev, ops.SEED_DEV = ops.SEED_DEV, self.seed_dev
        try:
            with ops.no_rounding():
                return self._forward_losses(images, answers, response, target, kliter, eps, backward)
        finally:
            ops.SEED_DEV = prev

    def unscaled_grads(self):
        """the flat gradient arena with the loss scale of the pass that filled it divided out (tests / inspection)"""
        return self.grad_arena / getattr(self, "_grad_scale", 1.0)

    def _forward_losses(self, images, answers, response, target, kliter, eps, backward):
        dev = self.flat.device
        feats = self._features(images)
        cids, tids = _ids(answers, dev), _ids(target, dev)
        rids = _ids(response, dev) if self.latent_transformer else None
        B, S, T = feats.shape[0], cids.shape[1], tids.shape[1]
        R = rids.shape[1] if rids is not None else 0
        c = self._make_ctx(B, S, R, T, self.training)
        self._ctx = c
        ws, P, G, H, V = c.ws, c.P, c.G, self.H, self.V
        terms = self.terms
        terms.zero_()
        exact = ops.exact_mode()
        SC = self._grad_scale = self.scale(B * T)
        pdl_prev = ops.set_pdl(True) if PDL_FORWARD else None
        with _nvtx("iq.forward"):
            dec_out = self._run_forward(c, feats, cids, rids, tids, eps)
        m = ws.meta
        c_rec, c_img, c_kld, c_aux = self.loss_coefs(kliter)
        M = B * T
        count = terms[4:5]
        ops.count_nonpad(tids, PAD, count)
        tflat = tids.view(-1)
        d_dec = ws.buf("d_dec_out", M, H)
        chunk = self._ce_chunk_rows(M)
        Vp = ops.pad8(V)   # row pitch of the chunk buffers: any vocabulary size keeps 16-byte rows
        logits = ws.buf("logits_chunk", chunk, Vp)[:, :V]
        # the chunk's gradient coef * (softmax - onehot) / N: an fp16 plane beside the fp32 logits (exact mode: in place)
        lgrad = ops.Act(f32=logits) if exact else ops.Act(hi=ws.buf("logits_grad", chunk, Vp, dtype=torch.float16)[:, :V])
        wout, bout = c.W["decoder.output.weight"], P["decoder.output.bias"]
        if NVTX:
            torch.cuda.nvtx.range_push("iq.vocab_ce")
        for r0 in range(0, M, chunk):
            rows = min(chunk, M - r0)
            lg, xs, gr = logits[:rows], dec_out.rows(r0, r0 + rows), lgrad.rows(0, rows)
            ops.gemm2(xs, wout, lg, bias=bout, terms=1)       # decoder_transformer.py:40
            ops.ce_rows_act(lg, tflat[r0:r0 + rows], PAD, count, c_rec * SC, terms[0:1], gr if backward else None)
            if backward:
                # dX on the critical path, the vocabulary weight gradient (+ bias column sum) on the side stream under it;
                # joined before the next chunk's gradient overwrites gr
                ev = c.mark()
                ops.gemm2(gr, wout, d_dec[r0:r0 + rows], b_t=True, auto_split=True, terms=1)
                c.on_side(lambda: tf.grad_w(c, gr, xs, G["decoder.output.weight"], G["decoder.output.bias"]), ev)
                c.join_side()
        if NVTX:
            torch.cuda.nvtx.range_pop()
        if pdl_prev is not None:
            ops.set_pdl(pdl_prev)
        d_img = ws.buf("d_img", B, H)
        d_z = None
        if self.latent_transformer:
            dzl = ws.buf("d_z_logit", B, V)
            ops.ce_aux(m["z_logit"], dzl, tids, PAD, count, c_aux * SC, terms[3:4])
            if backward:
                dzl_a = ops.to_act(ws, "d_z_logit.a", dzl, lo=False)
                d_z = tf.linear_bwd(c, "d_zimg", "decoder.z_classifier", dzl_a, m["zimg_a"])
                ops.axpby(d_z, d_z, d_img, 1.0, 0.0)
        elif backward:
            d_img.zero_()
        d_recon = ws.buf("d_recon", B, H)
        ops.mse(m["img"], m["recon"], c_img * SC, terms[1:2], d_img if backward else None, d_recon if backward else None)
        if backward:
            with _nvtx("iq.backward"):
                self._backward_tail(c, d_dec, d_img, d_z, d_recon, self.kld_coef_dev if self.kld_coef_dev is not None else c_kld * SC)
        return terms

    def total_loss(self, terms, kliter=0):
        c_rec, c_img, c_kld, c_aux = self.loss_coefs(kliter)
        return c_rec * terms[0] + c_img * terms[1] + c_kld * terms[2] + c_aux * terms[3]

    # ------------------------------------------------------------------ API-exact path (materialises logits)
    def forward(self, images, answers, response, target, eps=None):
        """models/iq.py:82-114. Returns (output (B,T,V), z_logit (B,V)|None, kld|None, (image_features, recon))."""
        dev = self.flat.device
        feats = self._features(images)
        cids, tids = _ids(answers, dev), _ids(target, dev)
        rids = _ids(response, dev) if self.latent_transformer else None
        if torch.is_grad_enabled() and self.flat.requires_grad:
            output, z_logit, kld, img, recon = _IQFunction.apply(self.flat, self, feats, cids, rids, tids, eps)
        else:
            output, z_logit, kld, img, recon = self._forward_api(feats, cids, rids, tids, eps)
        if not self.latent_transformer:
            z_logit, kld = None, None
        return output, z_logit, kld, (img, recon)

    def _forward_api(self, feats, cids, rids, tids, eps):
        with ops.no_rounding():
            B, S, T = feats.shape[0], cids.shape[1], tids.shape[1]
            R = rids.shape[1] if rids is not None else 0
            c = self._make_ctx(B, S, R, T, self.training)
            self._ctx = c
            self.terms.zero_()
            dec_out = self._run_forward(c, feats, cids, rids, tids, eps)
            m = c.ws.meta
            Vp = ops.pad8(self.V)
            output = torch.empty(B * T, Vp, device=feats.device)[:, :self.V]   # padded pitch (a view when V % 8 != 0)
            ops.gemm2(dec_out, c.W["decoder.output.weight"], output, bias=c.P["decoder.output.bias"], terms=1)
            if self.latent_transformer:
                z_logit, kld = m["z_logit"].clone(), self.terms[2].clone()
            else:
                z_logit, kld = torch.zeros(1, device=feats.device), torch.zeros((), device=feats.device)
            return output.unflatten(0, (B, T)), z_logit, kld, m["img"].clone(), m["recon"].clone()

    def _backward_api(self, d_output, d_z_logit, d_kld, d_img_in, d_recon_in):
        with ops.no_rounding():
            return self._backward_api_body(d_output, d_z_logit, d_kld, d_img_in, d_recon_in)

    def _backward_api_body(self, d_output, d_z_logit, d_kld, d_img_in, d_recon_in):
        """autograd's incoming gradients are multiplied by the loss scale as they become operands; the result is divided by it"""
        c = self._ctx
        ws, m, H = c.ws, c.ws.meta, self.H
        B, T = m["tids"].shape
        SC = self.scale(B * T)
        g = torch.zeros_like(self.flat.data)
        c.gflat = g
        c.G = self.arena.views_over(g, cache=False)
        d_dec = ws.buf("d_dec_out", B * T, H)
        if d_output is not None:
            dout = d_output.reshape(B * T, self.V).contiguous()
            if ops.exact_mode():
                dout_a = ops.Act(f32=dout)
            else:
                dout_a = ops.Act(hi=torch.empty(B * T, ops.pad8(self.V), device=dout.device, dtype=torch.float16)[:, :self.V])
                ops.split_f16(dout, dout_a.hi, None, SC)
            ops.gemm2(dout_a, c.W["decoder.output.weight"], d_dec, b_t=True, auto_split=True, terms=1)
            tf.grad_w(c, dout_a, m["dec_out"], c.G["decoder.output.weight"], c.G["decoder.output.bias"])
        else:
            d_dec.zero_()
        d_img = ws.buf("d_img", B, H)
        if d_img_in is not None:
            ops.axpby(d_img_in.contiguous(), d_img_in.contiguous(), d_img, SC, 0.0)
        else:
            d_img.zero_()
        d_z = None
        if self.latent_transformer:
            d_z = ws.buf("d_zimg", B, H)
            if d_z_logit is not None:
                dzl_a = ops.to_act(ws, "d_z_logit.a", d_z_logit.contiguous(), lo=False, scale=SC)
                tf.linear_bwd(c, "d_zimg", "decoder.z_classifier", dzl_a, m["zimg_a"], dx=d_z)
                ops.axpby(d_img, d_z, d_img)
            else:
                d_z.zero_()
        d_recon = ws.buf("d_recon", B, H)
        if d_recon_in is not None:
            ops.axpby(d_recon_in.contiguous(), d_recon_in.contiguous(), d_recon, SC, 0.0)
        else:
            d_recon.zero_()
        kld_coef = float(d_kld) * SC if (d_kld is not None and self.latent_transformer) else 0.0
        self._backward_tail(c, d_dec, d_img, d_z, d_recon, kld_coef)
        c.gflat, c.G = self.grad_arena, None
        if SC != 1.0:
            g.mul_(1.0 / SC)
        return g

    # ------------------------------------------------------------------ greedy decoding (models/iq.py:117-152)
    @torch.no_grad()
    def decode_greedy(self, images, answers, max_decode_length=50, eps=None, exact=True, kv_cache=True):
        """models/iq.py:117-152: encoder once, prior-only latent (a fresh eps: Philox here, or injected), then
        max_decode_length+1 decoder passes over the growing prefix that starts with PAD (so key 0 is pad-masked and the first
        query row is fully masked -> uniform attention), argmax + top-6 softmax per step.
        Runs in EXACT fp32 (CUDA-core GEMMs, no TF32 rounding) so the token ids are bit-identical to an fp32 run.
        kv_cache=True (default) feeds only the newest token through the decoder and keeps every layer's self-attention K/V
        (and the cross-attention K/V of the 5-token context) in a cache: the decoder is causal and pre-LN, so earlier rows
        never change and the last row comes out bit-identical to the reference's O(T^2) recomputation (kv_cache=False).
        Returns (sentences, top_args (B, L+1, 6) float32, top_vals (B, L+1, 6) float32) like the reference."""
        with ops.gemm_mode(simt=exact):
            return self._decode_greedy(images, answers, max_decode_length, eps, kv_cache)

    def _decode_steps_cached(self, c, ys, top_i, top_v, enc_out, enc_a, cids, img, z, B, S, steps):
        """inference_forward (decoder_transformer.py:43-48) for the newest token only, with per-layer K/V caches"""
        ws, P, W, H, V, NL, nh, d = c.ws, c.P, c.W, self.H, self.V, self.NL, self.nh, self.H // self.nh
        pfx = "decoder.decoder."
        A = lambda name, t, lo=True: ops.to_act(ws, name, t, lo=lo)   # fp32 buffer -> GEMM operand (free in exact mode)
        kc = [ws.buf("kv.k%d" % i, B, steps, H) for i in range(NL)]
        vc = [ws.buf("kv.v%d" % i, B, steps, H) for i in range(NL)]
        ckv = []
        for i in range(NL):   # cross-attention K | V of the context: once per layer instead of once per step
            lp = "%sdec.%d.multi_head_attention_enc_dec." % (pfx, i)
            ckv.append(ops.gemm2(enc_a, c.pspan(lp + "key_linear.weight", lp + "value_linear.weight", H), ws.buf("kv.c%d" % i, B * S, 2 * H)))
        e1, x = c.act("kv.e", B, self.E), ws.buf("kv.x", B, H)
        xn, qkv, ctx = c.act("kv.xn", B, H), ws.buf("kv.qkv", B, 3 * H), c.act("kv.ctx", B, H)
        a, q, b_, h, f2 = ws.buf("kv.a", B, H), ws.buf("kv.q", B, H), ws.buf("kv.b", B, H), c.act("kv.h", B, self.F), ws.buf("kv.f2", B, H)
        logit = ws.buf("g.logit", B, V)
        ti, tv = ws.buf("g.ti", B, 6, dtype=torch.int32), ws.buf("g.tv", B, 6)

        def ln(name, src, dst):
            ops.layernorm_fwd_act(src, P[name + ".weight"], P[name + ".bias"], None, None, act=dst)
        for p in range(steps):
            n = p + 1
            tok = ys[:, p].contiguous()
            prefix = ys[:, :n].contiguous()
            ops.embed_gather_act(tok, P["embedding.0.weight"], e1)
            ops.gemm2(e1, W["embedding.1.weight"], x, bias=P["embedding.1.bias"])
            # timing signal of position p; token 0 also gets z + image_features (decoder_transformer.py:46)
            ops.add_timing(x, self.timing_dec[p:p + 1], z if p == 0 else None, img if p == 0 else None, B, 1, H)
            for i in range(NL):
                lp = "%sdec.%d." % (pfx, i)
                sa, ca, ff = lp + "multi_head_attention_dec.", lp + "multi_head_attention_enc_dec.", lp + "positionwise_feed_forward."
                ln(lp + "layer_norm_mha_dec", x, xn)
                ops.gemm2(xn, c.pspan(sa + "query_linear.weight", sa + "value_linear.weight", H), qkv)
                kc[i][:, p].copy_(qkv[:, H:2 * H])
                vc[i][:, p].copy_(qkv[:, 2 * H:])
                # every cached key is at or before the query's position: the causal part of the mask is implied
                ops.attention_fwd(qkv[:, :H], kc[i].view(B * steps, H), vc[i].view(B * steps, H), ctx, prefix, PAD, False, B, nh, 1, n, d,
                                  kv_batch_stride=steps * H)
                ops.gemm2(ctx, W[sa + "output_linear.weight"], a, residual=x)
                ln(lp + "layer_norm_mha_enc", a, xn)
                ops.gemm2(xn, W[ca + "query_linear.weight"], q)
                ops.attention_fwd(q, ckv[i][:, :H], ckv[i][:, H:], ctx, cids, PAD, False, B, nh, 1, S, d)
                ops.gemm2(ctx, W[ca + "output_linear.weight"], b_, residual=a)
                ln(lp + "layer_norm_ffn", b_, xn)
                ops.gemm2(xn, W[ff + "layers.0.weight"], planes=h, bias=P[ff + "layers.0.bias"], relu=True)
                ops.gemm2(h, W[ff + "layers.1.weight"], f2, bias=P[ff + "layers.1.bias"], relu=True)
                ops.axpby(b_, f2, x)
            ln(pfx + "layer_norm", x, xn)
            ops.gemm2(xn, W["decoder.output.weight"], logit, bias=P["decoder.output.bias"], terms=1)
            ops.topk_softmax_rows(logit, 6, ti, tv)                           # iq.py:136-137
            ys[:, n] = ti[:, 0]
            top_i[:, p] = ti
            top_v[:, p] = tv

    def _decode_greedy(self, images, answers, max_decode_length, eps, kv_cache=True):
        with ops.no_rounding():
            return self._decode_greedy_body(images, answers, max_decode_length, eps, kv_cache)

    def _decode_greedy_body(self, images, answers, max_decode_length, eps, kv_cache=True):
        dev = self.flat.device
        was_training = self.training
        self.eval()
        wkey = None
        try:
            feats = self._features(images)
            cids = _ids(answers, dev)
            B, S, H, L, V = feats.shape[0], cids.shape[1], self.H, self.L, self.V
            steps = max_decode_length + 1
            c = self._make_ctx(B, S, 0, steps, False)
            wkey = (B, S, 0, steps, bool(self.latent_transformer), False, ops.exact_mode())
            ws, P = c.ws, c.P
            img_pre = ws.buf("img_pre", B, H)
            ops.gemm2(ops.to_act(ws, "feats.a", feats), c.W["encoder_cnn.cnn.fc.weight"], img_pre, bias=P["encoder_cnn.cnn.fc.bias"])
            img = ws.buf("img", B, H)
            ops.bn_fwd(img_pre, P["encoder_cnn.bn.weight"], P["encoder_cnn.bn.bias"], img, ws.buf("bn.mean", H), ws.buf("bn.rstd", H),
                       self.bn_running_mean, self.bn_running_var, False)
            x = tf.embed_fwd(c, "enc.emb", cids, B, S, self.timing_enc)
            enc_out = tf.encoder_fwd(c, "enc", "answer_encoder.encoder.", x, cids, B, S, self.NL, out_kind="f32")
            ops.add_timing(enc_out, None, img, None, B, S, H)
            enc_a = ops.to_act(ws, "enc_out.a", enc_out)
            z = None
            if self.latent_transformer:
                x0 = ws.buf("g.x0", B, H)
                ops.tok0_gather(enc_out, x0, B, S, H)
                pp = "latent_layer.mean_logvar_prior."
                p1 = tf.linear_fwd(c, "lat.p1", pp + "0", ops.to_act(ws, "g.x0.a", x0), relu=True, planes=True)
                p2 = tf.linear_fwd(c, "lat.p2", pp + "3", p1, relu=True, planes=True)
                prior = tf.linear_fwd(c, "lat.prior", pp + "6", p2)
                e = ws.buf("lat.eps", B, L)
                if eps is not None:
                    e.copy_(eps)
                else:
                    ops.randn_(e, c.seed, 11 << 40)
                zl = ws.buf("lat.z", B, L)
                ops.latent_fwd(prior, None, e, zl, None)                     # Latent(x, None): sample from the prior (iq.py:126)
                z = tf.linear_fwd(c, "z", "latent_projection", ops.to_act(ws, "lat.z.a", zl))
            ys = torch.full((B, steps + 1), PAD, dtype=torch.int32, device=dev)  # prefix: ys[:, 0] = PAD (iq.py:129)
            top_i = torch.zeros(B, steps, 6, dtype=torch.int32, device=dev)
            top_v = torch.zeros(B, steps, 6, device=dev)
            last = ws.buf("g.last", B, H)
            logit = ws.buf("g.logit", B, V)
            ti, tv = ws.buf("g.ti", B, 6, dtype=torch.int32), ws.buf("g.tv", B, 6)
            if kv_cache:
                self._decode_steps_cached(c, ys, top_i, top_v, enc_out, enc_a, cids, img, z, B, S, steps)
            else:
                for i in range(steps):
                    n = i + 1
                    prefix = ys[:, :n].contiguous()
                    xd = tf.embed_fwd(c, "g.emb", prefix, B, n, self.timing_dec, add0=z, add1=img)   # decoder_transformer.py:46
                    dec = tf.decoder_fwd(c, "g", "decoder.decoder.", xd, prefix, enc_a, cids, B, n, S, self.NL, out_kind="f32")
                    ops.tok0_gather(dec[n - 1:], last, B, n, H)                    # logits[:, -1]
                    ops.gemm2(ops.to_act(ws, "g.last.a", last, lo=False), c.W["decoder.output.weight"], logit, bias=P["decoder.output.bias"], terms=1)
                    ops.topk_softmax_rows(logit, 6, ti, tv)                       # iq.py:136-137
                    ys[:, n] = ti[:, 0]
                    top_i[:, i] = ti
                    top_v[:, i] = tv
            ids = ys[:, 1:].cpu()
        finally:
            self.train(was_training)
            if wkey is not None:
                self._ws.pop(wkey, None)  # decode buffers are not kept
        eos = self.vocab.word2idx[self.vocab.SYM_EOS]
        sentences = []
        for row in ids.tolist():                                              # iq.py:144-151
            st = ""
            for t in row:
                if t == eos:
                    break
                st += self.vocab.idx2word[t] + " "
            sentences.append(st)
        self.last_greedy_ids = ids.long()
        return sentences, top_i.float(), top_v
