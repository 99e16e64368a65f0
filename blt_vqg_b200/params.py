"""Flat fp32 parameter arena with reference-compatible state_dict names.

All trainable tensors of the IQ head live in ONE contiguous fp32 buffer (`arena.flat`, the only
nn.Parameter), so gradient all-reduce, global-norm clipping and Adam each touch one flat range.
`arena.views[name]` are views into it under the reference's own parameter names
(models/iq.py:32-48 attribute tree), and `state_dict()` / `load_state_dict()` speak exactly those
names — including the duplicate keys the reference produces for shared sub-modules
(`embedding`, `answer_encoder.embedding`, `decoder.embedding`; `latent_layer`,
`answer_encoder.latent_layer`; SURVEY.md §5 checkpoint row).
"""
import math
from collections import OrderedDict

import torch

ALIGN = 64  # floats (256 B): keeps every tensor TMA-aligned


def _linear_init(w, b, fan_in):
    bound = 1.0 / math.sqrt(fan_in)
    w.uniform_(-bound, bound)  # == kaiming_uniform_(a=sqrt(5)) of nn.Linear
    if b is not None:
        b.uniform_(-bound, bound)


class Arena:
    def __init__(self):
        self.specs = []  # (name, shape, init)
        self.offsets = {}
        self.flat = None
        self.views = {}

    def add(self, name, shape, init):
        self.specs.append((name, tuple(shape), init))

    def linear(self, prefix, out_f, in_f, bias=True, init="linear"):
        self.add(prefix + ".weight", (out_f, in_f), (init, in_f))
        if bias:
            self.add(prefix + ".bias", (out_f,), (init + "_bias", in_f))

    def layernorm(self, prefix, h):
        self.add(prefix + ".weight", (h,), ("ones", 0))
        self.add(prefix + ".bias", (h,), ("zeros", 0))

    def finalize(self, generator=None):
        off = 0
        for name, shape, _ in self.specs:
            self.offsets[name] = off
            n = 1
            for s in shape:
                n *= s
            # tensors that must stay contiguous for fused GEMMs (q|k|v) are sized in multiples of 4 and
            # registered back-to-back by the caller via `pack=True` names (no padding between them)
            # "#pack": the next tensor follows without padding; otherwise the next OFFSET is rounded up to ALIGN (so that a short
            # packed tensor, e.g. a [Z] bias, cannot shift everything behind it off the 16-byte grid of the fp16 plane arenas)
            off = off + n if name.endswith("#pack") else (off + n + ALIGN - 1) // ALIGN * ALIGN
        self.numel = off
        flat = torch.zeros(off, dtype=torch.float32)
        self.bind(flat)
        with torch.no_grad():
            for name, shape, (kind, fan) in self.specs:
                v = self.views[name.replace("#pack", "")]
                if kind == "linear":
                    bound = 1.0 / math.sqrt(fan)
                    v.uniform_(-bound, bound, generator=generator)
                elif kind == "linear_bias":
                    bound = 1.0 / math.sqrt(fan)
                    v.uniform_(-bound, bound, generator=generator)
                elif kind == "he":  # models/mlp.py:38
                    v.normal_(0.0, math.sqrt(2.0 / fan), generator=generator)
                elif kind == "he_bias" or kind == "zeros":
                    v.zero_()
                elif kind == "ones":
                    v.fill_(1.0)
                elif kind == "normal":  # fan carries the std
                    v.normal_(0.0, fan, generator=generator)
                elif kind == "normal_bias":
                    v.zero_()
                elif kind == "uniform":  # fan carries the bound
                    v.uniform_(-fan, fan, generator=generator)
                else:
                    raise ValueError(kind)
        return flat

    def bind(self, flat):
        """(re)create the named views over `flat` (after .to(device))"""
        self.flat = flat
        self.views = {}
        for name, shape, _ in self.specs:
            n = 1
            for s in shape:
                n *= s
            o = self.offsets[name]
            self.views[name.replace("#pack", "")] = flat.detach()[o:o + n].view(shape)

    def view_of(self, other_flat, name):
        """the same named slice inside another flat buffer (gradients, Adam moments)"""
        for nm, shape, _ in self.specs:
            if nm.replace("#pack", "") == name:
                n = 1
                for s in shape:
                    n *= s
                o = self.offsets[nm]
                return other_flat[o:o + n].view(shape)
        raise KeyError(name)

    def views_over(self, other_flat, cache=True):
        """all named slices of another flat buffer laid out like the arena (TF32 operand copy, gradients), built once
        per buffer: a training step asks for them every step"""
        store = self.__dict__.setdefault("_views_over", {})
        key = (other_flat.data_ptr(), other_flat.numel(), other_flat.dtype, str(other_flat.device))
        v = store.get(key) if cache else None
        if v is None:
            v = {}
            for nm, shape, _ in self.specs:
                n = 1
                for s in shape:
                    n *= s
                o = self.offsets[nm]
                v[nm.replace("#pack", "")] = other_flat[o:o + n].view(shape)
            if cache:   # only long-lived buffers (the cached views keep the storage alive)
                if len(store) > 8:
                    store.clear()
                store[key] = v
        return v

    def span(self, first, last):
        """flat range covering consecutive packed tensors first..last as one [rows, cols] matrix"""
        o0 = self.offsets.get(first + "#pack", self.offsets.get(first))
        v1 = self.views[last]
        o1 = self.offsets.get(last + "#pack", self.offsets.get(last)) + v1.numel()
        return o0, o1

    def names(self):
        return [n.replace("#pack", "") for n, _, _ in self.specs]


def mha_specs(a, prefix, H, fused_qkv=True):
    """q,k,v weights back-to-back so [3H,H] (self-attention) or [2H,H] (k|v, cross-attention) is one GEMM operand"""
    a.add(prefix + "query_linear.weight#pack", (H, H), ("linear", H))
    a.add(prefix + "key_linear.weight#pack", (H, H), ("linear", H))
    a.add(prefix + "value_linear.weight", (H, H), ("linear", H))
    a.linear(prefix + "output_linear", H, H, bias=False)


def transformer_iq_specs(a, V, E, H, L, F, NL, FEAT):
    """parameter inventory of models/iq.py:32-48 in forward order (reverse = gradient-ready order)"""
    a.add("embedding.0.weight", (V, E), ("normal", 0.01))            # iq.py:58,73
    a.linear("embedding.1", H, E)                                     # iq.py:76
    a.add("encoder_cnn.cnn.fc.weight", (H, FEAT), ("normal", 0.02))   # encoder_cnn.py:20,27
    a.add("encoder_cnn.cnn.fc.bias", (H,), ("zeros", 0))
    a.layernorm("encoder_cnn.bn", H)                                   # BatchNorm1d affine: ones / zeros
    for enc in ("answer_encoder.encoder.", "answer_encoder.r_encoder."):
        for i in range(NL):
            lp = "%senc.%d." % (enc, i)
            a.layernorm(lp + "layer_norm_mha", H)
            mha_specs(a, lp + "multi_head_attention.", H)
            a.layernorm(lp + "layer_norm_ffn", H)
            a.linear(lp + "positionwise_feed_forward.layers.0", F, H)
            a.linear(lp + "positionwise_feed_forward.layers.1", H, F)
        a.layernorm(enc + "layer_norm", H)
    for which, in_f in (("mean_logvar_prior", H), ("mean_logvar_posterior", 2 * H)):
        a.linear("latent_layer.%s.0" % which, 2 * L, in_f)
        a.linear("latent_layer.%s.3" % which, 2 * L, 2 * L)
        a.linear("latent_layer.%s.6" % which, 2 * L, 2 * L)
    a.linear("latent_projection", H, L)
    for i in range(NL):
        lp = "decoder.decoder.dec.%d." % i
        a.layernorm(lp + "layer_norm_mha_dec", H)
        mha_specs(a, lp + "multi_head_attention_dec.", H)
        a.layernorm(lp + "layer_norm_mha_enc", H)
        mha_specs(a, lp + "multi_head_attention_enc_dec.", H)
        a.layernorm(lp + "layer_norm_ffn", H)
        a.linear(lp + "positionwise_feed_forward.layers.0", F, H)
        a.linear(lp + "positionwise_feed_forward.layers.1", H, F)
    a.layernorm("decoder.decoder.layer_norm", H)
    a.linear("decoder.z_classifier", V, H)
    a.linear("decoder.output", V, H)
    a.add("image_reconstructor.layers.fc0.weight", (F, H), ("he", H))   # mlp.py:37-39
    a.add("image_reconstructor.layers.fc0.bias", (F,), ("zeros", 0))
    a.add("image_reconstructor.layers.fc1.weight", (H, F), ("he", F))
    a.add("image_reconstructor.layers.fc1.bias", (H,), ("zeros", 0))


# reference state_dict aliases: shared sub-modules are reachable under several attribute paths
TRANSFORMER_ALIASES = (
    ("answer_encoder.embedding.", "embedding."),
    ("decoder.embedding.", "embedding."),
    ("answer_encoder.latent_layer.", "latent_layer."),
)


def export_state_dict(arena, buffers, aliases, prefix=""):
    out = OrderedDict()
    for name in arena.names():
        out[prefix + name] = arena.views[name]
    for k, v in buffers.items():
        out[prefix + k] = v
    for alias, canon in aliases:
        for name in arena.names():
            if name.startswith(canon):
                out[prefix + alias + name[len(canon):]] = arena.views[name]
    return out


def import_state_dict(arena, buffers, aliases, sd, strict=True):
    """copy a reference-style state_dict into the arena; returns (missing, unexpected)"""
    seen = set()
    unexpected = []
    with torch.no_grad():
        for k, v in sd.items():
            name = k
            for alias, canon in aliases:
                if k.startswith(alias):
                    name = canon + k[len(alias):]
            if name in arena.views:
                arena.views[name].copy_(v.to(arena.views[name].dtype))
                seen.add(name)
            elif name in buffers:
                buffers[name].copy_(v)
                seen.add(name)
            else:
                unexpected.append(k)
    missing = [n for n in arena.names() if n not in seen]
    if strict and missing:
        raise KeyError("missing keys in state_dict: %s" % missing[:8])
    return missing, unexpected
