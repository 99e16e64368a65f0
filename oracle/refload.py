"""ORACLE support (test infrastructure only): import the REAL reference modules from /root/reference.

Usable where the reference tree exists: /root/reference in this container, or its copy under baseline/_ref on the GPU box
(baseline/install_ref.py; bench.py's reference arm runs there).  Used by
oracle/gen_golden.py to pin oracle/iq_oracle.py and to write tests/golden/*.pt.  Shims
(nothing under /root/reference is edited):
  1. numpy.float = float            (models/transformer_layers.py:551 uses the removed alias)
  2. torchvision resnet18(pretrained=True) -> weights=None (no network), then the trunk is CUT:
     encoder_cnn.cnn = Linear(FEAT, H) initialised like encoder_cnn.py:27-28, fed pooled features.
  3. `eps` injection: torch.randn is patched while Latent.forward runs (transformer_layers.py:45).
  4. dropout: every nn.Dropout p := 0 for parity runs.
"""
import contextlib
import os
import sys

import numpy as np
import torch
import torch.nn as nn

_HERE = os.path.dirname(os.path.abspath(__file__))


def _find_root():
    """the reference tree: $B200IQ_REFERENCE, /root/reference (this container), else the copy baseline/install_ref.py made under
    baseline/_ref (what the GPU box has)"""
    for p in (os.environ.get("B200IQ_REFERENCE"), "/root/reference", os.path.join(os.path.dirname(_HERE), "baseline", "_ref")):
        if p and os.path.isdir(os.path.join(p, "models")):
            return p
    return "/root/reference"


REF_ROOT = _find_root()


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "models"))


def import_models():
    if not hasattr(np, "float"):
        np.float = float
    import torchvision.models as tvm
    if not getattr(tvm, "_b200iq_patched", False):
        orig = tvm.resnet18

        def resnet18(pretrained=False, **kw):
            return orig(weights=None, **kw)

        tvm.resnet18 = resnet18
        tvm._b200iq_patched = True
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import models  # noqa: the reference package
    return models


def build_reference_iq(vocab, args, latent_mode, seed=0, dropout=False):
    """models/iq.py:22 IQ with the trunk cut at pooled features; weights from the reference's own inits"""
    models = import_models()
    torch.manual_seed(seed)
    np.random.seed(seed)
    with open(os.devnull, "w") as dn, contextlib.redirect_stdout(dn):
        m = models.IQ(latent_mode, vocab, args)
    fc = nn.Linear(args.feat_dim, args.hidden_dim)
    fc.weight.data.normal_(0.0, 0.02)  # encoder_cnn.py:27-28
    fc.bias.data.fill_(0)
    m.encoder_cnn.cnn = fc
    if not dropout:
        for mod in m.modules():
            if isinstance(mod, nn.Dropout):
                mod.p = 0.0
    return m


def head_state_dict(m):
    """state_dict with the cut trunk renamed to the reference's own key (encoder_cnn.cnn.fc.*)"""
    sd = {}
    for k, v in m.state_dict().items():
        if k.startswith("encoder_cnn.cnn."):
            k = k.replace("encoder_cnn.cnn.", "encoder_cnn.cnn.fc.")
        sd[k] = v.detach().clone()
    return sd


def head_grads(m):
    g = {}
    for k, p in m.named_parameters():
        if k.startswith("encoder_cnn.cnn."):
            k = k.replace("encoder_cnn.cnn.", "encoder_cnn.cnn.fc.")
        g[k] = None if p.grad is None else p.grad.detach().clone()
    return g


@contextlib.contextmanager
def inject_eps(eps):
    """make torch.randn(size) return `eps` (Latent.forward draws exactly once per call)"""
    orig = torch.randn

    def fake(*size, **kw):
        if len(size) == 1 and not isinstance(size[0], int):
            size = tuple(size[0])
        assert tuple(size) == tuple(eps.shape), (size, eps.shape)
        return eps.clone()

    torch.randn = fake
    try:
        yield
    finally:
        torch.randn = orig


def import_calculate_losses():
    """train_iq.TrainIQ.calculate_losses, usable verbatim after stubbing pytorch_lightning / utils"""
    import types
    if "train_iq" in sys.modules:
        return sys.modules["train_iq"].TrainIQ.calculate_losses
    pl = types.ModuleType("pytorch_lightning")

    class LightningModule(nn.Module):
        def log(self, *a, **k):
            pass

    pl.LightningModule = LightningModule
    pl.Trainer = object
    pl.Callback = object
    pl.seed_everything = lambda *a, **k: None
    cb = types.ModuleType("pytorch_lightning.callbacks")
    cb.Callback = object
    cb.ModelCheckpoint = object
    es = types.ModuleType("pytorch_lightning.callbacks.early_stopping")
    es.EarlyStopping = object
    sys.modules.setdefault("pytorch_lightning", pl)
    sys.modules.setdefault("pytorch_lightning.callbacks", cb)
    sys.modules.setdefault("pytorch_lightning.callbacks.early_stopping", es)

    class _Any(types.ModuleType):
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return lambda *a, **k: None

    for name in ("utils", "utils.vocab", "utils.data_loader", "utils.train_utils", "utils.nlg_eval", "nlgeval", "nltk",
                 "h5py", "torchtext", "progressbar"):
        sys.modules.setdefault(name, _Any(name))
    import_models()
    import train_iq
    return train_iq.TrainIQ.calculate_losses
