"""CPU: libb200iq.so loads and exports every symbol include/b200iq.h declares (no compute calls without a GPU);
the host-side parameter arena speaks the reference's state_dict names."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b200iq.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200iq_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(built_lib)
    names = _declared()
    assert len(names) >= 28
    for n in names:
        assert hasattr(lib, n), "libb200iq.so does not export %s" % n
    assert lib.b200iq_abi_version() == 3


def test_argument_errors_are_reported_without_a_launch(built_lib):
    lib = ctypes.CDLL(built_lib)
    lib.b200iq_last_error.restype = ctypes.c_char_p
    rc = lib.b200iq_gemm(None, None)
    assert rc < 0 and b"null" in lib.b200iq_last_error()
    rc = lib.b200iq_layernorm_fwd(None, None, None, None, None, None, 0, 0, ctypes.c_float(1e-5), None)
    assert rc < 0


def test_no_cpu_fallback():
    """the product path fails loudly without a B200"""
    from blt_vqg_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.B200iqError):
        _lib.require_device()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "blt_vqg_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b|import_module\(.oracle|__import__\(.oracle", src, flags=re.M), fn


def test_state_dict_round_trip_with_reference_names(golden_dir):
    from blt_vqg_b200 import synth
    from blt_vqg_b200.iq import IQ
    gold = torch.load(os.path.join(golden_dir, "iq_transformer_tiny_latent.pt"), weights_only=False)
    cfg = gold["cfg"]
    m = IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg))
    missing, unexpected = m.load_state_dict(gold["state_dict"])
    assert not missing and not unexpected
    sd = m.state_dict()
    assert set(sd.keys()) == set(gold["state_dict"].keys())
    for k, v in gold["state_dict"].items():
        assert torch.equal(sd[k].cpu(), v), k
    # a single flat parameter carries everything trainable
    assert [n for n, _ in m.named_parameters()] == ["flat"]


def test_noam_and_kl_weight_schedule():
    from blt_vqg_b200 import synth
    from blt_vqg_b200.iq import IQ
    from blt_vqg_b200.train import noam_lr
    from oracle import iq_oracle
    for s in (0, 1, 100, 4000, 20000):
        assert noam_lr(s, 512) == pytest.approx(iq_oracle.noam_lr(s, 512))
    cfg = synth.CONFIGS["tiny"]
    m = IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg))
    c = m.loss_coefs(kliter=0)
    assert c[2] == pytest.approx(0.5 * (1 - 0.9950547536867305), rel=1e-6)  # tanh(-3)+1
    assert m.loss_coefs(kliter=10 ** 6)[2] == pytest.approx(0.5)
    m.switch_GVT_train_mode(False)
    assert m.loss_coefs(5) == (1.0, 0.1, 0.0, 0.0)
