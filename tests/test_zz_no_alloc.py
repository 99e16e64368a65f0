"""GPU: b200iq_lstm_seq_bwd_static — the allocation-free BPTT entry point (scratch in a file-scope device array of the library) —
against b200iq_lstm_seq_bwd with a caller-owned workspace, through the C ABI, on two calls in a row.  Both run the same kernel in
the same (fixed) reduction order, so the outputs are compared bit for bit.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(built_lib):
    from blt_vqg_b200 import _lib
    _lib.require_device()
    L = _lib.lib()
    L.b200iq_lstm_seq_bwd_workspace_bytes.restype = C.c_size_t
    L.b200iq_lstm_seq_bwd_static_max_bytes.restype = C.c_size_t
    return L


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _case(lib, T, B, H, ragged, seed):
    """activated gates / cell states of a real forward pass (b200iq_lstm_seq_fwd), then the BPTT through both entry points"""
    from blt_vqg_b200 import _lib
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(seed)
    rn = lambda *s: torch.randn(*s, device="cuda", generator=g)
    whh = rn(4 * H, H) * H ** -0.5
    gates = rn(T, B, 4 * H)
    hbuf, cbuf = torch.zeros(T + 1, B, H, device="cuda"), torch.zeros(T + 1, B, H, device="cuda")
    hbuf[0].copy_(rn(B, H) * 0.5)
    cbuf[0].copy_(rn(B, H) * 0.5)
    lengths = torch.randint(1, T + 1, (B,), device="cuda", dtype=torch.int32, generator=g) if ragged else None
    _lib.check(lib.b200iq_lstm_seq_fwd(_p(gates), _p(whh), _p(hbuf), _p(cbuf), _p(lengths), T, B, H, st), "lstm_seq_fwd")
    dh_last, d_out = rn(B, H), rn(T, B, H)
    out = []
    for static in (False, True):
        # NaN-filled outputs: every element must be written by the call itself
        dg, dh0, dc0 = (torch.full(s, float("nan"), device="cuda") for s in ((T, B, 4 * H), (B, H), (B, H)))
        if static:
            rc = lib.b200iq_lstm_seq_bwd_static(_p(gates), _p(cbuf), _p(whh), _p(dh_last), _p(d_out), _p(dg), _p(dh0), _p(dc0),
                                                _p(lengths), T, B, H, st)
        else:
            ws = torch.empty(lib.b200iq_lstm_seq_bwd_workspace_bytes(B, H) // 4, device="cuda")
            rc = lib.b200iq_lstm_seq_bwd(_p(gates), _p(cbuf), _p(whh), _p(dh_last), _p(d_out), _p(dg), _p(dh0), _p(dc0), _p(ws),
                                         _p(lengths), T, B, H, st)
        _lib.check(rc, "lstm_seq_bwd%s" % ("_static" if static else ""))
        torch.cuda.synchronize()
        out.append((dg, dh0, dc0))
    for a, b, name in zip(out[0], out[1], ("dgates", "dh0", "dc0")):
        assert torch.isfinite(b).all(), name
        assert float(b.abs().max()) > 0, name
        assert torch.equal(a, b), (name, float((a - b).abs().max()))


def test_static_scratch_bptt_matches_the_workspace_entry_point_twice_in_a_row(lib):
    assert lib.b200iq_lstm_seq_bwd_static_max_bytes() == 5 << 20
    # first call: the full-size LSTM config's shape (B = 256, H = 300: 4.98 MB of the 5 MiB array); second call: a smaller ragged
    # batch that finds the first call's partial products still in the array
    _case(lib, 19, 256, 300, False, seed=11)
    _case(lib, 6, 37, 64, True, seed=12)
    _case(lib, 4, 256, 304, True, seed=13)


def test_static_scratch_rejects_sizes_beyond_the_array(lib):
    B, H = 272, 300  # 17 clusters: 5.3 MB of partial products
    assert lib.b200iq_lstm_seq_bwd_workspace_bytes(B, H) > lib.b200iq_lstm_seq_bwd_static_max_bytes()
    x = torch.zeros(16, device="cuda")
    rc = lib.b200iq_lstm_seq_bwd_static(_p(x), _p(x), _p(x), None, None, _p(x), _p(x), _p(x), None, 2, B, H, None)
    assert rc < 0
    msg = lib.b200iq_last_error().decode()
    assert "bytes of scratch" in msg and "b200iq_lstm_seq_bwd" in msg
