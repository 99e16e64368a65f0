"""GPU: every C-ABI kernel against a plain torch fp32/fp64 statement of the same op (called through ctypes)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(built_lib, request):
    from blt_vqg_b200 import _lib, ops as o
    _lib.require_device()
    prev = _lib.lib().b200iq_set_tf32_rounding(0)  # kernels are checked in exact mode; rounding has its own test
    request.addfinalizer(lambda: _lib.lib().b200iq_set_tf32_rounding(prev))
    torch.manual_seed(0)
    return o


def rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (256, 384, 300), (1000, 200, 1024), (37, 52, 20), (640, 15000, 512)])
@pytest.mark.parametrize("a_t,b_t", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("tile", [0, 512])
def test_gemm_all_layouts(ops, M, N, K, a_t, b_t, tile, request):
    """tile=512 forces the CTA-pair kernel (tcgen05.mma.cta_group::2, 256 x 256 per pair) wherever M, N > 128"""
    from blt_vqg_b200 import _lib
    prev = _lib.lib().b200iq_set_gemm_tile(tile)
    request.addfinalizer(lambda: _lib.lib().b200iq_set_gemm_tile(prev))
    if (a_t and M % 4) or (b_t and N % 4) or (not a_t and K % 4) or (not b_t and K % 4):
        pytest.skip("TMA needs 16-byte row pitch")
    dev = "cuda"
    A, B = torch.randn(M, K, device=dev), torch.randn(N, K, device=dev)
    Aa = A.t().contiguous() if a_t else A
    Bb = B.t().contiguous() if b_t else B
    bias, res = torch.randn(N, device=dev), torch.randn(M, N, device=dev)
    ref = A.double() @ B.double().t()
    for simt, tol in ((True, 1e-5), (False, 2e-3)):  # exact fp32 CUDA cores / tcgen05 TF32 (10-bit mantissa)
        assert rel(ops.gemm(Aa, Bb, a_t=a_t, b_t=b_t, simt=simt), ref) < tol
        out = ops.gemm(Aa, Bb, a_t=a_t, b_t=b_t, bias=bias, residual=res, relu=True, alpha=0.5, simt=simt)
        assert rel(out, torch.relu(0.5 * ref + bias.double() + res.double())) < tol
        acc = torch.ones(M, N, device=dev)
        ops.gemm(Aa, Bb, acc, a_t=a_t, b_t=b_t, accumulate=True, split_k=4, simt=simt)
        assert rel(acc, ref + 1.0) < tol
        gate = torch.randn(M, N, device=dev)
        out = ops.gemm(Aa, Bb, a_t=a_t, b_t=b_t, relu_mask=gate, simt=simt)
        assert rel(out, ref * (gate > 0)) < tol


@pytest.mark.parametrize("tile", [128, 256, 512])
def test_gemm_tile_policies_agree_on_multi_wave_shapes(ops, tile, request):
    """every tile policy on shapes with several persistent waves, ragged edges and split-K (the cfg3 GEMM shapes, scaled)"""
    from blt_vqg_b200 import _lib
    prev = _lib.lib().b200iq_set_gemm_tile(tile)
    request.addfinalizer(lambda: _lib.lib().b200iq_set_gemm_tile(prev))
    for M, N, K, a_t, b_t, sk in ((10240, 2048, 512, False, False, 1), (10240, 512, 2048, False, True, 1), (5000, 1000, 512, False, False, 1),
                                  (512, 2048, 10240, True, True, 4), (768, 15000, 512, False, False, 1), (300, 260, 96, False, True, 1)):
        A, B = torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda")
        Aa = A.t().contiguous() if a_t else A
        Bb = B.t().contiguous() if b_t else B
        ref = A.double() @ B.double().t()
        if sk > 1:
            acc = torch.ones(M, N, device="cuda")
            ops.gemm(Aa, Bb, acc, a_t=a_t, b_t=b_t, accumulate=True, split_k=sk)
            assert rel(acc, ref + 1.0) < 2e-3
        else:
            bias, gate = torch.randn(N, device="cuda"), torch.randn(M, N, device="cuda")
            assert rel(ops.gemm(Aa, Bb, a_t=a_t, b_t=b_t, bias=bias, relu=True), torch.relu(ref + bias.double())) < 2e-3
            assert rel(ops.gemm(Aa, Bb, a_t=a_t, b_t=b_t, relu_mask=gate), ref * (gate > 0)) < 2e-3


def test_gemm_rejects_bad_arguments(ops):
    from blt_vqg_b200._lib import B200iqError
    a, b = torch.randn(8, 6, device="cuda"), torch.randn(8, 6, device="cuda")
    with pytest.raises(B200iqError):
        ops.gemm(a, b)  # K=6: leading dimension not a multiple of 4 floats
    with pytest.raises(B200iqError):
        ops.gemm(torch.randn(8, 8, device="cuda"), torch.randn(8, 8, device="cuda"), torch.zeros(8, 8, device="cuda"), split_k=2)


def test_embedding_gather_scatter(ops):
    V, E, M = 300, 52, 1000
    table = torch.randn(V, E, device="cuda")
    ids = torch.randint(0, V, (M,), device="cuda", dtype=torch.int32)
    ids[::7] = 0
    out = torch.empty(M, E, device="cuda")
    ops.embed_gather(ids, table, out)
    assert torch.equal(out, table[ids.long()])
    dout = torch.randn(M, E, device="cuda")
    dt = torch.zeros(V, E, device="cuda")
    ops.embed_scatter_add(ids, dout, dt, 0)
    ref = torch.zeros(V, E, device="cuda", dtype=torch.float64).index_add_(0, ids.long(), dout.double())
    ref[0] = 0  # padding_idx row gets no gradient
    assert rel(dt, ref) < 1e-6


def test_layernorm_fwd_bwd(ops):
    for M, H in ((333, 64), (1000, 512), (77, 300), (10240, 512), (5, 128), (3001, 1024), (100, 256)):
        x = torch.randn(M, H, device="cuda", requires_grad=True)
        g, b = torch.randn(H, device="cuda", requires_grad=True), torch.randn(H, device="cuda", requires_grad=True)
        y = torch.empty(M, H, device="cuda")
        mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
        ops.layernorm_fwd(x.detach(), g.detach(), b.detach(), y, mean, rstd)
        yr = torch.nn.functional.layer_norm(x, (H,), g, b, 1e-5)
        assert rel(y, yr) < 1e-5
        dy = torch.randn(M, H, device="cuda")
        yr.backward(dy)
        dx = torch.ones(M, H, device="cuda")
        dg, db = torch.zeros(H, device="cuda"), torch.zeros(H, device="cuda")
        ops.layernorm_bwd(dy, x.detach(), g.detach(), mean, rstd, dx, dg, db, True)
        assert rel(dx - 1.0, x.grad) < 1e-4
        assert rel(dg, g.grad) < 1e-4 and rel(db, b.grad) < 1e-4


def test_add_layernorm_is_axpby_then_layernorm_bit_for_bit(ops):
    """the fused residual-add + LayerNorm (one pass) writes exactly what b200iq_axpby followed by b200iq_layernorm_fwd write"""
    for M, H in ((1000, 512), (10240, 512), (5, 128), (3001, 1024), (100, 256)):
        assert ops.add_layernorm_supported(torch.empty(M, H))
        x, f = torch.randn(M, H, device="cuda"), torch.randn(M, H, device="cuda").relu_()
        g, b = torch.randn(H, device="cuda"), torch.randn(H, device="cuda")
        s0, y0 = torch.empty(M, H, device="cuda"), torch.empty(M, H, device="cuda")
        m0, r0 = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
        ops.axpby(x, f, s0)
        ops.layernorm_fwd(s0, g, b, y0, m0, r0)
        s1, y1 = torch.empty(M, H, device="cuda"), torch.empty(M, H, device="cuda")
        m1, r1 = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
        ops.add_layernorm_fwd(x, f, s1, g, b, y1, m1, r1)
        assert torch.equal(s1, s0) and torch.equal(y1, y0) and torch.equal(m1, m0) and torch.equal(r1, r0)
        assert rel(y1, torch.nn.functional.layer_norm(x + f, (H,), g, b, 1e-5)) < 2e-3   # TF32-rounded sum and output
    assert not ops.add_layernorm_supported(torch.empty(8, 300))


def test_small_elementwise(ops):
    B, n, H = 9, 5, 64
    x = torch.randn(B * n, H, device="cuda")
    timing, a0, a1 = torch.randn(n, H, device="cuda"), torch.randn(B, H, device="cuda"), torch.randn(B, H, device="cuda")
    ref = x.view(B, n, H) + timing[None]
    ref[:, 0] += a0 + a1
    ops.add_timing(x, timing, a0, a1, B, n, H)
    assert rel(x.view(B, n, H), ref) < 1e-6
    wide = torch.zeros(B, 2 * H, device="cuda")
    ops.tok0_gather(x, wide[:, H:], B, n, H)
    assert torch.equal(wide[:, H:], x.view(B, n, H)[:, 0]) and float(wide[:, :H].abs().max()) == 0
    dst = torch.zeros(B * n, H, device="cuda")
    ops.tok0_scatter_add(wide[:, H:], dst, B, n, H)
    assert torch.equal(dst.view(B, n, H)[:, 0], wide[:, H:]) and float(dst.view(B, n, H)[:, 1:].abs().max()) == 0
    y = torch.randn(500, 96, device="cuda")
    cs = torch.ones(96, device="cuda")
    ops.colsum_add(y, cs)
    assert rel(cs, 1 + y.double().sum(0)) < 1e-5
    dy = torch.randn(500, 96, device="cuda")
    dpre, dbias = torch.empty_like(dy), torch.zeros(96, device="cuda")
    ops.relu_bwd_colsum(dy, y, dpre, dbias)
    assert torch.equal(dpre, dy * (y > 0)) and rel(dbias, (dy * (y > 0)).double().sum(0)) < 1e-5
    for M, N in ((10240, 2048), (37, 512), (5000, 100)):   # float4 path incl. ragged row blocks
        yy, dd = torch.randn(M, N, device="cuda"), torch.randn(M, N, device="cuda")
        cs = torch.zeros(N, device="cuda")
        ops.colsum_add(yy, cs)
        assert rel(cs, yy.double().sum(0)) < 1e-5
        dp, dbb = torch.empty_like(dd), torch.zeros(N, device="cuda")
        ops.relu_bwd_colsum(dd, yy, dp, dbb)
        assert torch.equal(dp, dd * (yy > 0)) and rel(dbb, (dd * (yy > 0)).double().sum(0)) < 1e-5
    out = torch.empty_like(y)
    ops.axpby(y, dy, out, 0.5, -2.0)
    assert rel(out, 0.5 * y - 2.0 * dy) < 1e-6


def test_philox_randn_and_dropout(ops):
    n = 1 << 20
    a, b = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
    ops.randn_(a, 1234, 0)
    ops.randn_(b, 1234, 0)
    assert torch.equal(a, b)  # counter-based: same key, same numbers
    ops.randn_(b, 1235, 0)
    assert not torch.equal(a, b)
    assert abs(float(a.mean())) < 5e-3 and abs(float(a.std()) - 1) < 5e-3
    x = torch.ones(n, device="cuda")
    ops.dropout_(x, 0.1, 7, 1 << 40)
    keep = float((x > 0).float().mean())
    assert abs(keep - 0.9) < 3e-3 and float(x.max()) == pytest.approx(1 / 0.9)
    g = torch.ones(n, device="cuda")
    ops.dropout_(g, 0.1, 7, 1 << 40)
    assert torch.equal(g, x)  # backward re-creates the same mask


@pytest.mark.parametrize("simt", [False, True])
@pytest.mark.parametrize("M,N,K", [(300, 256, 96), (1000, 2048, 64), (77, 36, 40)])
def test_gemm_fused_relu_dropout_matches_the_standalone_mask(ops, M, N, K, simt):
    """GEMM epilogue dropout == the b200iq_dropout mask of the same (seed, offset) on the contiguous output; and the
    backward needs no Philox: (y > 0) of the post-dropout tensor masks relu and dropout at once, times 1/(1-p)."""
    a, w, b = torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda"), torch.randn(N, device="cuda")
    key = (0.25, 4321, 3 << 40)
    y0 = ops.gemm(a, w, bias=b, relu=True, simt=simt)
    ops.dropout_(y0, *key)
    y1 = ops.gemm(a, w, bias=b, relu=True, simt=simt, dropout=key)
    assert torch.equal(y1 > 0, y0 > 0)
    assert rel(y1, y0) < 1e-6
    keep = float((y1 > 0).sum()) / float((ops.gemm(a, w, bias=b, relu=True, simt=simt) > 0).sum())
    assert abs(keep - 0.75) < 0.03
    dy = torch.randn(M, N, device="cuda")
    dpre, db = torch.empty_like(dy), torch.zeros(N, device="cuda")
    ops.relu_bwd_colsum(dy, y1, dpre, db, scale=1 / 0.75)
    want = dy.clone()
    ops.dropout_(want, *key)          # dropout backward with the regenerated mask ...
    want = want * (y0 > 0)            # ... then relu backward
    assert rel(dpre, want) < 1e-6 and rel(db, want.double().sum(0)) < 1e-5


def test_tf32_rounding_mode(ops):
    """b200iq_set_tf32_rounding(1): producers write round-to-nearest TF32 values, so the tensor core's operand truncation
    is a no-op and GEMM results are unbiased; with truncation alone every dot product shrinks by ~7e-4."""
    from blt_vqg_b200 import _lib
    x = torch.rand(1 << 16, device="cuda") + 0.5
    r = ops.round_tf32(x, torch.empty_like(x))
    assert float((r.view(torch.int32) & 0x1FFF).abs().max()) == 0           # low 13 mantissa bits cleared
    assert float(((r - x) / x).abs().max()) <= 2.0 ** -11 + 1e-9            # nearest, not truncated
    assert abs(float(((r - x) / x).mean())) < 2e-5                          # unbiased
    A, B = torch.rand(512, 1024, device="cuda") + 0.5, torch.rand(256, 1024, device="cuda") + 0.5
    ref = A.double() @ B.double().t()
    trunc = ops.gemm(A, B, simt=False)
    bias_trunc = float(((trunc.double() - ref) / ref).mean())
    Ar, Br = ops.round_tf32(A, torch.empty_like(A)), ops.round_tf32(B, torch.empty_like(B))
    rn = ops.gemm(Ar, Br, simt=False)
    bias_rn = float(((rn.double() - ref) / ref).mean())
    assert bias_trunc < -3e-4 and abs(bias_rn) < 3e-5, (bias_trunc, bias_rn)
    _lib.lib().b200iq_set_tf32_rounding(1)
    try:
        y = torch.empty(64, 96, device="cuda")
        xin = torch.randn(64, 96, device="cuda")
        ops.layernorm_fwd(xin, torch.ones(96, device="cuda"), torch.zeros(96, device="cuda"), y, torch.empty(64, device="cuda"), torch.empty(64, device="cuda"))
        assert float((y.view(torch.int32) & 0x1FFF).abs().max()) == 0
        out = ops.gemm(Ar, Br, simt=False)
        assert float((out.view(torch.int32) & 0x1FFF).abs().max()) == 0
        out = ops.gemm(Ar, Br, simt=False, fp32_out=True)
        assert float((out.view(torch.int32) & 0x1FFF).abs().max()) != 0
    finally:
        _lib.lib().b200iq_set_tf32_rounding(0)
