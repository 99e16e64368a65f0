"""Time the cfg3 GEMM shapes under every tile policy (128 / 256 / CTA pair) with rotating operands (no L2 reuse between
launches).   python scripts/gpu_gemm_sweep.py [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from blt_vqg_b200 import _lib, ops  # noqa: E402

_lib.require_device()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
# (M, N, K, a_t, b_t, kind, split_k)
SHAPES = [
    (10240, 1536, 512, 0, 0, "plain", 1), (10240, 512, 512, 0, 0, "res", 1), (10240, 512, 512, 0, 1, "plain", 1),
    (10240, 2048, 512, 0, 0, "drop", 1), (10240, 2048, 512, 0, 0, "plain", 1), (10240, 512, 2048, 0, 0, "drop", 1),
    (10240, 2048, 512, 0, 1, "gate", 1), (10240, 512, 2048, 0, 1, "plain", 1), (10240, 512, 1536, 0, 1, "plain", 1),
    (512, 2048, 10240, 1, 1, "acc", 4), (2048, 512, 10240, 1, 1, "acc", 4), (512, 512, 10240, 1, 1, "acc", 18),
    (1536, 512, 10240, 1, 1, "acc", 6), (768, 15000, 512, 0, 0, "plain", 1), (15000, 512, 768, 1, 1, "acc", 1),
    (768, 512, 15000, 0, 1, "acc", 6), (2560, 1536, 512, 0, 0, "plain", 1), (2560, 2048, 512, 0, 0, "drop", 1),
    (2560, 512, 512, 0, 0, "res", 1),
]
NSET = 6
print("%-44s %9s %9s %9s   best" % ("shape", "t128 us", "t256 us", "pair us"))
for (M, N, K, a_t, b_t, kind, sk) in SHAPES:
    As = [torch.randn((K, M) if a_t else (M, K), device="cuda") for _ in range(NSET)]
    Bs = [torch.randn((K, N) if b_t else (N, K), device="cuda") for _ in range(NSET)]
    Cs = [torch.zeros(M, N, device="cuda") for _ in range(NSET)]
    R = [torch.randn(M, N, device="cuda") for _ in range(NSET)] if kind in ("res", "gate") else None
    bias = torch.randn(N, device="cuda")

    def run(i):
        kw = dict(a_t=bool(a_t), b_t=bool(b_t))
        if kind == "plain": kw.update(bias=bias, relu=True)
        elif kind == "drop": kw.update(bias=bias, relu=True, dropout=(0.1, 5, 1 << 40))
        elif kind == "res": kw.update(residual=R[i])
        elif kind == "gate": kw.update(relu_mask=R[i])
        elif kind == "acc": kw.update(accumulate=True, split_k=sk)
        ops.gemm(As[i], Bs[i], Cs[i], **kw)
    res = []
    for tile in (128, 256, 512):
        prev = _lib.lib().b200iq_set_gemm_tile(tile)
        for i in range(NSET):
            run(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for r in range(reps):
            for i in range(NSET):
                run(i)
        e1.record()
        torch.cuda.synchronize()
        _lib.lib().b200iq_set_gemm_tile(prev)
        res.append(1e3 * e0.elapsed_time(e1) / (reps * NSET))
    best = min(res)
    print("%6d %6d %6d a_t=%d b_t=%d %-5s sk=%-2d   %9.1f %9.1f %9.1f   %s %.0f TF/s" % (
        M, N, K, a_t, b_t, kind, sk, res[0], res[1], res[2], ("128", "256", "pair")[res.index(best)], 2e-6 * M * N * K / best))
    del As, Bs, Cs, R
