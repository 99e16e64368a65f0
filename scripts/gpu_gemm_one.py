"""run one GEMM shape a few times (ncu target): python scripts/gpu_gemm_one.py M N K [a_t b_t split_k]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import ops, _lib
_lib.require_device()
M, N, K = (int(x) for x in sys.argv[1:4])
a_t, b_t, sk = (int(x) for x in sys.argv[4:7]) if len(sys.argv) > 6 else (0, 0, 1)
A = torch.randn((K, M) if a_t else (M, K), device="cuda"); B = torch.randn((K, N) if b_t else (N, K), device="cuda")
out = torch.zeros(M, N, device="cuda"); bias = torch.randn(N, device="cuda")
for _ in range(5):
    ops.gemm(A, B, out, a_t=bool(a_t), b_t=bool(b_t), bias=None if sk > 1 else bias, relu=sk == 1, accumulate=sk > 1, split_k=sk)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    ops.gemm(A, B, out, a_t=bool(a_t), b_t=bool(b_t), bias=None if sk > 1 else bias, relu=sk == 1, accumulate=sk > 1, split_k=sk)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
print("M=%d N=%d K=%d a_t=%d b_t=%d sk=%d: %.4f ms  %.1f TF/s" % (M, N, K, a_t, b_t, sk, ms, 2e-9 * M * N * K / ms))
