"""ncu target: the FFN-1 GEMM (10240 x 2048 x 512, bias + relu + dropout epilogue); argv[1] = tile policy, argv[2] = kind"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import _lib, ops
_lib.require_device()
tile = int(sys.argv[1]) if len(sys.argv) > 1 else 0
kind = sys.argv[2] if len(sys.argv) > 2 else "drop"
_lib.lib().b200iq_set_gemm_tile(tile)
M, N, K = 10240, int(sys.argv[3]) if len(sys.argv) > 3 else 2048, 512
A, B, C = torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda"), torch.empty(M, N, device="cuda")
bias, R = torch.randn(N, device="cuda"), torch.randn(M, N, device="cuda")
for _ in range(5):
    if kind == "drop": ops.gemm(A, B, C, bias=bias, relu=True, dropout=(0.1, 5, 1 << 40))
    elif kind == "gate": ops.gemm(A, B, C, relu_mask=R)
    else: ops.gemm(A, B, C, bias=bias, relu=True)
torch.cuda.synchronize()
