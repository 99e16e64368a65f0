"""Operand-rounding error of the tensor-core attention kernels (n <= 32: register kernels; 32 < n <= 80: row-block kernels) against the
exact SIMT kernels on the same TF32-rounded inputs, for plain random inputs and for keys / values with a large common component
(what LayerNorm + a bias-free projection produce at initialisation):   python scripts/gpu_attention_err.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import _lib, ops
_lib.require_device()
torch.manual_seed(0)
B, nh, d = 16, 8, 64
H = nh * d
rel = lambda a, b: float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))
rt = lambda x: ops.round_tf32(x, torch.empty_like(x))
for common in (0.0, 10.0):
    for causal in (False, True):
        for n in (20, 32, 33, 48, 64, 65, 80):
            q = rt(torch.randn(B * n, H, device="cuda") * 0.3)
            k = rt(torch.randn(B * n, H, device="cuda") * 0.3 + common * torch.randn(1, H, device="cuda"))
            v = rt(torch.randn(B * n, H, device="cuda") + common * torch.randn(1, H, device="cuda"))
            do = rt(torch.randn(B * n, H, device="cuda"))
            ids = torch.ones(B, n, device="cuda", dtype=torch.int32)
            def run():
                out = torch.empty(B * n, H, device="cuda"); dq, dk, dv = (torch.empty(B * n, H, device="cuda") for _ in range(3))
                ops.attention_fwd(q, k, v, out, ids, 0, causal, B, nh, n, n, d)
                ops.attention_bwd(q, k, v, do, dq, dk, dv, ids, 0, causal, B, nh, n, n, d)
                return out, dq, dk, dv
            prev = _lib.lib().b200iq_set_tf32_rounding(0)
            ex = run()
            _lib.lib().b200iq_set_tf32_rounding(1)
            tc = run()
            _lib.lib().b200iq_set_tf32_rounding(prev)
            print("common %4.1f causal %d n=%2d  out %.2e dq %.2e dk %.2e dv %.2e" % ((common, causal, n) + tuple(rel(a, b) for a, b in zip(tc, ex))))
