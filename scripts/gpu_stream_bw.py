"""HBM streaming rates of the arena-wide kernels (round_tf32, grad_sumsq, adam) next to torch's copy / sum on the same sizes:
   python scripts/gpu_stream_bw.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import _lib, ops
_lib.require_device()
n = 32_600_000
NS = 4
a = [torch.randn(n, device="cuda") for _ in range(NS)]
b = [torch.empty(n, device="cuda") for _ in range(NS)]
m = [torch.zeros(n, device="cuda") for _ in range(NS)]
v = [torch.zeros(n, device="cuda") for _ in range(NS)]
hyper = torch.zeros(4, device="cuda"); ops.set_hyper(hyper, 1e-4, 0.1, 0.001, 1.0)
nsq = torch.zeros(1, device="cuda")
def t(name, fn, bytes_):
    for i in range(NS): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(5):
        for i in range(NS): fn(i)
    e1.record(); torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / (5 * NS)
    print("%-28s %8.1f us  %6.2f TB/s" % (name, us, bytes_ / us * 1e-6))
t("torch copy_", lambda i: b[i].copy_(a[i]), 8 * n)
t("torch sum", lambda i: a[i].sum(), 4 * n)
t("torch mul_", lambda i: b[i].mul_(1.0001), 8 * n)
t("round_tf32", lambda i: ops.round_tf32(a[i], b[i]), 8 * n)
t("grad_sumsq", lambda i: ops.grad_sumsq(a[i], nsq), 4 * n)
t("adam_step", lambda i: ops.adam_step(b[i], a[i], m[i], v[i], hyper, nsq, 0.9, 0.98, 1e-9, 5.0, zero_grad=False), 28 * n)
t("torch addcmul_ (3r 1w)", lambda i: b[i].addcmul_(m[i], v[i]), 16 * n)
