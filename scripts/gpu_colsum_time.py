"""time b200iq_colsum_add_f16 / b200iq_colsum_add on the step's shapes (graph-replayed, rotating inputs)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import _lib, ops
if os.environ.get("B200IQ_LIB"): _lib.LIB_PATH = os.environ["B200IQ_LIB"]
_lib.require_device()
for (M, N, f16) in [(10240, 2048, True), (10752, 2048, True), (1280, 15000, True), (10240, 512, True), (10240, 512, False), (10240, 2048, False)]:
    xs = [torch.randn(M, N, device="cuda").half() if f16 else torch.randn(M, N, device="cuda") for _ in range(4)]
    out = torch.zeros(N, device="cuda")
    def run(i):
        if f16: ops.colsum_add_act(ops.Act(hi=xs[i]), out)
        else: ops.colsum_add(xs[i], out)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st), _lib.pinned_stream():
        for i in range(4): run(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for r in range(24): run(r % 4)
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): g.replay()
        e1.record(); torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / 120
    by = M * N * (2 if f16 else 4)
    print("colsum %s M=%d N=%d: %.1f us  %.0f GB/s" % ("f16" if f16 else "f32", M, N, us, by / us * 1e-3))
    ref = xs[3].float().sum(0)
    out.zero_(); run(3); torch.cuda.synchronize()
    assert float((out - ref).abs().max()) < 1e-2 * float(ref.abs().max() + 1), "mismatch"
