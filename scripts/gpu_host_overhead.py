"""host-side cost per launch: python scripts/gpu_host_overhead.py  (prints us per call for a small GEMM / elementwise op, and a
cProfile of one LSTM training step)"""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import _lib, ops, synth
from blt_vqg_b200.lstm_iq import LSTMIQ
from blt_vqg_b200.train import TrainStep
_lib.require_device()
a, b, c = torch.randn(256, 256, device="cuda"), torch.randn(256, 256, device="cuda"), torch.empty(256, 256, device="cuda")
for name, fn in (("gemm", lambda: ops.gemm(a, b, c)), ("axpby", lambda: ops.axpby(a, b, c))):
    for _ in range(50): fn()
    torch.cuda.synchronize()
    with _lib.pinned_stream():
        t0 = time.perf_counter()
        for _ in range(2000): fn()
        t1 = time.perf_counter()
    torch.cuda.synchronize()
    print("%s: %.2f us per call (host, stream cached)" % (name, 1e6 * (t1 - t0) / 2000))
cfg = dict(synth.LSTM_CONFIGS["cfg2"])
m = LSTMIQ.from_config(cfg).cuda().train()
ts = TrainStep(m, lr_fn=lambda s: 1e-4)
bt = {k: v.cuda() for k, v in synth.make_batch(cfg, seed=1).items()}
args = (bt["feats"], bt["answers"], bt["alengths"], bt["questions"])
for _ in range(5): ts.step(*args)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20): ts.step(*args)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print("LSTM step: host %.3f ms, with final sync %.3f ms" % (1e3 * (t1 - t0) / 20, 1e3 * (t2 - t0) / 20))
pr = cProfile.Profile(); pr.enable()
for _ in range(10): ts.step(*args)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
