"""host issue time vs device time of one transformer training step (cfg3 by default):
   python scripts/gpu_host_step.py [config] [profile]"""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import _lib, synth
from blt_vqg_b200.iq import IQ
from blt_vqg_b200.train import TrainStep
_lib.require_device()
name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
cfg = dict(synth.CONFIGS[name])
m = IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg)).cuda().train()
ts = TrainStep(m)
bt = {k: v.cuda() for k, v in synth.make_batch(cfg, seed=1).items()}
args = (bt["feats"], bt["answers"], bt["posteriors"], bt["questions"])
for _ in range(5): ts.step(*args)
torch.cuda.synchronize()
N = 20
t0 = time.perf_counter()
for _ in range(N): ts.step(*args)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print("%s step: host issue %.3f ms, with final sync %.3f ms" % (name, 1e3 * (t1 - t0) / N, 1e3 * (t2 - t0) / N))
iso = []
for _ in range(10):   # one step issued into an idle queue: never blocks on queue depth, so this is the true host cost
    torch.cuda.synchronize(); t0 = time.perf_counter(); ts.step(*args); iso.append(time.perf_counter() - t0)
torch.cuda.synchronize()
print("isolated host issue per step: min %.3f ms, median %.3f ms" % (1e3 * min(iso), 1e3 * sorted(iso)[5]))
if len(sys.argv) > 2:
    pr = cProfile.Profile(); pr.enable()
    for _ in range(10): ts.step(*args)
    pr.disable(); torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("tottime").print_stats(28)
