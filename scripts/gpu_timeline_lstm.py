"""kernel timeline of one steady-state LSTM-variant training step (cfg2 by default) through torch.profiler (CUPTI)
-> gpurun_out/timeline_lstm_<tag>.json and a per-kernel / per-phase summary.   python scripts/gpu_timeline_lstm.py <tag> [config]"""
import collections, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from blt_vqg_b200 import _lib, synth
from blt_vqg_b200.lstm_iq import LSTMIQ
from blt_vqg_b200.train import TrainStep
_lib.require_device()
tag = sys.argv[1] if len(sys.argv) > 1 else "t"
name = sys.argv[2] if len(sys.argv) > 2 else "cfg2"
cfg = dict(synth.LSTM_CONFIGS[name])
torch.manual_seed(0)
m = LSTMIQ.from_config(cfg).cuda().train()
ts = TrainStep(m, lr_fn=lambda s: 1e-4)
bt = {k: v.cuda() for k, v in synth.make_batch(cfg, seed=1).items()}
args = (bt["feats"], bt["answers"].int(), bt["alengths"].int(), bt["questions"].int())
for _ in range(6): ts.step(*args)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): ts.step(*args)
    torch.cuda.synchronize()
rows = []
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        rows.append((e.time_range.start, e.time_range.end - e.time_range.start, getattr(e, "device_resource_id", 0), e.name))
rows.sort()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/timeline_lstm_%s.json" % tag, "w"))
ad = [i for i, r in enumerate(rows) if "adam" in r[3]]
s = rows[ad[0] + 1:ad[1] + 1]
t0 = s[0][0]
print("step span %.1f us, %d kernels, sum of durations %.1f us" % (s[-1][0] + s[-1][1] - t0, len(s), sum(r[1] for r in s)))
agg = collections.defaultdict(lambda: [0, 0.0])
for r in s:
    n = r[3].replace("void ", "").replace("b200iq::", "").split("(")[0][:60]
    agg[n][0] += 1
    agg[n][1] += r[1]
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:16]:
    print("%-62s %4d %8.1f" % (n, c, t))
print("sequence of kernels >= 25 us or recurrences (start ms, dur us, stream):")
for r in s:
    if r[1] >= 25 or "lstm_seq" in r[3]:
        print("  %6.3f %7.1f  s%-3d %s" % ((r[0] - t0) / 1e3, r[1], r[2], r[3].replace("void ", "").replace("b200iq::", "").split("(")[0][:60]))
