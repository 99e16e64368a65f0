"""kernel timeline of one steady-state cfg3 training step through torch.profiler (CUPTI): stream, start, duration per
launch -> gpurun_out/timeline_<tag>.json   python scripts/gpu_timeline.py <tag> [config]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from blt_vqg_b200 import _lib, synth
from blt_vqg_b200.iq import IQ
from blt_vqg_b200.train import TrainStep
_lib.require_device()
tag = sys.argv[1] if len(sys.argv) > 1 else "t"
name = sys.argv[2] if len(sys.argv) > 2 else "cfg3"
cfg = dict(synth.CONFIGS[name])
dp = None
if "RANK" in os.environ:   # under torchrun: data parallel, rank 0 writes its timeline (NCCL kernels included)
    import torch.distributed as dist
    from blt_vqg_b200.parallel import GradReducer
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl")
    dp = GradReducer()
torch.manual_seed(0)
m = IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg)).cuda().train()
ts = TrainStep(m, dp=dp)
bt = {k: v.cuda() for k, v in synth.make_batch(cfg, seed=1).items()}
args = (bt["feats"], bt["answers"], bt["posteriors"], bt["questions"])
for _ in range(6): ts.step(*args)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): ts.step(*args)
    torch.cuda.synchronize()
rows = []
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        rows.append((e.time_range.start, e.time_range.end - e.time_range.start, getattr(e, "device_resource_id", getattr(e, "thread", 0)), e.name))
rows.sort()
os.makedirs("gpurun_out", exist_ok=True)
if int(os.environ.get("RANK", "0")) == 0:
    json.dump(rows, open("gpurun_out/timeline_%s.json" % tag, "w"))
print(len(rows), "device events; span %.3f ms" % ((rows[-1][0] + rows[-1][1] - rows[0][0]) / 1e3))
