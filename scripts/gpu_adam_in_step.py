"""why are sumsq / adam / round_tf32 2.5x slower inside a step than alone?  times adam on the model's own arenas (a) after a
synchronize, (b) queued right behind a full step"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import _lib, ops, synth
from blt_vqg_b200.iq import IQ
from blt_vqg_b200.train import TrainStep
_lib.require_device()
cfg = dict(synth.CONFIGS["cfg3"])
m = IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg)).cuda().train()
ts = TrainStep(m)
bt = {k: v.cuda() for k, v in synth.make_batch(cfg, seed=1).items()}
args = (bt["feats"], bt["answers"], bt["posteriors"], bt["questions"])
for _ in range(5): ts.step(*args)
torch.cuda.synchronize()
print("arena floats", m.flat.numel(), "ptrs %x %x %x %x" % (m.flat.data_ptr(), m.grad_arena.data_ptr(), ts.exp_avg.data_ptr(), ts.exp_avg_sq.data_ptr()))
def adam():
    ops.adam_step(m.flat.data, m.grad_arena, ts.exp_avg, ts.exp_avg_sq, ts.hyper, ts.normsq, 0.9, 0.98, 1e-9, 5.0, zero_grad=True)
def timed(fn, pre=None):
    out = []
    for _ in range(5):
        if pre: pre()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        out.append(1e3 * e0.elapsed_time(e1))
    return sorted(out)[2]
print("adam alone after sync: %.1f us" % timed(adam, torch.cuda.synchronize))
print("adam right behind a step (same stream): %.1f us" % timed(adam, lambda: ts.step(*args)))
print("sumsq alone: %.1f us" % timed(lambda: ops.grad_sumsq(m.grad_arena, ts.normsq), torch.cuda.synchronize))
print("sumsq behind a step: %.1f us" % timed(lambda: ops.grad_sumsq(m.grad_arena, ts.normsq), lambda: ts.step(*args)))
big = torch.empty(64 << 20, device="cuda")
print("torch copy 256 MB behind a step: %.1f us (alone %.1f us)" % (timed(lambda: big[:32 << 20].copy_(big[32 << 20:]), lambda: ts.step(*args)),
      timed(lambda: big[:32 << 20].copy_(big[32 << 20:]), torch.cuda.synchronize)))
