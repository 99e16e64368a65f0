"""the small workload scripts/sanitize.sh runs under compute-sanitizer: pair GEMMs of every layout (1- and 3-term, plane and
fp32 epilogues, split-K reduce-add), the persistent LSTM recurrence forward + BPTT, one tiny transformer and LSTM training step"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from blt_vqg_b200 import _lib, ops, synth
from blt_vqg_b200.iq import IQ
from blt_vqg_b200.lstm_iq import LSTMIQ
from scripts.gpu_f16_check import planes_of

_lib.require_device()
dev = "cuda"
torch.manual_seed(0)
for (M, N, K) in [(512, 512, 256), (300, 200, 136)]:
    A, B = torch.randn(M, K, device=dev), torch.randn(N, K, device=dev)
    for terms in (1, 3):
        a, b = ops.Act(*(None,) + planes_of(A, lo=terms == 3)), ops.Act(*(None,) + planes_of(B, lo=terms == 3))
        out = torch.zeros(M, N, device=dev)
        ops.gemm2(a, b, out, terms=terms)
        pl = ops.Act(*(None,) + planes_of(torch.zeros(M, N, device=dev)))
        ops.gemm2(a, b, planes=pl, relu=True, terms=terms)
    Bt = torch.randn(K, N, device=dev)
    bt = ops.Act(hi=planes_of(Bt, lo=False)[0])
    ops.gemm2(ops.Act(hi=planes_of(A, lo=False)[0]), bt, torch.zeros(M, N, device=dev), b_t=True, terms=1)
    At = torch.randn(K, M, device=dev)
    ops.gemm2(ops.Act(hi=planes_of(At, lo=False)[0]), bt, torch.zeros(M, N, device=dev), a_t=True, b_t=True, accumulate=True, split_k=2, terms=1)
torch.cuda.synchronize()
for name in ("small",):
    cfg = synth.LSTM_CONFIGS[name]
    m = LSTMIQ.from_config(cfg).cuda().train()
    b = synth.make_batch(cfg)
    t = m.forward_losses(b["feats"], b["answers"], b["alengths"], b["questions"])
    torch.cuda.synchronize()
    print("lstm", name, t[:4].tolist())
cfg = synth.CONFIGS["small"]
m = IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg)).cuda().train()
b = synth.make_batch(cfg)
t = m.forward_losses(b["feats"], b["answers"], b["posteriors"], b["questions"])
torch.cuda.synchronize()
print("transformer small", t[:4].tolist())
