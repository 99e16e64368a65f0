"""do two GEMMs on two streams overlap?  dX (10240x512x2048, 80 pair tiles = 2 waves, second wave 6 pairs) with the dW
of the same block (2048x512x10240 split-K 4 = 64 pair units):  python scripts/gpu_gemm_overlap.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import _lib, ops
_lib.require_device()
M, N, K = 10240, 512, 2048
NS = 6
dy = [torch.randn(M, K, device="cuda") for _ in range(NS)]      # dh [M, F]
w = [torch.randn(K, N, device="cuda") for _ in range(NS)]        # W1 [F, H] used b_t
dx = [torch.zeros(M, N, device="cuda") for _ in range(NS)]
x = [torch.randn(M, N, device="cuda") for _ in range(NS)]
gw = [torch.zeros(K, N, device="cuda") for _ in range(NS)]
side = torch.cuda.Stream()
ev = torch.cuda.Event()
def dX(i): ops.gemm(dy[i], w[i], dx[i], b_t=True)
def dW(i): ops.gemm(dy[i], x[i], gw[i], a_t=True, b_t=True, accumulate=True, split_k=4)
def seq(i): dX(i); dW(i)
def conc_main_first(i):
    ev.record(); dX(i); side.wait_event(ev)
    with torch.cuda.stream(side): dW(i)
    torch.cuda.current_stream().wait_stream(side)
def conc_side_first(i):
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side): dW(i)
    dX(i)
    torch.cuda.current_stream().wait_stream(side)
for name, fn in (("dX alone", dX), ("dW alone", dW), ("sequential", seq), ("two streams, dX issued first", conc_main_first),
                 ("two streams, dW issued first", conc_side_first)):
    for i in range(NS): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(10):
        for i in range(NS): fn(i)
    e1.record(); torch.cuda.synchronize()
    print("%-32s %7.1f us" % (name, 1e3 * e0.elapsed_time(e1) / (10 * NS)))
