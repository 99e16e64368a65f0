"""HBM rate of the feature-cache kernels (b200iq_feature_gather / _store) on a COCO-sized cache (123287 x 2048), CUDA events on the
launching stream, rotating index sets so that no batch is served from L2:   python scripts/gpu_feature_cache_bw.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import _lib
from blt_vqg_b200.batch import FeatureCache
_lib.require_device()
N, F = 123287, 2048
peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
peak = peaks.get("hbm_gbs_burst") or peaks.get("hbm_gbs") or peaks.get("hbm_gbs_sustained")
print("peaks:", {k: v for k, v in peaks.items() if "hbm" in k})
for dtype in (torch.bfloat16, torch.float32):
    cache = FeatureCache(N, F, dtype=dtype)
    cache.rows.copy_(torch.randn(4096, F, device="cuda").abs().to(dtype).repeat(N // 4096 + 1, 1)[:N])
    es = 2 if dtype == torch.bfloat16 else 4
    for B in (512, 4096, 32768):
        NS = 8
        rows = [torch.randperm(N, device="cuda")[:B].int() for _ in range(NS)]
        outs = [torch.empty(B, F, device="cuda") for _ in range(NS)]
        for what in ("gather", "store"):
            fn = (lambda i: cache.gather(rows[i], outs[i])) if what == "gather" else (lambda i: cache.store(rows[i], outs[i]))
            for i in range(NS): fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for r in range(5):
                for i in range(NS): fn(i)
            e1.record(); torch.cuda.synchronize()
            us = 1e3 * e0.elapsed_time(e1) / (5 * NS)
            by = B * F * (es + 4) + B * 4
            print("%-7s %-8s B=%-6d %8.2f us  %7.1f GB/s  %5.1f %% of %s" % (what, str(dtype)[6:], B, us, by / us * 1e-3, 100 * by / us * 1e-3 / peak if peak else 0, peak))
    del cache
