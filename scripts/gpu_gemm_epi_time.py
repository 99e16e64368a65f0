"""Time the cfg3 step's b200iq_gemm_f16 variants one by one: every case is captured in a CUDA graph of REPS launches over rotating
operand sets (no L2 reuse between launches, no host pacing) and replayed.   python scripts/gpu_gemm_epi_time.py [case-substring]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from blt_vqg_b200 import _lib, ops  # noqa: E402

if os.environ.get("B200IQ_LIB"):   # A/B against another build of the library
    _lib.LIB_PATH = os.environ["B200IQ_LIB"]
_lib.require_device()
_lib.lib().b200iq_set_tf32_rounding(0)
dev = "cuda"
NSET, REPS = 4, 24
only = sys.argv[1] if len(sys.argv) > 1 else ""


def planes(M, K, lo=True):
    hi = (torch.randn(M, K, device=dev) * 0.5).half()
    return ops.Act(hi=hi, lo=(torch.randn(M, K, device=dev) * 2e-4).half() if lo else None)


def out_planes(M, N, lo=True):
    return ops.Act(hi=torch.empty(M, N, device=dev, dtype=torch.float16), lo=torch.empty(M, N, device=dev, dtype=torch.float16) if lo else None)


# name, M, N, K, terms, kind
CASES = [
    ("fwd3 qkv        fp32 out", 10240, 1536, 512, 3, "f32"),
    ("fwd3 out-proj   fp32 out", 10240, 512, 512, 3, "f32"),
    ("fwd3 ffn1 bias relu drop -> planes", 10240, 2048, 512, 3, "drop_planes"),
    ("fwd3 ffn1 bias relu      -> planes", 10240, 2048, 512, 3, "relu_planes"),
    ("fwd3 ffn2 bias relu drop -> fp32", 10240, 512, 2048, 3, "drop_f32"),
    ("fwd3 ffn2 bias relu      -> fp32", 10240, 512, 2048, 3, "relu_f32"),
    ("dx1 ffn2 gate -> hi plane", 10240, 2048, 512, 1, "gate"),
    ("dx1 ffn2 plain -> hi plane", 10240, 2048, 512, 1, "dx_plane"),
    ("dx1 ffn1 fp32 out", 10240, 512, 2048, 1, "dx"),
    ("dx1 qkv  fp32 out", 10240, 512, 1536, 1, "dx"),
    ("dx1 out-proj fp32 out", 10240, 512, 512, 1, "dx"),
    ("dw1 512x512 sk18", 512, 512, 10240, 1, "dw18"),
    ("dw1 2048x512 sk4", 2048, 512, 10240, 1, "dw4"),
    ("vocab logits 1T bias", 1280, 15000, 512, 1, "f32b"),
    ("vocab dx 1T sk3", 1280, 512, 15000, 1, "vdx"),
]
for name, M, N, K, terms, kind in CASES:
    if only and only not in name:
        continue
    a_t = kind.startswith("dw")
    b_t = kind in ("gate", "gate_bits", "dx_plane", "dx", "vdx") or a_t
    lo = terms == 3
    As = [planes(K, M, lo) if a_t else planes(M, K, lo) for _ in range(NSET)]
    Bs = [planes(K, N, lo) if b_t else planes(N, K, lo) for _ in range(NSET)]
    bias = torch.randn(N, device=dev)
    Cs = [torch.zeros(M, N, device=dev) for _ in range(NSET)] if kind in ("f32", "drop_f32", "relu_f32", "dx", "dw18", "dw4", "f32b", "vdx") else None
    Ps = [out_planes(M, N, lo=kind.startswith(("drop_planes", "relu_planes"))) for _ in range(NSET)] if kind in ("drop_planes", "drop_planes_bits", "relu_planes", "gate", "gate_bits", "dx_plane") else None
    Gs = [ops.Act(hi=torch.relu(torch.randn(M, N, device=dev)).half()) for _ in range(NSET)] if kind in ("gate", "gate_bits") else None

    def run(i):
        kw = dict(a_t=a_t, b_t=b_t, terms=terms)
        if kind in ("drop_planes", "relu_planes", "drop_planes_bits"):
            kw.update(planes=Ps[i], bias=bias, relu=True)
        if kind in ("drop_f32", "relu_f32"):
            kw.update(bias=bias, relu=True)
        if kind in ("drop_planes", "drop_f32"):
            kw.update(dropout=(0.1, 5, 1 << 40))
        if kind == "gate":
            kw.update(planes=Ps[i], relu_mask=Gs[i], alpha=1.1)
        if kind == "dx_plane":
            kw.update(planes=Ps[i])
        if kind == "dw18":
            kw.update(accumulate=True, split_k=18)
        if kind == "dw4":
            kw.update(accumulate=True, split_k=4)
        if kind == "vdx":
            kw.update(accumulate=True, split_k=3)
        if kind == "f32b":
            kw.update(bias=bias)
        ops.gemm2(As[i], Bs[i], Cs[i] if Cs is not None else None, **kw)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st), _lib.pinned_stream():
        for i in range(NSET):
            run(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for r in range(REPS):
                run(r % NSET)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / (5 * REPS)
    fl = 2.0 * M * N * K
    print("%-38s M=%-6d N=%-6d K=%-6d terms=%d  %7.1f us  %6.0f TFLOP/s algorithmic (%6.0f hw)" % (name, M, N, K, terms, us, fl / us * 1e-6, fl * terms / us * 1e-6), flush=True)
    del As, Bs, Cs, Ps, Gs
