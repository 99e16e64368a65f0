"""GPU bring-up check for b200iq_gemm: tcgen05 TF32 + SIMT paths vs torch fp64, all operand layouts."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from blt_vqg_b200 import ops, _lib
_lib.require_device()
torch.manual_seed(0)
dev = "cuda"
ok = True
def rel(a, b): return ((a.double() - b).norm() / b.norm()).item()
for (M, N, K) in [(128, 128, 32), (128, 128, 64), (128, 128, 512), (256, 384, 300), (1000, 200, 1024), (37, 50, 20), (10240, 512, 512), (512, 15000, 512)]:
    for a_t in (False, True):
        for b_t in (False, True):
            if (a_t and M % 4) or (b_t and N % 4) or (not a_t and K % 4) or (not b_t and K % 4):
                continue
            A = torch.randn(M, K, device=dev); B = torch.randn(N, K, device=dev)
            Aa = A.t().contiguous() if a_t else A
            Bb = B.t().contiguous() if b_t else B
            bias = torch.randn(N, device=dev); res = torch.randn(M, N, device=dev)
            ref = A.double() @ B.double().t()
            for simt in (True, False):
                try:
                    out = ops.gemm(Aa, Bb, a_t=a_t, b_t=b_t, simt=simt)
                    torch.cuda.synchronize()
                    e0 = rel(out, ref)
                    out2 = ops.gemm(Aa, Bb, a_t=a_t, b_t=b_t, bias=bias, residual=res, relu=True, simt=simt, alpha=0.5)
                    e1 = rel(out2, torch.relu(0.5 * ref + bias.double() + res.double()))
                    acc = torch.ones(M, N, device=dev)
                    ops.gemm(Aa, Bb, acc, a_t=a_t, b_t=b_t, accumulate=True, split_k=4, simt=simt)
                    e2 = rel(acc, ref + 1.0)
                    torch.cuda.synchronize()
                except Exception as ex:
                    print("EXC", M, N, K, a_t, b_t, simt, ex); ok = False; continue
                tol = 1e-5 if simt else 2e-3
                good = e0 < tol and e1 < tol and e2 < tol
                ok &= good
                print("%s M=%d N=%d K=%d a_t=%d b_t=%d simt=%d  err plain=%.2e epi=%.2e splitk=%.2e" % ("ok " if good else "BAD", M, N, K, a_t, b_t, simt, e0, e1, e2), flush=True)
# timing
for (M, N, K, a_t, b_t, sk) in [(10752, 1536, 512, False, False, 1), (10752, 512, 2048, False, False, 1), (10752, 2048, 512, False, False, 1), (10240, 15000, 512, False, False, 1), (10752, 512, 512, False, True, 1), (512, 2048, 10752, True, True, 8), (8192, 8192, 8192, False, False, 1)]:
    A = torch.randn((K, M) if a_t else (M, K), device=dev); B = torch.randn((K, N) if b_t else (N, K), device=dev)
    out = torch.zeros(M, N, device=dev)
    for _ in range(3): ops.gemm(A, B, out, a_t=a_t, b_t=b_t, accumulate=sk > 1, split_k=sk)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): ops.gemm(A, B, out, a_t=a_t, b_t=b_t, accumulate=sk > 1, split_k=sk)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    torch.backends.cuda.matmul.allow_tf32 = True
    Am = A.t() if a_t else A; Bm = B if b_t else B.t()
    for _ in range(3): torch.matmul(Am, Bm)
    e0.record()
    for _ in range(10): torch.matmul(Am, Bm)
    e1.record(); torch.cuda.synchronize()
    ms2 = e0.elapsed_time(e1) / 10
    print("time M=%d N=%d K=%d a_t=%d b_t=%d sk=%d: %.3f ms %.1f TF/s | cublas tf32 %.3f ms %.1f TF/s" % (M, N, K, a_t, b_t, sk, ms, 2e-9 * M * N * K / ms, ms2, 2e-9 * M * N * K / ms2), flush=True)
print("ALL OK" if ok else "FAILURES")
sys.exit(0 if ok else 1)
