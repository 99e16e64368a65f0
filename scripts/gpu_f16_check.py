"""GPU check of the fp16-plane kernels: b200iq_gemm_f16 (every layout / term count / epilogue), the plane-writing row kernels
and the precise attention mode, each against an fp64 torch computation on the same (already rounded) operands.
Each case runs in its own process (a bad descriptor would poison the CUDA context), with a timeout.

    python scripts/gpu_f16_check.py            # all cases
    python scripts/gpu_f16_check.py --case gemm_kk1
"""
import argparse
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def planes_of(x, lo=True, ld=None):
    import torch
    M, K = x.shape
    ld = ld or (K + 7) // 8 * 8
    hi = torch.zeros(M, ld, device=x.device, dtype=torch.float16)
    hi[:, :K] = x.half()
    l = None
    if lo:
        l = torch.zeros(M, ld, device=x.device, dtype=torch.float16)
        l[:, :K] = (x - hi[:, :K].float()).half()
    return hi[:, :K], (l[:, :K] if lo else None)


def val(a):
    """the fp64 value an operand's planes represent"""
    v = a.hi.double()
    if a.lo is not None:
        v = v + a.lo.double()
    return v


def relerr(got, want):
    return float((got.double() - want).norm() / (want.norm() + 1e-300))


def time_call(fn, reps=20):
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def gemm_case(layout, terms):
    import torch
    from blt_vqg_b200 import ops
    torch.manual_seed(0)
    dev = "cuda"
    shapes = [(10240, 512, 512), (10240, 1536, 512), (10240, 2048, 512), (10240, 512, 2048), (512, 512, 2048), (2560, 1024, 512),
              (10240, 512, 1536), (300, 200, 304), (129, 136, 72), (1280, 15000, 512), (10240, 304, 512)]
    if layout == "mnmn":   # dW = dY^T X: M = N_out, N = K_in, reduction over tokens
        shapes = [(512, 512, 10240), (2048, 512, 10240), (512, 2048, 10240), (1536, 512, 10240), (512, 304, 10240), (15000, 512, 1280), (200, 136, 300)]
    if layout == "kmn":    # dX = dY W
        shapes = [(10240, 512, 512), (10240, 2048, 512), (10240, 512, 2048), (10240, 512, 1536), (1280, 512, 15000), (10240, 304, 512), (300, 200, 304)]
    ok = True
    for (M, N, K) in shapes:
        a_t, b_t = layout == "mnmn", layout in ("mnmn", "kmn")
        A = torch.randn(K, M, device=dev) if a_t else torch.randn(M, K, device=dev)
        B = torch.randn(K, N, device=dev) if b_t else torch.randn(N, K, device=dev)
        B *= 0.05
        ah, al = planes_of(A, lo=terms == 3)
        bh, bl = planes_of(B, lo=terms == 3)
        a, b = ops.Act(hi=ah, lo=al), ops.Act(hi=bh, lo=bl)
        av, bv = val(a), val(b)
        want = (av.t() if a_t else av) @ (bv if b_t else bv.t())
        if terms == 3:   # the lo.lo term is dropped by design
            want = want - (al.double() @ bl.double().t())
        out = torch.full((M, N), float("nan"), device=dev)
        ops.gemm2(a, b, out, a_t=a_t, b_t=b_t, terms=terms)
        torch.cuda.synchronize()
        e = relerr(out, want)
        ms = time_call(lambda: ops.gemm2(a, b, out, a_t=a_t, b_t=b_t, terms=terms))
        tf = 2.0 * M * N * K / ms * 1e-9
        tol = 1e-6 + 1.5e-9 * K * terms   # fp32 accumulation in TMEM truncates: the error grows linearly with the number of MMAs
        good = e < tol
        ok = ok and good
        print("  %-5s terms=%d M=%-6d N=%-6d K=%-6d err %.2e  %.1f us  %.0f TFLOP/s algorithmic (%.0f hw)  %s" % (
            layout, terms, M, N, K, e, ms * 1e3, tf, tf * terms, "ok" if good else "FAIL"), flush=True)
        if layout == "mnmn" and terms == 1:
            # split-K accumulate into fp32 (the weight-gradient form)
            for sk in (4, 16):
                acc = torch.ones(M, N, device=dev)
                ops.gemm2(a, b, acc, a_t=True, b_t=True, accumulate=True, split_k=sk, terms=1)
                e = relerr(acc, want + 1.0)
                ms = time_call(lambda: ops.gemm2(a, b, acc, a_t=True, b_t=True, accumulate=True, split_k=sk, terms=1))
                good = e < tol
                ok = ok and good
                print("        split_k=%d accumulate err %.2e  %.1f us  %.0f TFLOP/s %s" % (sk, e, ms * 1e3, 2.0 * M * N * K / ms * 1e-9, "ok" if good else "FAIL"), flush=True)
    return ok


def epilogue_case():
    import torch
    from blt_vqg_b200 import _lib, ops
    torch.manual_seed(1)
    dev = "cuda"
    ok = True
    _lib.lib().b200iq_set_tf32_rounding(0)   # the standalone dropout kernel used as the mask reference must not round
    for (M, N, K) in [(10240, 2048, 512), (640, 512, 2048), (300, 200, 304), (100, 64, 64), (77, 203, 72), (300, 100, 64)]:   # the last two: generic epilogue
        A, B = torch.randn(M, K, device=dev), torch.randn(N, K, device=dev) * 0.05
        bias = torch.randn(N, device=dev)
        res = torch.randn(M, N, device=dev)
        gate_src = torch.relu(torch.randn(M, N, device=dev))
        gh, _ = planes_of(gate_src, lo=False, ld=N if N % 8 else None)
        for terms in (1, 3):
            a, b = ops.Act(*(None,) + planes_of(A, lo=terms == 3)), ops.Act(*(None,) + planes_of(B, lo=terms == 3))
            base = val(a) @ val(b).t()
            if terms == 3:
                base = base - a.lo.double() @ b.lo.double().t()
            # bias + relu -> planes (hi, lo)
            pl = ops.Act(*(None,) + planes_of(torch.zeros(M, N, device=dev)))
            ops.gemm2(a, b, planes=pl, bias=bias, relu=True, terms=terms)
            want = torch.relu(base + bias.double())
            e1 = relerr(val(pl), want)
            ehi = relerr(pl.hi, want)
            # bias + relu -> fp32 AND planes
            out = torch.zeros(M, N, device=dev)
            pl2 = ops.Act(*(None,) + planes_of(torch.zeros(M, N, device=dev), lo=False))
            ops.gemm2(a, b, out, planes=pl2, bias=bias, relu=True, terms=terms)
            e2 = max(relerr(out, want), relerr(pl2.hi, pl.hi.double()))
            # residual -> fp32
            out3 = torch.zeros(M, N, device=dev)
            ops.gemm2(a, b, out3, residual=res, alpha=0.5, terms=terms)
            e3 = relerr(out3, 0.5 * base + res.double())
            # gate (fp16 plane) * alpha -> hi plane
            pl4 = ops.Act(*(None,) + planes_of(torch.zeros(M, N, device=dev), lo=False))
            ops.gemm2(a, b, planes=pl4, relu_mask=ops.Act(hi=gh), alpha=1.25, terms=terms)
            want4 = torch.where(gh.double() > 0, 1.25 * base, torch.zeros_like(base))
            e4 = relerr(pl4.hi, want4)
            # dropout epilogue = the standalone dropout kernel's mask
            e5 = 0.0
            if N % 8 == 0:
                pl5 = ops.Act(*(None,) + planes_of(torch.zeros(M, N, device=dev)))
                ops.gemm2(a, b, planes=pl5, bias=bias, relu=True, dropout=(0.1, 77, 5 << 40), terms=terms)
                ref = want.float().contiguous()
                ops.dropout_(ref, 0.1, 77, 5 << 40)
                e5 = relerr(val(pl5), ref.double())
                o6 = torch.zeros(M, N, device=dev)
                ops.gemm2(a, b, o6, bias=bias, relu=True, dropout=(0.1, 77, 5 << 40), terms=terms)
                e5 = max(e5, relerr(o6, ref.double()))
            tol = 2e-6 + 1.5e-9 * K * terms
            good = e1 < tol and ehi < 6e-4 and e2 < tol and e3 < tol and e4 < 6e-4 and e5 < tol
            ok = ok and good
            print("  epi M=%d N=%d K=%d terms=%d: planes %.1e (hi only %.1e) f32+planes %.1e residual %.1e gate %.1e dropout %.1e %s" % (
                M, N, K, terms, e1, ehi, e2, e3, e4, e5, "ok" if good else "FAIL"), flush=True)
    return ok


def rows_case():
    import torch
    import torch.nn.functional as F
    from blt_vqg_b200 import ops
    torch.manual_seed(2)
    dev = "cuda"
    ok = True

    def chk(name, e, tol):
        nonlocal ok
        good = e < tol
        ok = ok and good
        print("  %-28s err %.2e %s" % (name, e, "ok" if good else "FAIL"), flush=True)
    with ops.no_rounding():
        for (M, H) in [(10240, 512), (77, 64), (300, 300)]:
            x, add = torch.randn(M, H, device=dev), torch.randn(M, H, device=dev)
            g, b = torch.randn(H, device=dev), torch.randn(H, device=dev)
            mean, rstd = torch.zeros(M, device=dev), torch.zeros(M, device=dev)
            ld = (H + 7) // 8 * 8
            act = ops.Act(hi=torch.zeros(M, ld, device=dev, dtype=torch.float16)[:, :H], lo=torch.zeros(M, ld, device=dev, dtype=torch.float16)[:, :H])
            y = torch.zeros(M, H, device=dev)
            ops.layernorm_fwd_act(x, g, b, mean, rstd, y=y, act=act)
            want = F.layer_norm(x.double(), (H,), g.double(), b.double())
            chk("ln_fwd planes H=%d" % H, max(relerr(val(act), want), relerr(y, want)), 2e-6)
            s = torch.zeros(M, H, device=dev)
            ops.layernorm_fwd_act(x, g, b, mean, rstd, act=act, add=add, sum_out=s)
            want = F.layer_norm((x + add).double(), (H,), g.double(), b.double())
            chk("add+ln_fwd planes H=%d" % H, max(relerr(val(act), want), relerr(s, (x + add).double())), 2e-6)
            # backward with the fp16 copy of dx
            dy = torch.randn(M, H, device=dev)
            dx, dxh = torch.randn(M, H, device=dev), torch.zeros(M, H, device=dev, dtype=torch.float16)
            dx0 = dx.clone()
            dg, db = torch.zeros(H, device=dev), torch.zeros(H, device=dev)
            xs = (x + add)
            ops.layernorm_bwd_act(dy, xs, g, mean, rstd, dx, dg, db, True, dx_hi=dxh)
            xr = xs.double().requires_grad_(True)
            gr = g.double().requires_grad_(True)
            F.layer_norm(xr, (H,), gr, b.double()).backward(dy.double())
            chk("ln_bwd + hi H=%d" % H, max(relerr(dx, xr.grad + dx0.double()), relerr(dg, gr.grad), relerr(dxh, dx.double()) / 300), 3e-6)
        for (M, N) in [(10240, 512), (100, 60)]:
            dy, yv = torch.randn(M, N, device=dev), torch.relu(torch.randn(M, N, device=dev))
            hi = torch.zeros(M, N, device=dev, dtype=torch.float16)
            dbias = torch.zeros(N, device=dev)
            ops.relu_bwd_colsum_act(dy, yv, ops.Act(hi=hi), dbias, scale=1.1)
            want = torch.where(yv > 0, dy.double() * 1.1, torch.zeros_like(dy, dtype=torch.float64))
            chk("relu_bwd_colsum_f16 N=%d" % N, max(relerr(hi, want) / 300, relerr(dbias, want.sum(0))), 3e-6)
            cs = torch.zeros(N, device=dev)
            ops.colsum_add_act(ops.Act(hi=hi), cs)
            chk("colsum_add_f16 N=%d" % N, relerr(cs, hi.double().sum(0)), 3e-6)
        for (R, V) in [(1280, 15000), (256, 32000), (50, 200)]:
            lg = torch.randn(R, V, device=dev) * 3
            tg = torch.randint(0, V, (R,), device=dev, dtype=torch.int32)
            tg[::7] = 0
            cnt = torch.tensor([float((tg != 0).sum())], device=dev)
            ls = torch.zeros(1, device=dev)
            gh = torch.zeros(R, V, device=dev, dtype=torch.float16)
            S = 4096.0
            ops.ce_rows_act(lg, tg, 0, cnt, S, ls[0:1], ops.Act(hi=gh))
            lr = lg.double().requires_grad_(True)
            loss = F.cross_entropy(lr, tg.long(), ignore_index=0)
            loss.backward()
            chk("ce_rows_f16 V=%d" % V, max(abs(float(ls) - float(loss)) / float(loss), relerr(gh, lr.grad * S) / 300), 3e-6)
        tab = torch.randn(1000, 300, device=dev)
        ids = torch.randint(0, 1000, (777,), device=dev, dtype=torch.int32)
        act = ops.Act(hi=torch.zeros(777, 304, device=dev, dtype=torch.float16)[:, :300], lo=torch.zeros(777, 304, device=dev, dtype=torch.float16)[:, :300])
        ops.embed_gather_act(ids, tab, act)
        chk("embed_gather_f16", relerr(val(act), tab[ids.long()].double()), 2e-6)
        flat = torch.randn(1 << 22, device=dev)
        hi, lo = torch.zeros(1 << 22, device=dev, dtype=torch.float16), torch.zeros(1 << 22, device=dev, dtype=torch.float16)
        ops.split_f16(flat, hi, lo)
        chk("split_f16 flat", relerr(hi.double() + lo.double(), flat.double()), 2e-6)
        ms = time_call(lambda: ops.split_f16(flat, hi, lo))
        print("  split_f16 4M floats: %.1f us = %.0f GB/s" % (ms * 1e3, (1 << 22) * 8 / ms * 1e-6))
    return ok


def attention_case():
    import torch
    from blt_vqg_b200 import ops
    torch.manual_seed(3)
    dev = "cuda"
    ok = True
    with ops.no_rounding():
        for (B, nh, nq, nk, d, causal) in [(512, 8, 20, 20, 64, True), (512, 8, 21, 21, 64, False), (512, 8, 20, 5, 64, False), (512, 8, 5, 5, 64, False),
                                           (64, 8, 64, 64, 64, True), (64, 8, 65, 65, 64, False), (8, 4, 12, 12, 16, True)]:
            H = nh * d
            # a common component in the keys (what LayerNorm + a bias-free projection give at initialisation)
            q = torch.randn(B * nq, H, device=dev)
            k = torch.randn(B * nk, H, device=dev) + 2.0
            v = torch.randn(B * nk, H, device=dev)
            ids = torch.randint(1, 50, (B, nk), device=dev, dtype=torch.int32)
            ids[:, nk - 1] = 0
            if nk > 3:
                ids[::2, nk - 2] = 0
            qd, kd, vd = (t.double().requires_grad_(True) for t in (q, k, v))
            Q = qd.view(B, nq, nh, d).permute(0, 2, 1, 3) * d ** -0.5
            K = kd.view(B, nk, nh, d).permute(0, 2, 1, 3)
            V = vd.view(B, nk, nh, d).permute(0, 2, 1, 3)
            lg = Q @ K.transpose(-1, -2)
            mask = (ids == 0)[:, None, None, :].expand(B, nh, nq, nk)
            if causal:
                mask = mask | torch.triu(torch.ones(nq, nk, dtype=torch.bool, device=dev), 1)[None, None]
            lg = lg.masked_fill(mask, -1e18)
            want = (torch.softmax(lg, -1) @ V).permute(0, 2, 1, 3).reshape(B * nq, H)
            act = ops.Act(hi=torch.zeros(B * nq, H, device=dev, dtype=torch.float16), lo=torch.zeros(B * nq, H, device=dev, dtype=torch.float16))
            ops.attention_fwd(q, k, v, act, ids, 0, causal, B, nh, nq, nk, d)
            e_f = relerr(val(act), want.detach())
            ops.PRECISE_ATTENTION = 1
            o32 = torch.zeros(B * nq, H, device=dev)
            ops.attention_fwd(q, k, v, o32, ids, 0, causal, B, nh, nq, nk, d)
            e_f32 = relerr(o32, want.detach())
            dout = torch.randn(B * nq, H, device=dev) * 1e-2
            want.backward(dout.double())
            dq, dk, dv = (torch.zeros_like(t, dtype=torch.float16) for t in (q, k, v))
            ops.attention_bwd(q, k, v, dout, dq, dk, dv, ids, 0, causal, B, nh, nq, nk, d)
            dq32, dk32, dv32 = (torch.zeros_like(t) for t in (q, k, v))
            ops.attention_bwd(q, k, v, dout, dq32, dk32, dv32, ids, 0, causal, B, nh, nq, nk, d)
            ops.PRECISE_ATTENTION = 0
            eb = [relerr(a, w.grad) for a, w in ((dq32, qd), (dk32, kd), (dv32, vd))]
            eh = [relerr(a, w.grad) for a, w in ((dq, qd), (dk, kd), (dv, vd))]
            ms_f = time_call(lambda: ops.attention_fwd(q, k, v, act, ids, 0, causal, B, nh, nq, nk, d))
            ms_b = time_call(lambda: ops.attention_bwd(q, k, v, dout, dq, dk, dv, ids, 0, causal, B, nh, nq, nk, d))
            good = e_f < 3e-6 and e_f32 < 3e-6 and max(eb) < 1.5e-3 and max(eh) < 2e-3
            ok = ok and good
            print("  att B=%d nh=%d nq=%d nk=%d d=%d causal=%d: fwd planes %.1e f32 %.1e | bwd f32 dq %.1e dk %.1e dv %.1e | f16 %.1e %.1e %.1e | %.1f / %.1f us %s" % (
                B, nh, nq, nk, d, causal, e_f, e_f32, eb[0], eb[1], eb[2], eh[0], eh[1], eh[2], ms_f * 1e3, ms_b * 1e3, "ok" if good else "FAIL"), flush=True)
    return ok


CASES = {
    "gemm_kk1": lambda: gemm_case("kk", 1), "gemm_kk3": lambda: gemm_case("kk", 3), "gemm_kmn": lambda: gemm_case("kmn", 1),
    "gemm_mnmn": lambda: gemm_case("mnmn", 1), "epilogue": epilogue_case, "rows": rows_case, "attention": attention_case,
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="")
    args = ap.parse_args()
    if args.case:
        from blt_vqg_b200 import _lib
        _lib.require_device()
        ok = CASES[args.case]()
        print("CASE %s: %s" % (args.case, "PASS" if ok else "FAIL"), flush=True)
        return 0 if ok else 1
    bad = []
    for name in CASES:
        print("== %s" % name, flush=True)
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", name], timeout=300)
            if r.returncode != 0:
                bad.append(name)
        except subprocess.TimeoutExpired:
            print("CASE %s: TIMEOUT" % name, flush=True)
            bad.append(name)
    print("SUMMARY: %d/%d cases pass%s" % (len(CASES) - len(bad), len(CASES), (" ; failing: " + ", ".join(bad)) if bad else ""))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
