"""Thin Python wrappers: one function per C-ABI entry point of libb200iq.so.

Every function launches on torch's current CUDA stream and returns immediately.
Tensors are fp32 CUDA tensors with unit inner stride; token ids are int32.
"""
import ctypes as C
import os

import torch

from . import _lib as L

# global switch: exact-fp32 CUDA-core GEMMs instead of tcgen05 TF32 (parity debugging, greedy decode)
_FORCE_SIMT = os.environ.get("B200IQ_GEMM", "") == "simt"


class gemm_mode:
    """context manager: `with gemm_mode(simt=True): ...` forces the exact-fp32 GEMM path"""

    def __init__(self, simt):
        self.simt = simt

    def __enter__(self):
        global _FORCE_SIMT
        self.prev = _FORCE_SIMT
        _FORCE_SIMT = self.simt

    def __exit__(self, *a):
        global _FORCE_SIMT
        _FORCE_SIMT = self.prev


def _ld(t):
    assert t.dim() == 2 and t.stride(1) == 1, (t.shape, t.stride())
    return t.stride(0) if t.shape[0] > 1 else max(t.stride(0), t.shape[1])


def gemm(a, b, out=None, *, a_t=False, b_t=False, bias=None, residual=None, relu=False, relu_mask=None,
         accumulate=False, split_k=1, alpha=1.0, simt=None):
    """out[M,N] = epi(alpha * A @ B^T); a is [M,K] ([K,M] if a_t), b is [N,K] ([K,N] if b_t).
    mirrors torch.nn.functional.linear(a, b, bias) (+ fused relu / residual / relu-backward mask)."""
    M, K = (a.shape[1], a.shape[0]) if a_t else a.shape
    N, Kb = (b.shape[1], b.shape[0]) if b_t else b.shape
    assert K == Kb, (a.shape, b.shape, a_t, b_t)
    if out is None:
        assert not accumulate
        out = torch.empty(M, N, device=a.device, dtype=torch.float32)
    assert out.shape == (M, N)
    flags = 0
    if a_t: flags |= L.GEMM_A_MN
    if b_t: flags |= L.GEMM_B_MN
    if bias is not None: flags |= L.GEMM_BIAS
    if relu: flags |= L.GEMM_RELU
    if residual is not None: flags |= L.GEMM_RESIDUAL
    if relu_mask is not None:
        flags |= L.GEMM_RELU_MASK
        residual = relu_mask
    if accumulate: flags |= L.GEMM_ACCUM
    if simt if simt is not None else _FORCE_SIMT: flags |= L.GEMM_SIMT
    p = L.GemmParams(L.fptr(a), L.fptr(b), L.fptr(out), L.fptr(bias), L.fptr(residual), M, N, K,
                     _ld(a), _ld(b), _ld(out), _ld(residual) if residual is not None else 0,
                     flags, split_k, float(alpha))
    L.check(L.lib().b200iq_gemm(C.byref(p), L.stream_ptr()), "b200iq_gemm")
    return out
