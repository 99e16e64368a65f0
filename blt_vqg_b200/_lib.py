"""ctypes loader for libb200iq.so — the only way Python reaches the kernels.

There is no fallback: if the shared library is missing or the device is not a B200
(compute capability 10.0) the ops raise.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200iq.so")
_lib = None


class B200iqError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200iqError(
                "libb200iq.so not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                "there is no CPU fallback")
        _lib = C.CDLL(LIB_PATH)
        _lib.b200iq_last_error.restype = C.c_char_p
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().b200iq_last_error().decode()
        raise B200iqError("%s failed (rc=%d): %s" % (what, rc, msg))


def require_device():
    if not torch.cuda.is_available():
        raise B200iqError("no CUDA device: libb200iq has no CPU fallback")
    if not lib().b200iq_device_supported():
        raise B200iqError("libb200iq only contains sm_100a code (B200, compute capability 10.0)")


_STREAM = None  # cached c_void_p of the launching stream while a `pinned_stream()` block is active


def stream_ptr():
    if _STREAM is not None:
        return _STREAM
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class pinned_stream:
    """`with pinned_stream():` looks torch's current stream up once for the whole block (a training step issues
    ~650 launches; the per-launch lookup was a measurable share of the host time).  Do not switch streams inside."""

    def __enter__(self):
        global _STREAM
        self.prev = _STREAM
        _STREAM = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def __exit__(self, *a):
        global _STREAM
        _STREAM = self.prev


def ptr(t):
    """device pointer of a tensor (or NULL)"""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def fptr(t):
    if t is None:
        return C.cast(None, C.POINTER(C.c_float))
    assert t.dtype == torch.float32 and t.is_cuda, (t.dtype, t.device)
    return C.cast(C.c_void_p(t.data_ptr()), C.POINTER(C.c_float))


class GemmParams(C.Structure):
    _fields_ = [
        ("A", C.c_void_p), ("B", C.c_void_p), ("C", C.c_void_p), ("bias", C.c_void_p), ("residual", C.c_void_p),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("lda", C.c_int), ("ldb", C.c_int), ("ldc", C.c_int), ("ldr", C.c_int),
        ("flags", C.c_int), ("split_k", C.c_int), ("alpha", C.c_float),
        ("drop_p", C.c_float), ("drop_seed", C.c_uint64), ("drop_offset", C.c_uint64),
    ]


GEMM_A_MN, GEMM_B_MN, GEMM_BIAS, GEMM_RELU = 1, 2, 4, 8
GEMM_RESIDUAL, GEMM_ACCUM, GEMM_SIMT, GEMM_RELU_MASK, GEMM_FP32_OUT, GEMM_DROPOUT = 16, 32, 64, 128, 256, 512
