"""ctypes binding of libldnn_b200.so (see include/ldnn_b200.h).

The library is plain CUDA with a C ABI: tensors cross as raw device pointers
(``tensor.data_ptr()``) plus the current CUDA stream, so there is no torch C++
ABI coupling and the in-tree ``.so`` is visible to the driver's "was native code
loaded" check.  On a GPU box a missing library is a hard error — there is no
silent eager fallback for the ops that have kernels.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(_PKG, "lib", "libldnn_b200.so")
MAX_RANKS = 8
SIGNAL_WORDS = 65536

_lib = None


class LdnnError(RuntimeError):
    pass


class ConvDesc(C.Structure):
    _fields_ = [("mode", C.c_int), ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("Cin", C.c_int),
                ("Cout", C.c_int), ("R", C.c_int), ("stride", C.c_int), ("pad", C.c_int), ("block_n", C.c_int),
                ("splits", C.c_int), ("accumulate", C.c_int), ("x", C.c_void_p), ("w", C.c_void_p),
                ("y", C.c_void_p), ("dw", C.c_void_p), ("addend", C.c_void_p), ("scratch", C.c_void_p),
                ("tickets", C.c_void_p), ("a_lbo", C.c_uint), ("a_sbo", C.c_uint), ("b_lbo", C.c_uint),
                ("b_sbo", C.c_uint), ("cluster", C.c_int), ("lite", C.c_int), ("atomic_out", C.c_int), ("bn_sums", C.c_void_p), ("fuse_out", C.c_void_p), ("fuse_y", C.c_void_p),
                ("fuse_mean", C.c_void_p), ("fuse_invstd", C.c_void_p), ("fuse_sums", C.c_void_p),
                ("fuse_relu", C.c_int), ("dtype", C.c_int)]


class Optim(C.Structure):
    _fields_ = [("kind", C.c_int), ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float),
                ("eps", C.c_float), ("momentum", C.c_float), ("weight_decay", C.c_float), ("bc1", C.c_float),
                ("bc2", C.c_float), ("grad_scale", C.c_float)]


class Comm(C.Structure):
    _fields_ = [("rank", C.c_int), ("world", C.c_int), ("emulate", C.c_int), ("blocks", C.c_int),
                ("epoch", C.c_uint), ("timeout_cycles", C.c_ulonglong), ("signal", C.c_void_p * MAX_RANKS),
                ("error_flag", C.c_void_p), ("debug_delay_block", C.c_int), ("debug_delay_cycles", C.c_uint)]


class Exchange(C.Structure):
    _fields_ = [("grads", C.c_void_p * MAX_RANKS), ("params", C.c_void_p * MAX_RANKS),
                ("recv", C.c_void_p * MAX_RANKS), ("m", C.c_void_p), ("v", C.c_void_p), ("shadow", C.c_void_p),
                ("n", C.c_longlong), ("weight", C.c_float * MAX_RANKS), ("self_coef", C.c_float), ("mode", C.c_int),
                ("opt", Optim), ("ll", C.c_void_p * MAX_RANKS), ("ll_phase4", C.c_longlong),
                ("ll_clear4", C.c_longlong), ("ll_epoch", C.c_uint)]


X_UPDATE, X_SELFMIX, X_CAST = 1, 2, 4


def available() -> bool:
    """True when the native library exists or can be built here (nvcc present)."""
    if os.path.exists(LIB_PATH):
        return True
    import shutil
    return shutil.which("nvcc") is not None or os.path.exists("/usr/local/cuda/bin/nvcc")


def _ensure_built() -> None:
    """In-tree build on first use (nvcc needed, ~1 min), safe when N ranks start together: one process builds under an
    exclusive file lock while the others wait; `build()` itself compares the source digest with the stamp (so a stale
    library is rebuilt after csrc/ changes) and installs the new .so with an atomic rename."""
    import fcntl
    from ..build import build as _build, needs_build
    if not needs_build():
        return
    os.makedirs(os.path.dirname(LIB_PATH), exist_ok=True)
    with open(os.path.join(os.path.dirname(LIB_PATH), ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if needs_build():                      # another rank may have finished while we waited
                _build(verbose=False)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def lib():
    global _lib
    if _lib is None:
        try:
            _ensure_built()
        except Exception as e:
            if not os.path.exists(LIB_PATH):
                raise LdnnError(f"{LIB_PATH} is missing and could not be built ({e}) — run "
                                f"`python -m ldnn_b200.build` (or __graft_entry__.build())")
        _lib = C.CDLL(LIB_PATH)
        _lib.ldnn_last_error.restype = C.c_char_p
        _lib.ldnn_launch_count.restype = C.c_ulonglong
        _lib.ldnn_ring_ll_slots.restype = C.c_longlong
        _lib.ldnn_ring_ll_slots.argtypes = [C.c_longlong, C.c_int, C.c_int]
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise LdnnError(f"libldnn_b200 error {rc}: {lib().ldnn_last_error().decode()}")


def stream_ptr() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def launch_count() -> int:
    return int(lib().ldnn_launch_count())


def make_optim(opt, step: int, lr: float, grad_scale: float = 1.0) -> Optim:
    o = Optim()
    o.kind = 0 if opt.kind == "adam" else 1
    o.lr, o.beta1, o.beta2, o.eps = lr, opt.beta1, opt.beta2, opt.eps
    o.momentum, o.weight_decay = opt.momentum, opt.weight_decay
    o.bc1 = 1.0 - opt.beta1 ** step
    o.bc2 = 1.0 - opt.beta2 ** step
    o.grad_scale = grad_scale
    return o
