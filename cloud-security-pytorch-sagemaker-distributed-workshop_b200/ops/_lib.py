"""ctypes binding of ``libb200ddp.so`` (C ABI declared in ``include/b200ddp.h``).

The library is built in-tree by ``ops/build.py`` (``__graft_entry__.build()``).  On a machine with a
GPU a missing library is a hard error — the native kernels are the product, there is no silent eager
fallback on CUDA tensors.  On the CPU dev box the library still loads (static cudart, driver API
resolved lazily), which is what lets ``build()`` double as an import check.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_uint8,
                    c_uint64, c_ulonglong, c_void_p)

import torch

from .build import LIB_PATH

MAX_RANKS = 8
MAX_COMM_BLOCKS = 160
SIGNAL_BYTES = 1 << 20
DTYPE_F32, DTYPE_BF16 = 0, 1
ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS = 0, 1, 2, 3
ALGO_NAMES = {ALGO_AUTO: "auto", ALGO_ONESHOT: "oneshot", ALGO_TWOSHOT: "twoshot", ALGO_NVLS: "nvls"}


class CommStruct(Structure):
    _fields_ = [("rank", c_int), ("world", c_int), ("peer", c_void_p * MAX_RANKS), ("mc", c_void_p),
                ("err_flag", c_void_p), ("timeout_ns", c_ulonglong)]


_lib = None


class NativeError(RuntimeError):
    pass


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> ctypes.CDLL:
    """Load (once) and return the native library; build it first if the .so is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        from .build import build
        build()
    L = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    L.b200_last_error.restype = c_char_p
    L.b200_bn_scratch_bytes.restype = c_size_t
    L.b200_bn_scratch_bytes.argtypes = [c_size_t, c_int]
    _lib = L
    return L


LAUNCHES = [0]          # native entry points that returned success (one kernel launch each, adam = 2)


def check(rc: int, what: str = "") -> None:
    LAUNCHES[0] += 1
    if rc != 0:
        msg = lib().b200_last_error().decode(errors="replace")
        raise NativeError("b200ddp native call %s failed (rc=%d): %s" % (what, rc, msg))


def ptr(t) -> c_void_p:
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(stream=None) -> c_void_p:
    """The launch stream as a ``cudaStream_t``.  Called once per native op: the current stream is read through torch's raw
    accessor (0.2 us) rather than by building a ``torch.cuda.Stream`` object (2 us) -- the eager path is host-bound."""
    if stream is not None:
        return c_void_p(stream.cuda_stream)
    if _raw_stream is not None:
        return c_void_p(_raw_stream(torch.cuda.current_device()))
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return DTYPE_F32
    if dt == torch.bfloat16:
        return DTYPE_BF16
    raise TypeError("b200ddp kernels take float32 or bfloat16, got %s" % dt)


def require_cuda(*tensors) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NativeError("native op called with a CPU tensor")


class DevicePtrTensorHolder:
    """Expose a raw device allocation (symmetric memory) to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr_value: int, nbytes: int, typestr: str = "|u1", shape=None):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape) if shape is not None else (nbytes,),
            "typestr": typestr,
            "data": (int(ptr_value), False),
            "version": 2,
            "strides": None,
        }


def tensor_from_ptr(ptr_value: int, nbytes: int, device: torch.device) -> torch.Tensor:
    """uint8 tensor aliasing ``nbytes`` of device memory at ``ptr_value`` (no ownership)."""
    holder = DevicePtrTensorHolder(ptr_value, nbytes)
    t = torch.as_tensor(holder, device=device)
    t._b200_holder = holder
    return t
