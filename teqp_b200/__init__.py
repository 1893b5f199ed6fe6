"""teqp_b200 -- host-side mirror of teqp's Python surface (`teqp.make_model(json)` -> model with `get_Ar..` methods,
/root/reference/teqp/__init__.py:42-49, /root/reference/interface/pybind11_wrapper.cpp:496-573) over the C ABI of
libteqpb200.so (include/teqpb200.h).  All evaluation happens in the CUDA kernels of that library; if the library cannot be
loaded, importing the model factory fails loudly -- there is no CPU evaluation path in this package.

Scalar methods keep the reference's names and argument order (T, rho, molefrac).  Batched methods take arrays:
numpy arrays (host; the library streams them through the GPU(s)) or torch CUDA tensors (device-resident, zero-copy,
enqueued on torch's current stream).
"""
import ctypes as C
import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libteqpb200.so")
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_lib = None

DEVICE_PTRS = 1
BROADCAST_X = 2


class TeqpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("[%d] %s" % (code, msg))
        self.code = code
        self.msg = msg


def lib():
    """Loads libteqpb200.so (built in-tree by __graft_entry__.build() / teqp_b200/csrc/Makefile)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libteqpb200.so is not built (run `make -C teqp_b200/csrc`); teqp_b200 has no CPU fallback")
        _lib = C.CDLL(LIB_PATH)
        _lib.teqpb200_host_alloc.restype = C.c_void_p
        _lib.teqpb200_host_alloc.argtypes = [C.c_ulonglong]
        _lib.teqpb200_host_free.argtypes = [C.c_void_p]
        _lib.teqpb200_kernel_launches.restype = C.c_longlong
        _lib.teqpb200_set_stream.argtypes = [C.c_void_p]
    return _lib


def _check(rc, err):
    if rc != 0:
        raise TeqpError(rc, err.value.decode(errors="replace"))


def _is_torch(a):
    return type(a).__module__.startswith("torch")


class _Arg:
    """Normalises one array argument to (pointer, keep-alive object)."""

    def __init__(self, a, device):
        if device:
            import torch
            assert a.is_cuda and a.dtype == torch.float64 and a.is_contiguous(), "device arrays must be contiguous float64 CUDA tensors"
            self.obj = a
            self.ptr = C.cast(C.c_void_p(a.data_ptr()), _dp)
        else:
            self.obj = np.ascontiguousarray(a, dtype=np.float64)
            self.ptr = self.obj.ctypes.data_as(_dp)


def pinned_empty(shape):
    """float64 numpy array backed by pinned host memory from the library (fast asynchronous H2D/D2H)."""
    n = int(np.prod(shape))
    p = lib().teqpb200_host_alloc(C.c_ulonglong(max(n, 1) * 8))
    if not p:
        raise MemoryError("teqpb200_host_alloc failed")
    buf = (C.c_double * n).from_address(p)
    arr = np.frombuffer(buf, dtype=np.float64).reshape(shape)
    _PINNED[arr.ctypes.data] = p
    return arr


_PINNED = {}


def pinned_free(arr):
    p = _PINNED.pop(arr.ctypes.data, None)
    if p:
        lib().teqpb200_host_free(C.c_void_p(p))


def kernel_launches():
    return int(lib().teqpb200_kernel_launches())


def set_devices(ids):
    err = C.create_string_buffer(512)
    arr = (C.c_int * len(ids))(*ids)
    _check(lib().teqpb200_set_devices(arr, C.c_int(len(ids)), err, C.c_int(512)), err)


def device_count():
    return int(lib().teqpb200_device_count())


from .model import Model, make_model  # noqa: E402
