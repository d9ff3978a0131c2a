"""ctypes binding of libdpvo_b200.so (the C ABI declared in include/dpvo_b200.h).

The product path has no CPU or eager fallback: if the shared library is missing
the import fails loudly with the build instruction.
"""
import ctypes
import os

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
# DPVO_B200_LIB: a differently built copy of the library (development: tools/build_prof.sh)
LIB_PATH = os.environ.get("DPVO_B200_LIB") or os.path.join(_PKG, "libdpvo_b200.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build the sm_100a extension first "
        "(python -c 'import __graft_entry__ as g; g.build()' or "
        "python deep-visual-odometry-slam_b200/build.py)")

lib = ctypes.CDLL(LIB_PATH)
lib.dpvo_last_error.restype = ctypes.c_char_p
lib.dpvo_abi_version.restype = ctypes.c_int

ABI_VERSION = 5
if lib.dpvo_abi_version() != ABI_VERSION:
    raise ImportError(f"libdpvo_b200.so ABI {lib.dpvo_abi_version()} != expected {ABI_VERSION}: rebuild")

F16, F32, F64 = 0, 1, 2
_DT = {torch.float16: F16, torch.float32: F32, torch.float64: F64}

c_i32 = ctypes.c_int32
c_i64 = ctypes.c_int64
c_f32 = ctypes.c_float
c_p = ctypes.c_void_p


class CorrLevel(ctypes.Structure):
    """Mirror of dpvo_corr_level."""
    _fields_ = [("fmap", c_p),
                ("sb", c_i64), ("sn", c_i64), ("sc", c_i64), ("sy", c_i64), ("sx", c_i64),
                ("N2", c_i32), ("H2", c_i32), ("W2", c_i32),
                ("coord_scale", c_f32),
                ("out", c_p),
                ("out_edge_stride", c_i64), ("out_elem_stride", c_i64)]


def dtype_code(t):
    try:
        return _DT[t.dtype]
    except KeyError:
        raise RuntimeError(f"dpvo_b200: unsupported dtype {t.dtype} (expected float16/float32/float64)")


def ptr(t):
    return c_p(t.data_ptr()) if t is not None else c_p(0)


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None) or (
    lambda idx: torch.cuda.current_stream(idx).cuda_stream)


def stream_of(t):
    """Current torch stream of the tensor's device (the reference uses the legacy default stream)."""
    idx = t.device.index
    # the raw handle, without building a torch.cuda.Stream object per launch (10 us each, ~20 launches per frame)
    return c_p(_raw_stream(idx if idx is not None else torch.cuda.current_device()))


def check(status, what):
    if status != 0:
        msg = lib.dpvo_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what}: {msg}")


def require_cuda(*tensors):
    for t in tensors:
        if not t.is_cuda:
            raise RuntimeError("dpvo_b200: expected CUDA tensors (no CPU path is provided for this operator)")


class device_guard:
    """Make the tensor's device current for the duration of a ctypes launch (kernels are launched on
    that device's current torch stream; the reference extensions rely on the caller for this)."""
    __slots__ = ("idx", "prev")

    def __init__(self, t):
        self.idx = t.device.index if t is not None and t.is_cuda else None
        self.prev = None

    def __enter__(self):
        if self.idx is not None:
            cur = torch.cuda.current_device()
            if cur != self.idx:
                self.prev = cur
                torch.cuda.set_device(self.idx)
        return self

    def __exit__(self, *exc):
        if self.prev is not None:
            torch.cuda.set_device(self.prev)
        return False
