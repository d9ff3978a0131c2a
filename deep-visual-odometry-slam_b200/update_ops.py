"""Host side of the Update-operator kernels (csrc/update_gemm.cu): thin ctypes wrappers.

`linear384` is one nn.Linear(K, 384) over the edge rows with the row-local work that follows it in
Update.forward (net.py:69-88) / GatedResidual.forward (blocks.py:30-31) fused into the epilogue.
CUDA only; there is no CPU fallback.
"""
import ctypes

import torch

from . import _lib

DIM = 384


class _LinearParams(ctypes.Structure):
    _fields_ = [("A", ctypes.c_void_p), ("lda", ctypes.c_int64),
                ("W", ctypes.c_void_p), ("ldw", ctypes.c_int64),
                ("bias", ctypes.c_void_p),
                ("gate", ctypes.c_void_p),
                ("res1", ctypes.c_void_p),
                ("res2", ctypes.c_void_p), ("res2_index", ctypes.c_void_p),
                ("ln_gamma", ctypes.c_void_p), ("ln_beta", ctypes.c_void_p), ("ln_eps", ctypes.c_float),
                ("act", ctypes.c_int32),
                ("Y", ctypes.c_void_p),
                ("M", ctypes.c_int64), ("K", ctypes.c_int32),
                ("head_w", ctypes.c_void_p), ("head_acc", ctypes.c_void_p)]


_lib.lib.dpvo_linear384_f16.argtypes = [ctypes.POINTER(_LinearParams), ctypes.c_void_p]
_lib.lib.dpvo_linear384_f16.restype = ctypes.c_int


def _rows384(t, name, M=None):
    if t is None:
        return None
    if t.dtype != torch.float16 or t.dim() != 2 or t.shape[1] != DIM or not t.is_contiguous():
        raise RuntimeError(f"linear384: {name} must be a contiguous float16 [rows, {DIM}] tensor")
    if M is not None and t.shape[0] != M:
        raise RuntimeError(f"linear384: {name} must have {M} rows, got {t.shape[0]}")
    return t


def _vec384(t, name):
    if t is None:
        return None
    if t.dtype != torch.float32 or t.numel() != DIM or not t.is_contiguous():
        raise RuntimeError(f"linear384: {name} must be a contiguous float32 [{DIM}] tensor")
    return t


def _dp(t):
    return 0 if t is None else t.data_ptr()


def linear384_trusted(A, W, bias, stream, gate=None, res1=None, res2=None, idx=None, ln=None, act=0, head=None):
    """linear384 for callers that have already validated their operands and made the device current (update.Update
    issues 19 of these per call: the checks, the device guard and the stream lookup of the public entry cost more host
    time than the launch)."""
    M, K = A.shape
    Y = torch.empty((M, DIM), dtype=torch.float16, device=A.device)
    gamma, beta, eps = ln if ln is not None else (None, None, 0.0)
    hw, ha = head if head is not None else (None, None)
    p = _LinearParams(_dp(A), A.stride(0) if M > 1 else max(A.stride(0), K), _dp(W), W.stride(0), _dp(bias), _dp(gate),
                      _dp(res1), _dp(res2), _dp(idx), _dp(gamma), _dp(beta), float(eps), int(act), _dp(Y), M, K, _dp(hw),
                      _dp(ha))
    if _lib.lib.dpvo_linear384_f16(ctypes.byref(p), stream):
        _lib.check(1, "linear384")
    return Y


def linear384(A, W, bias=None, gate=None, res1=None, res2=None, idx=None, ln=None, act=0, out=None, head=None):
    """Y = act(LN((A W^T + bias) * gate + res1 + res2[idx])) as float16 [M, 384].

    A: float16 [M, K] (row stride a multiple of 8 elements), W: float16 [384, K] (likewise);
    ln = (gamma, beta, eps) or None; act: 0 none, 1 ReLU, 2 sigmoid.
    head = (w [4, 384] float32, acc [M, 4] float32 zeroed): acc += ReLU(Y) w^T (K = 384, no ln)."""
    _lib.require_cuda(A, W)
    if A.dtype != torch.float16 or W.dtype != torch.float16 or A.dim() != 2 or W.dim() != 2:
        raise RuntimeError("linear384: A and W must be 2-D float16")
    M, K = A.shape
    if W.shape[0] != DIM or W.shape[1] != K:
        raise RuntimeError(f"linear384: W must be [{DIM}, {K}], got {tuple(W.shape)}")
    if A.stride(1) != 1 or W.stride(1) != 1:
        raise RuntimeError("linear384: A and W must have unit stride along K")
    _rows384(gate, "gate", M); _rows384(res1, "res1", M)
    _rows384(res2, "res2", None if idx is not None else M)
    if idx is not None:
        if res2 is None or idx.dtype != torch.int64 or idx.numel() != M or not idx.is_contiguous():
            raise RuntimeError("linear384: idx must be a contiguous int64 [M] tensor indexing res2 rows")
    gamma = beta = None
    eps = 0.0
    if ln is not None:
        gamma, beta, eps = ln
        _vec384(gamma, "ln gamma"); _vec384(beta, "ln beta")
    _vec384(bias, "bias")
    head_w = head_acc = None
    if head is not None:
        head_w, head_acc = head
        if head_w.dtype != torch.float32 or tuple(head_w.shape) != (4, DIM) or not head_w.is_contiguous():
            raise RuntimeError(f"linear384: head weights must be contiguous float32 [4, {DIM}]")
        if head_acc.dtype != torch.float32 or tuple(head_acc.shape) != (M, 4) or not head_acc.is_contiguous():
            raise RuntimeError("linear384: head accumulator must be contiguous float32 [M, 4]")
    Y = out if out is not None else torch.empty((M, DIM), dtype=torch.float16, device=A.device)
    _rows384(Y, "out", M)
    p = _LinearParams(A=_lib.ptr(A).value, lda=A.stride(0) if M > 1 else max(A.stride(0), K), W=_lib.ptr(W).value,
                      ldw=W.stride(0), bias=_lib.ptr(bias).value, gate=_lib.ptr(gate).value,
                      res1=_lib.ptr(res1).value, res2=_lib.ptr(res2).value, res2_index=_lib.ptr(idx).value,
                      ln_gamma=_lib.ptr(gamma).value, ln_beta=_lib.ptr(beta).value, ln_eps=float(eps), act=int(act),
                      Y=_lib.ptr(Y).value, M=M, K=K, head_w=_lib.ptr(head_w).value, head_acc=_lib.ptr(head_acc).value)
    with torch.cuda.device(A.device):
        _lib.check(_lib.lib.dpvo_linear384_f16(ctypes.byref(p), _lib.stream_of(A)), "linear384")
    return Y
