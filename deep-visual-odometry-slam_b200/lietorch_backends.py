"""Drop-in for the reference's `lietorch_backends` pybind module (lietorch/src/lietorch.cpp:18-316).

Same function names, argument order and return conventions: every function takes the group id first
(1 = SO3, 2 = RxSO3, 3 = SE3, 4 = Sim3 — lietorch/groups.py:236,252,268,290), tensors are 2-D
contiguous `[batch, width]` float32 / float64 (CHECK_CONTIGUOUS, lietorch.cpp:7), forward ops return
a tensor, backward ops return a list of tensors in input order.  Outputs are allocated through
torch's caching allocator; the arithmetic runs in libdpvo_b200.so (csrc/lietorch.cu) on torch's
current stream.  SO3 and SE3 are provided (DPVO only instantiates SE3, projective_ops.py:60-65,
dpvo.py:419-420); other groups and CPU tensors raise RuntimeError — there is no CPU fallback.
"""
import torch

from . import _lib

_L = _lib.lib
_DIMS = {1: (3, 4), 3: (6, 7)}  # group id -> (K tangent, N embedding)


def _dims(group):
    try:
        return _DIMS[int(group)]
    except KeyError:
        raise RuntimeError(f"lietorch_backends: group {group} is not provided (SO3=1, SE3=3 are)")


def _check(group, *specs):
    """specs: (tensor, name, width).  Returns (batch, dtype code)."""
    first = specs[0][0]
    _lib.require_cuda(*[s[0] for s in specs])
    if first.dtype not in (torch.float32, torch.float64):
        raise RuntimeError(f"lietorch_backends: expected float32/float64, got {first.dtype}")
    n = first.shape[0] if first.dim() == 2 else -1
    for t, name, w in specs:
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")
        if t.dim() != 2 or t.shape[0] != n or t.shape[1] != w:
            raise RuntimeError(f"lietorch_backends: {name} must be [{n}, {w}], got {tuple(t.shape)}")
        if t.dtype != first.dtype or t.device != first.device:
            raise RuntimeError(f"lietorch_backends: {name} dtype/device differs from the first argument")
    return n, _lib.dtype_code(first)


def _call(fn, what, group, dt, n, ref, *ptr_tensors):
    args = [_lib.c_i32(int(group)), _lib.c_i32(dt)] + [_lib.ptr(t) for t in ptr_tensors] + \
           [_lib.c_i64(n), _lib.stream_of(ref)]
    with torch.cuda.device(ref.device):
        _lib.check(fn(*args), what)


def _new(ref, *shape):
    return torch.empty(shape, dtype=ref.dtype, device=ref.device)


# ----------------------------------------------------------------------------- unary
def expm(group, a):
    K, N = _dims(group)
    n, dt = _check(group, (a, "a", K))
    X = _new(a, n, N)
    _call(_L.dpvo_lie_expm, "expm", group, dt, n, a, a, X)
    return X


def expm_backward(group, grad, a):
    K, N = _dims(group)
    n, dt = _check(group, (a, "a", K), (grad, "grad", N))
    da = _new(grad, n, K)
    _call(_L.dpvo_lie_expm_backward, "expm_backward", group, dt, n, a, grad, a, da)
    return [da]


def logm(group, X):
    K, N = _dims(group)
    n, dt = _check(group, (X, "X", N))
    a = _new(X, n, K)
    _call(_L.dpvo_lie_logm, "logm", group, dt, n, X, X, a)
    return a


def logm_backward(group, grad, X):
    K, N = _dims(group)
    n, dt = _check(group, (X, "X", N), (grad, "grad", K))
    dX = _new(grad, n, N)
    _call(_L.dpvo_lie_logm_backward, "logm_backward", group, dt, n, X, grad, X, dX)
    return [dX]


def inv(group, X):
    K, N = _dims(group)
    n, dt = _check(group, (X, "X", N))
    Y = _new(X, n, N)
    _call(_L.dpvo_lie_inv, "inv", group, dt, n, X, X, Y)
    return Y


def inv_backward(group, grad, X):
    K, N = _dims(group)
    n, dt = _check(group, (X, "X", N), (grad, "grad", N))
    dX = _new(grad, n, N)
    _call(_L.dpvo_lie_inv_backward, "inv_backward", group, dt, n, X, grad, X, dX)
    return [dX]


# ----------------------------------------------------------------------------- binary
def mul(group, X, Y):
    K, N = _dims(group)
    n, dt = _check(group, (X, "X", N), (Y, "Y", N))
    Z = _new(X, n, N)
    _call(_L.dpvo_lie_mul, "mul", group, dt, n, X, X, Y, Z)
    return Z


def mul_backward(group, grad, X, Y):
    K, N = _dims(group)
    n, dt = _check(group, (X, "X", N), (Y, "Y", N), (grad, "grad", N))
    dX, dY = _new(grad, n, N), _new(grad, n, N)
    _call(_L.dpvo_lie_mul_backward, "mul_backward", group, dt, n, X, grad, X, Y, dX, dY)
    return [dX, dY]


def _binary(fn, what, group, X, b, wb):
    K, N = _dims(group)
    w = K if wb is None else wb
    n, dt = _check(group, (X, "X", N), (b, what + " operand", w))
    out = _new(b, n, w)
    _call(fn, what, group, dt, n, X, X, b, out)
    return out


def _binary_backward(fn, what, group, grad, X, b, wb):
    K, N = _dims(group)
    w = K if wb is None else wb
    n, dt = _check(group, (X, "X", N), (b, what + " operand", w), (grad, "grad", w))
    dX, db = _new(grad, n, N), _new(grad, n, w)
    _call(fn, what, group, dt, n, X, grad, X, b, dX, db)
    return [dX, db]


def adj(group, X, a):
    return _binary(_L.dpvo_lie_adj, "adj", group, X, a, None)


def adj_backward(group, grad, X, a):
    return _binary_backward(_L.dpvo_lie_adj_backward, "adj_backward", group, grad, X, a, None)


def adjT(group, X, a):
    return _binary(_L.dpvo_lie_adjT, "adjT", group, X, a, None)


def adjT_backward(group, grad, X, a):
    return _binary_backward(_L.dpvo_lie_adjT_backward, "adjT_backward", group, grad, X, a, None)


def act(group, X, p):
    return _binary(_L.dpvo_lie_act, "act", group, X, p, 3)


def act_backward(group, grad, X, p):
    return _binary_backward(_L.dpvo_lie_act_backward, "act_backward", group, grad, X, p, 3)


def act4(group, X, p):
    return _binary(_L.dpvo_lie_act4, "act4", group, X, p, 4)


def act4_backward(group, grad, X, p):
    return _binary_backward(_L.dpvo_lie_act4_backward, "act4_backward", group, grad, X, p, 4)


# ----------------------------------------------------------------------------- matrices
def as_matrix(group, X):
    K, N = _dims(group)
    n, dt = _check(group, (X, "X", N))
    T = _new(X, n, 4, 4)
    _call(_L.dpvo_lie_as_matrix, "as_matrix", group, dt, n, X, X, T)
    return T


def projector(group, X):
    K, N = _dims(group)
    n, dt = _check(group, (X, "X", N))
    P = _new(X, n, N, N)
    _call(_L.dpvo_lie_projector, "projector", group, dt, n, X, X, P)
    return P


def Jinv(group, X, a):
    return _binary(_L.dpvo_lie_jinv, "Jinv", group, X, a, None)
