"""Drop-in for the reference's `lietorch_backends` pybind module (lietorch/src/lietorch.cpp:18-316).

Same function names, argument order and return conventions: every function takes the group id first
(1 = SO3, 2 = RxSO3, 3 = SE3, 4 = Sim3 — lietorch/groups.py:236,252,268,290), tensors are 2-D
contiguous `[batch, width]` float32 / float64 (CHECK_CONTIGUOUS, lietorch.cpp:7), forward ops return
a tensor, backward ops return a list of tensors in input order.  Outputs are allocated through
torch's caching allocator; the arithmetic runs in libdpvo_b200.so (csrc/lietorch.cu) on torch's
current stream.  All four groups are provided.  Like the reference (lietorch/include/dispatch.h:12-43) the module
dispatches on the tensors' device: CUDA tensors run the sm_100a kernels, CPU tensors (data readers, evaluation scripts)
run the same templates compiled for the host (`dpvo_lie_host`) — that is the reference's own device dispatch, not a
fallback: a CUDA tensor never takes the host path.
"""
import torch

from . import _lib

_L = _lib.lib
_DIMS = {1: (3, 4), 2: (4, 5), 3: (6, 7), 4: (7, 8)}  # group id -> (K tangent, N embedding)
# device entry point -> (dpvo_lie_op code, number of input pointers) for the host form
_HOST = {"dpvo_lie_expm": (0, 1), "dpvo_lie_expm_backward": (1, 2), "dpvo_lie_logm": (2, 1), "dpvo_lie_logm_backward": (3, 2),
         "dpvo_lie_inv": (4, 1), "dpvo_lie_inv_backward": (5, 2), "dpvo_lie_mul": (6, 2), "dpvo_lie_mul_backward": (7, 3),
         "dpvo_lie_adj": (8, 2), "dpvo_lie_adj_backward": (9, 3), "dpvo_lie_adjT": (10, 2), "dpvo_lie_adjT_backward": (11, 3),
         "dpvo_lie_act": (12, 2), "dpvo_lie_act_backward": (13, 3), "dpvo_lie_act4": (14, 2), "dpvo_lie_act4_backward": (15, 3),
         "dpvo_lie_as_matrix": (16, 1), "dpvo_lie_projector": (17, 1), "dpvo_lie_jinv": (18, 2)}


def _dims(group):
    try:
        return _DIMS[int(group)]
    except KeyError:
        raise RuntimeError(f"lietorch_backends: unknown group {group} (SO3=1, RxSO3=2, SE3=3, Sim3=4)")


def _check(group, *specs):
    """specs: (tensor, name, width).  Returns (batch, dtype code)."""
    first = specs[0][0]
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
    if not ref.is_cuda:
        op, n_in = _HOST[fn.__name__]
        ins = [_lib.ptr(t) for t in ptr_tensors[:n_in]] + [_lib.c_p(0)] * (3 - n_in)
        outs = [_lib.ptr(t) for t in ptr_tensors[n_in:]]
        outs += [_lib.c_p(0)] * (2 - len(outs))
        _lib.check(_L.dpvo_lie_host(_lib.c_i32(op), _lib.c_i32(int(group)), _lib.c_i32(dt), *ins, *outs, _lib.c_i64(n)), what)
        return
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
