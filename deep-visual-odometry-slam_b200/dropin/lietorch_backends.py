"""Drop-in `lietorch_backends` for the unmodified reference Python code — imported by bare name at
lietorch/group_ops.py:1 (`import lietorch_backends`).  Put this directory FIRST on sys.path (or PYTHONPATH) instead of
installing the reference's extension; it re-exports the sm_100a implementation:
expm, logm, inv, mul, adj, adjT, act, act4 (+ _backward), as_matrix, projector, Jinv.
"""
from _bridge import load as _load

_m = _load("lietorch_backends")
globals().update({k: v for k, v in vars(_m).items() if not k.startswith("_")})
