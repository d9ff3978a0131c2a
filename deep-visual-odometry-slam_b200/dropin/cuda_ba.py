"""Drop-in `cuda_ba` for the unmodified reference Python code — imported by bare name at
fastba/ba.py:2 and loop_closure/optim_utils.py:1 (`import cuda_ba`).  Put this directory FIRST on sys.path (or PYTHONPATH) instead of
installing the reference's extension; it re-exports the sm_100a implementation:
forward, neighbors, reproject, solve_system.
"""
from _bridge import load as _load

_m = _load("cuda_ba")
globals().update({k: v for k, v in vars(_m).items() if not k.startswith("_")})
