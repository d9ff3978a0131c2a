"""Drop-in `cuda_corr` for the unmodified reference Python code — imported by bare name at
altcorr/correlation.py:2 (`import cuda_corr`).  Put this directory FIRST on sys.path (or PYTHONPATH) instead of
installing the reference's extension; it re-exports the sm_100a implementation:
forward, backward, patchify_forward, patchify_backward.
"""
from _bridge import load as _load

_m = _load("cuda_corr")
globals().update({k: v for k, v in vars(_m).items() if not k.startswith("_")})
