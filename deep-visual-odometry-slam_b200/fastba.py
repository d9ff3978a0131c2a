"""`fastba` — Python operator surface of the bundle adjustment
(mirrors /root/reference/methods/dpvo/dpvo/fastba/ba.py:1-8, including its argument reorder)."""
from . import cuda_ba

neighbors = cuda_ba.neighbors
reproject = cuda_ba.reproject


def BA(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, M, iterations, eff_impl=False,
       prepared=None):
    args = (poses.data, patches, intrinsics, target, weight, lmbda, ii, jj, kk, M, t0, t1, iterations, eff_impl)
    return cuda_ba.forward(*args) if prepared is None else cuda_ba.forward(*args, prepared=prepared)


def BA_prepare(poses, patches, ii, jj, kk, t0, t1, M, eff_impl=False, workspace=None):
    """Edge-list-only part of BA (see cuda_ba.prepare); hand the result to BA(..., prepared=ws)."""
    P = patches.shape[-1]
    return cuda_ba.prepare(ii, jj, kk, poses.numel() // 7, patches.numel() // (3 * P * P), P, M, t0, t1, eff_impl,
                           workspace=workspace)
