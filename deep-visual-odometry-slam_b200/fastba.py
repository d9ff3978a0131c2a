"""`fastba` — Python operator surface of the bundle adjustment
(mirrors /root/reference/methods/dpvo/dpvo/fastba/ba.py:1-8, including its argument reorder)."""
from . import cuda_ba

neighbors = cuda_ba.neighbors
reproject = cuda_ba.reproject


def BA(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, t0, t1, M, iterations, eff_impl=False):
    return cuda_ba.forward(poses.data, patches, intrinsics, target, weight, lmbda, ii, jj, kk, M, t0, t1,
                           iterations, eff_impl)
