"""Golden vectors for ALL FOUR Lie groups of lietorch_backends (SO3=1, RxSO3=2, SE3=3, Sim3=4) from the UNMODIFIED
reference: its CPU backend (lietorch/src/lietorch_cpu.cpp, compiled from the reference's own sources by
oracle/build_ref.py into oracle/_ref/lietorch_backends.so) executed in the build container.

    python tests/golden/make_golden_lie_groups.py        -> tests/golden/ref_cpu_lie_groups.npz (small, committed)

float64 inputs / outputs (the fp32 device and host paths are compared against them with fp32 tolerances).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import lietorch_backends as lb                                   # noqa: E402  (reference build)

DIMS = {1: (3, 4), 2: (4, 5), 3: (6, 7), 4: (7, 8)}


def tangent(group, n, g, scale=1.0):
    K, _ = DIMS[group]
    a = torch.randn(n, K, generator=g, dtype=torch.float64)
    if group == 1:
        a *= 0.9
    elif group == 2:
        a *= torch.tensor([0.9, 0.9, 0.9, 0.3], dtype=torch.float64)
    elif group == 3:
        a *= torch.tensor([0.5, 0.5, 0.5, 0.9, 0.9, 0.9], dtype=torch.float64)
    else:
        a *= torch.tensor([0.5, 0.5, 0.5, 0.9, 0.9, 0.9, 0.3], dtype=torch.float64)
    return a * scale


def vectors(group, n=40, seed=11):
    K, N = DIMS[group]
    g = torch.Generator().manual_seed(seed + group)
    a = tangent(group, n, g)
    rot = slice(0, 3) if group in (1, 2) else slice(3, 6)
    a[:3, rot] *= 1e-8                      # small-angle branches
    if group in (2, 4):
        a[2:5, K - 1] = 0.0                 # sigma == 0 branches of W
    a[5] = 0
    X = lb.expm(group, a)
    Y = lb.expm(group, tangent(group, n, g, 0.5))
    p3 = torch.randn(n, 3, generator=g, dtype=torch.float64)
    p4 = torch.cat([torch.randn(n, 3, generator=g, dtype=torch.float64), torch.rand(n, 1, generator=g, dtype=torch.float64)], 1)
    aK = torch.randn(n, K, generator=g, dtype=torch.float64)
    gN = torch.randn(n, N, generator=g, dtype=torch.float64)
    gK = torch.randn(n, K, generator=g, dtype=torch.float64)
    g3 = torch.randn(n, 3, generator=g, dtype=torch.float64)
    g4 = torch.randn(n, 4, generator=g, dtype=torch.float64)
    out = dict(a=a, X=X, Y=Y, p3=p3, p4=p4, aK=aK, gN=gN, gK=gK, g3=g3, g4=g4,
               log=lb.logm(group, X), inv=lb.inv(group, X), mul=lb.mul(group, X, Y), adj=lb.adj(group, X, aK),
               adjT=lb.adjT(group, X, aK), act=lb.act(group, X, p3), act4=lb.act4(group, X, p4),
               matrix=lb.as_matrix(group, X), projector=lb.projector(group, X), Jinv=lb.Jinv(group, X, aK))
    out["expm_bwd"] = lb.expm_backward(group, gN, a)[0]
    out["logm_bwd"] = lb.logm_backward(group, gK, X)[0]
    out["inv_bwd"] = lb.inv_backward(group, gN, X)[0]
    out["mul_bwd_X"], out["mul_bwd_Y"] = lb.mul_backward(group, gN, X, Y)
    out["act_bwd_X"], out["act_bwd_p"] = lb.act_backward(group, g3, X, p3)
    out["act4_bwd_X"], out["act4_bwd_p"] = lb.act4_backward(group, g4, X, p4)
    out["adj_bwd_X"], out["adj_bwd_a"] = lb.adj_backward(group, gK, X, aK)
    out["adjT_bwd_X"], out["adjT_bwd_a"] = lb.adjT_backward(group, gK, X, aK)
    return {f"g{group}_{k}": v.numpy() for k, v in out.items()}


if __name__ == "__main__":
    blob = {}
    for grp in (1, 2, 3, 4):
        blob.update(vectors(grp))
    path = os.path.join(HERE, "ref_cpu_lie_groups.npz")
    np.savez_compressed(path, **blob)
    print("wrote", path, len(blob), "arrays", os.path.getsize(path), "bytes")
