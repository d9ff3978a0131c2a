"""Generate golden vectors from the UNMODIFIED reference, executed on CPU in the build container.

    python tests/golden/make_golden_cpu.py        (needs /root/reference and oracle/_ref/lietorch_backends.so)

What can run here without a GPU: the reference's lietorch CPU backend (compiled from its
own sources by oracle/build_ref.py), dpvo/projective_ops.py and dpvo/ba.py on top of it.  The reference's
correlation and fastba are CUDA-only; their golden vectors come from the GPU box (make_golden_gpu.py).

Outputs (small, committed): tests/golden/ref_cpu_se3.npz, ref_cpu_transform.npz, ref_cpu_ba.npz
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))       # lietorch_backends (reference build)
sys.path.insert(0, "/root/reference/methods/dpvo")

import lietorch_backends as lb                                   # noqa: E402
from dpvo.lietorch import SE3                                    # noqa: E402
from dpvo import projective_ops as pops                          # noqa: E402
from dpvo import ba as refba                                     # noqa: E402

synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
SE3_ID = 3  # group index, lietorch/groups.py:268


def se3_vectors():
    g = torch.Generator().manual_seed(7)
    n = 48
    xi = torch.randn(n, 6, generator=g) * torch.tensor([0.5, 0.5, 0.5, 0.8, 0.8, 0.8])
    xi[:4, 3:] *= 1e-6            # small-angle branch
    xi[4] = 0
    xi2 = torch.randn(n, 6, generator=g) * 0.3
    X = lb.expm(SE3_ID, xi)
    Y = lb.expm(SE3_ID, xi2)
    p3 = torch.randn(n, 3, generator=g)
    p4 = torch.cat([torch.randn(n, 3, generator=g), torch.rand(n, 1, generator=g)], 1)
    a6 = torch.randn(n, 6, generator=g)
    out = dict(xi=xi, xi2=xi2, p3=p3, p4=p4, a6=a6, X=X, Y=Y,
               log=lb.logm(SE3_ID, X), inv=lb.inv(SE3_ID, X), mul=lb.mul(SE3_ID, X, Y),
               adj=lb.adj(SE3_ID, X, a6), adjT=lb.adjT(SE3_ID, X, a6),
               act=lb.act(SE3_ID, X, p3), act4=lb.act4(SE3_ID, X, p4),
               matrix=lb.as_matrix(SE3_ID, X), projector=lb.projector(SE3_ID, X), Jinv=lb.Jinv(SE3_ID, X, a6))
    # backward functions (grad wrt inputs for a fixed upstream gradient)
    gX = torch.randn(n, 7, generator=g)
    g3 = torch.randn(n, 3, generator=g)
    g4 = torch.randn(n, 4, generator=g)
    g6 = torch.randn(n, 6, generator=g)
    out["expm_bwd"] = lb.expm_backward(SE3_ID, gX, xi)[0]
    out["logm_bwd"] = lb.logm_backward(SE3_ID, g6, X)[0]
    out["inv_bwd"] = lb.inv_backward(SE3_ID, gX, X)[0]
    mb = lb.mul_backward(SE3_ID, gX, X, Y)
    out["mul_bwd_X"], out["mul_bwd_Y"] = mb[0], mb[1]
    ab = lb.act_backward(SE3_ID, g3, X, p3)
    out["act_bwd_X"], out["act_bwd_p"] = ab[0], ab[1]
    a4 = lb.act4_backward(SE3_ID, g4, X, p4)
    out["act4_bwd_X"], out["act4_bwd_p"] = a4[0], a4[1]
    adb = lb.adj_backward(SE3_ID, g6, X, a6)
    out["adj_bwd_X"], out["adj_bwd_a"] = adb[0], adb[1]
    atb = lb.adjT_backward(SE3_ID, g6, X, a6)
    out["adjT_bwd_X"], out["adjT_bwd_a"] = atb[0], atb[1]
    out.update(gX=gX, g3=g3, g4=g4, g6=g6)
    return {k: v.numpy() for k, v in out.items()}


def transform_vectors():
    d = synth.ba_inputs(9, seed=11, M=8)
    poses = SE3(torch.from_numpy(d["poses"])[None])
    patches = torch.from_numpy(d["patches"])[None]
    intr = torch.from_numpy(d["intrinsics"])[None]
    ii, jj, kk = (torch.from_numpy(d[k]) for k in ("ii", "jj", "kk"))
    coords = pops.transform(poses, patches, intr, ii, jj, kk)                     # [1,E,3,3,2]
    c2, valid, (Ji, Jj, Jz) = pops.transform(poses, patches, intr, ii, jj, kk, jacobian=True)
    out = {k: d[k] for k in ("poses", "patches", "intrinsics", "ii", "jj", "kk", "target", "weight")}
    out.update(coords=coords.numpy(), valid=valid.numpy(), Ji=Ji.numpy(), Jj=Jj.numpy(), Jz=Jz.numpy())
    return out, d


def ba_vectors(d):
    """One call of the reference's differentiable pure-torch BA (dpvo/ba.py:86-182), as net.py:254-257 uses it."""
    poses = SE3(torch.from_numpy(d["poses"])[None])
    patches = torch.from_numpy(d["patches"])[None]
    intr = torch.from_numpy(d["intrinsics"])[None]
    ii, jj, kk = (torch.from_numpy(d[k]) for k in ("ii", "jj", "kk"))
    target = torch.from_numpy(d["target"])[None]
    weight = torch.from_numpy(d["weight"])[None]
    bounds = [-64, -64, 160 + 64, 120 + 64]
    lmbda = 1e-4   # a python float: ba.py:155-156 only reshapes tensor lmbda, and only to C.shape
    Gs, ps = refba.BA(poses, patches, intr, target, weight, lmbda, ii, jj, kk, bounds, ep=10.0, fixedp=1,
                      structure_only=False)
    return dict(poses_out=Gs.data.numpy(), patches_out=ps.numpy(), bounds=np.array(bounds, np.float32),
                ep=np.float32(10.0), fixedp=np.int64(1))


def main():
    torch.set_num_threads(1)
    np.savez_compressed(os.path.join(HERE, "ref_cpu_se3.npz"), **se3_vectors())
    tv, d = transform_vectors()
    np.savez_compressed(os.path.join(HERE, "ref_cpu_transform.npz"), **tv)
    np.savez_compressed(os.path.join(HERE, "ref_cpu_ba.npz"), **ba_vectors(d))
    for f in ("ref_cpu_se3.npz", "ref_cpu_transform.npz", "ref_cpu_ba.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
