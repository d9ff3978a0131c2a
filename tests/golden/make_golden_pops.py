"""Golden vectors for the fused projective transform (N2) from the UNMODIFIED reference on CPU.

    python tests/golden/make_golden_pops.py     (needs /root/reference and oracle/_ref/lietorch_backends.so)

Runs dpvo/projective_ops.py (transform with depth / valid / tonly / jacobian, flow_mag, point_cloud)
on top of the reference's lietorch CPU backend, with per-frame intrinsics that differ between frames.
Output (small, committed): tests/golden/ref_cpu_pops.npz
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
sys.path.insert(0, "/root/reference/methods/dpvo")

from dpvo.lietorch import SE3                                    # noqa: E402
from dpvo import projective_ops as pops                          # noqa: E402

synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")


def main():
    torch.set_num_threads(1)
    d = synth.ba_inputs(10, seed=17, M=12)
    rng = np.random.default_rng(3)
    intr = d["intrinsics"].copy()
    intr *= rng.uniform(0.9, 1.1, intr.shape).astype(np.float32)          # per-frame intrinsics
    poses_np = d["poses"].copy()
    poses_np[:, 3:] *= rng.uniform(0.7, 1.4, (len(poses_np), 1)).astype(np.float32)   # un-normalised quaternions
    poses = SE3(torch.from_numpy(poses_np)[None])
    patches = torch.from_numpy(d["patches"])[None]
    K = torch.from_numpy(intr)[None]
    ii, jj, kk = (torch.from_numpy(d[k]) for k in ("ii", "jj", "kk"))
    out = dict(poses=poses_np, patches=d["patches"], intrinsics=intr, ii=d["ii"], jj=d["jj"], kk=d["kk"])
    out["coords"] = pops.transform(poses, patches, K, ii, jj, kk).numpy()
    out["coords_depth"] = pops.transform(poses, patches, K, ii, jj, kk, depth=True).numpy()
    c, v = pops.transform(poses, patches, K, ii, jj, kk, valid=True)
    out["valid"] = v.numpy()
    out["coords_tonly"] = pops.transform(poses, patches, K, ii, jj, kk, tonly=True).numpy()
    c, jv, (Ji, Jj, Jz) = pops.transform(poses, patches, K, ii, jj, kk, jacobian=True)
    out.update(jvalid=jv.numpy(), Ji=Ji.numpy(), Jj=Jj.numpy(), Jz=Jz.numpy())
    # flow_mag mutates Gij of its tonly call only; safe to reuse poses
    f, fv = pops.flow_mag(poses, patches, K, ii, jj, kk, beta=0.5)
    out["flow"] = f.numpy(); out["flow_valid"] = fv.numpy()
    ix = torch.from_numpy((np.arange(patches.shape[1]) // d["M"]).astype(np.int64))
    out["ix"] = ix.numpy()
    out["points"] = pops.point_cloud(poses, patches, K, ix).numpy()
    p = os.path.join(HERE, "ref_cpu_pops.npz")
    np.savez_compressed(p, **out)
    print(p, os.path.getsize(p), "bytes", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
