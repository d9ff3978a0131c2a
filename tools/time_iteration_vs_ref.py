"""One `DPVO.update` body (dpvo.py:328-354) on the 45,312-edge default window, two ways, in one process on one GPU:

  reference  the reference's OWN CUDA extensions rebuilt for sm_100 (oracle/_ref: cuda_ba.reproject,
             cuda_corr.forward x 2 levels + stack, cuda_ba.neighbors, cuda_ba.forward) around the torch
             formulation of its Update module under autocast fp16 (oracle/update_oracle.py), as stream
             launches with the reference's host synchronisations (neighbors' D2H/H2D) — how the reference runs;
  ours       hotpath.UpdateIteration, one CUDA graph.

The reference side is given every advantage that is cheap to give: cuda_ba.reproject instead of the
lietorch + projective_ops chain it really calls (dpvo.py:209-213), neighbours computed once per iteration.
Development-time measurement tool (like tools/time_corr.py --ref): it is not part of the package, the tests or bench.py, and
nothing in the product path loads oracle/.
"""
import importlib
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
hotpath = importlib.import_module("deep-visual-odometry-slam_b200.hotpath")
synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
from oracle import update_oracle as uo
from oracle.ref_ext import load_ref


def main():
    dev = torch.device("cuda", 0)
    it = hotpath.UpdateIteration(dev, n_frames=40, seed=0)
    for _ in range(3):
        it.step_full()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        it.step_full()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    K = 30
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    ms_ours = e0.elapsed_time(e1) / K
    out = dict(edges=it.E, ours_ms=round(ms_ours, 4), ours_frames_per_s=round(1e3 / ms_ours, 1))

    corr_ext, ba_ext = load_ref("cuda_corr"), load_ref("cuda_ba")
    if corr_ext is None or ba_ext is None:
        out["reference"] = "oracle/_ref not built"
        print(json.dumps(out)); return
    c = it.host["c"]
    f1 = torch.from_numpy(c["fmap1"]).to(dev); f2 = torch.from_numpy(c["fmap2"]).to(dev)      # NCHW rings, as the reference keeps them
    gmap = it.gmap
    sd = uo.state_dict_tensors(synth.update_state_dict(3), device=dev)
    us, vs = it.kk % it.n_patches, it.jj % it.mem
    inp32 = it.imap.float()
    net = it.net.float()
    stages = {}

    def stage(name, fn):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); r = fn(); b.record()
        stages.setdefault(name, []).append((a, b))
        return r

    def ref_iteration():
        poses, patches = it.poses0.clone(), it.patches0.clone()
        coords = stage("reproject", lambda: ba_ext.reproject(poses, patches, it.intrinsics, it.ii, it.jj, it.kk))
        def corr():
            c1, = corr_ext.forward(gmap, f1, coords, us, vs, 3)
            c2, = corr_ext.forward(gmap, f2, coords / 4, us, vs, 3)
            return torch.stack([c1, c2], -1).view(1, it.E, -1)
        cr = stage("corr", corr)
        neigh = stage("neighbors", lambda: tuple(ba_ext.neighbors(it.kk, it.jj)))
        with torch.no_grad():
            _, delta, weight = stage("update", lambda: uo.forward(sd, net[0], inp32[us], cr[0].float(), it.ii, it.jj, it.kk,
                                                                   autocast=True, neigh=neigh))
        target = coords[..., 1, 1] + delta.float()[None]
        stage("ba", lambda: ba_ext.forward(poses, patches, it.intrinsics, target.contiguous(), weight.float()[None].contiguous(),
                                           it.lmbda, it.ii, it.jj, it.kk, it.M, it.t0, it.t1, 2, False))

    for _ in range(2):
        ref_iteration()
    torch.cuda.synchronize()
    stages.clear()
    R = 5
    t0 = time.perf_counter()
    e0.record()
    for _ in range(R):
        ref_iteration()
    e1.record(); torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / R * 1e3
    ms_ref = e0.elapsed_time(e1) / R
    out.update(reference_ms=round(ms_ref, 3), reference_wall_ms=round(wall, 3), reference_frames_per_s=round(1e3 / ms_ref, 1),
               reference_stage_ms={k: round(sum(a.elapsed_time(b) for a, b in v) / len(v), 3) for k, v in stages.items()},
               speedup=round(ms_ref / ms_ours, 1))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
