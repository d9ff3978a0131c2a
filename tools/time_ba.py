"""Kernel-level timing of fastba.BA / neighbors / reproject (development tool).
    python tools/time_ba.py [--ref]
"""
import argparse, importlib, json, os, sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", action="store_true")
    a = ap.parse_args()
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
    ref = None
    if a.ref:
        from oracle.ref_ext import load_ref
        ref = load_ref("cuda_ba")
    dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
    for n_frames, win in [(40, 10), (40, 16), (40, 32), (70, 64)]:
        d = synth.ba_inputs(n_frames, seed=0, n_buf=max(n_frames + 2, 128))
        t0, t1 = max(n_frames - win, 1), n_frames
        P0 = dev(d["poses"])[None]; Z0 = dev(d["patches"])[None]
        args = (dev(d["intrinsics"])[None], dev(d["target"])[None], dev(d["weight"])[None],
                torch.tensor([1e-4], device="cuda"), dev(d["ii"]), dev(d["jj"]), dev(d["kk"]))
        P = P0.clone(); Z = Z0.clone()

        def ours():
            P.copy_(P0); Z.copy_(Z0)
            pkg.fastba.BA(P, Z, *args, t0, t1, d["M"], 2, False)
        def copy_only():
            P.copy_(P0); Z.copy_(Z0)
        base = timeit(copy_only)
        ms = timeit(ours) - base
        row = {"op": "BA", "edges": len(d["kk"]), "window": t1 - t0, "ms_b200": ms}
        if ref is not None:
            def theirs():
                P.copy_(P0); Z.copy_(Z0)
                ref.forward(P, Z, *args, d["M"], t0, t1, 2, False)
            row["ms_reference_ext"] = timeit(theirs, iters=10) - base
        print(json.dumps(row))
    kk, jj = args[6], args[5]
    d = synth.ba_inputs(40, seed=0)
    kk = dev(d["kk"]); jj = dev(d["jj"]); ii = dev(d["ii"])
    row = {"op": "neighbors", "edges": len(d["kk"]), "ms_b200": timeit(lambda: pkg.fastba.neighbors(kk, jj))}
    if ref is not None:
        row["ms_reference_ext"] = timeit(lambda: ref.neighbors(kk, jj), iters=10)
    print(json.dumps(row))
    Pp = dev(d["poses"])[None]; Zz = dev(d["patches"])[None]; In = dev(d["intrinsics"])[None]
    row = {"op": "reproject", "edges": len(d["kk"]), "ms_b200": timeit(lambda: pkg.fastba.reproject(Pp, Zz, In, ii, jj, kk))}
    if ref is not None:
        row["ms_reference_ext"] = timeit(lambda: ref.reproject(Pp, Zz, In, ii, jj, kk), iters=10)
    print(json.dumps(row))


if __name__ == "__main__":
    main()
