"""Timing of DPV-SLAM-style global BA (BASELINE config 5): N keyframes, every sequential edge kept,
proximity loop-closure frame pairs, eff_impl=True, 2 iterations (development / measurement tool).

    python tools/time_global_ba.py [--frames 1000] [--M 96] [--loops 200] [--ref] [--iters 3]

Prints one JSON line: ms per call for this repo and (with --ref) for the reference's own extension
(oracle/_ref/cuda_ba.so — test infrastructure) on identical inputs, plus the relative difference of
the two results and the weighted reprojection cost before / after.
"""
import argparse, importlib, json, os, sys, time
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--M", type=int, default=96)
    ap.add_argument("--loops", type=int, default=200)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--ref", action="store_true")
    a = ap.parse_args()
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
    nf, M = a.frames, a.M
    rng = np.random.default_rng(0)
    ii, jj, kk = synth.dpvo_edges(nf, M, removal=10 ** 9)
    n_loop = 0
    for _ in range(a.loops):
        i, j = int(rng.integers(0, nf)), int(rng.integers(0, nf))
        if abs(i - j) >= 30:
            ii = np.concatenate([ii, np.full(M, i)]); jj = np.concatenate([jj, np.full(M, j)])
            kk = np.concatenate([kk, i * M + np.arange(M)]); n_loop += 1
    d = synth.ba_inputs(nf, seed=1, M=M, n_buf=nf + 2, edges=(ii.astype(np.int64), jj.astype(np.int64), kk.astype(np.int64)))
    dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
    P0 = dev(d["poses"])[None]; Z0 = dev(d["patches"])[None]
    args = (dev(d["intrinsics"])[None], dev(d["target"])[None], dev(d["weight"])[None],
            torch.tensor([1e-4], device="cuda"), dev(d["ii"]), dev(d["jj"]), dev(d["kk"]))
    t0, t1 = 1, nf

    def run(fn):
        ts = []
        out = None
        for _ in range(a.iters + 1):
            P = P0.clone(); Z = Z0.clone()
            torch.cuda.synchronize()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            w0 = time.perf_counter()
            e0.record(); fn(P, Z); e1.record()
            torch.cuda.synchronize()
            ts.append((e0.elapsed_time(e1), (time.perf_counter() - w0) * 1e3))
            out = (P, Z)
        ts = ts[1:]
        return float(np.median([t[0] for t in ts])), float(np.median([t[1] for t in ts])), out

    ms, wall, (P, Z) = run(lambda P, Z: pkg.fastba.BA(P, Z, *args, t0, t1, M, 2, True))
    row = {"op": "global_BA_eff", "keyframes": nf, "M": M, "edges": int(len(kk)), "loop_pairs": n_loop,
           "free_poses": t1 - t0, "system": 6 * (t1 - t0), "ms_b200": ms, "wall_ms_b200": wall}
    rel = lambda x, y: float((x - y).norm() / y.norm())
    if a.ref:
        from oracle.ref_ext import load_ref
        ref = load_ref("cuda_ba")
        if ref is not None:
            msr, wallr, (Pr, Zr) = run(lambda P, Z: ref.forward(P, Z, *args, M, t0, t1, 2, True))
            row.update(ms_reference_ext=msr, wall_ms_reference_ext=wallr,
                       rel_diff_poses=rel(P, Pr), rel_diff_depths=rel(Z[0, :, 2], Zr[0, :, 2]),
                       rel_update_poses=rel(P - P0, Pr - P0))
    print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
