"""Run fastba.BA a few times on one problem (for ncu launch lists).  python tools/run_ba_once.py --frames 70 --window 64"""
import argparse, importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser(); ap.add_argument("--frames", type=int, default=70); ap.add_argument("--window", type=int, default=64)
ap.add_argument("--reps", type=int, default=3); a = ap.parse_args()
pkg = importlib.import_module("deep-visual-odometry-slam_b200"); synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
d = synth.ba_inputs(a.frames, seed=0, n_buf=max(a.frames + 2, 128))
dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
P0 = dev(d["poses"])[None]; Z0 = dev(d["patches"])[None]
args = (dev(d["intrinsics"])[None], dev(d["target"])[None], dev(d["weight"])[None], torch.tensor([1e-4], device="cuda"),
        dev(d["ii"]), dev(d["jj"]), dev(d["kk"]))
for _ in range(a.reps):
    P = P0.clone(); Z = Z0.clone()
    pkg.fastba.BA(P, Z, *args, max(a.frames - a.window, 1), a.frames, d["M"], 2, False)
torch.cuda.synchronize(); print("ok")
