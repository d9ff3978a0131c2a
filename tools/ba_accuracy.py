"""Accuracy report for fastba.BA (development tool): error of this repo's fp32 kernels and of the
reference's fp32 CUDA extension against the float64 restatement, on the test-suite problems.
    python tools/ba_accuracy.py
"""
import importlib, json, os, sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import fastba_oracle as fo          # noqa: E402  (checker)
from oracle.ref_ext import load_ref             # noqa: E402


def main():
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
    ref = load_ref("cuda_ba")
    dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
    rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64) - b) / max(np.linalg.norm(b), 1e-30))
    cases = [("smoke", 12, 3, 96, 2, 12), ("local14", 14, 14, 96, 4, 14), ("local9", 9, 9, 32, 1, 9),
             ("local30", 30, 30, 24, 20, 30), ("wide40", 40, 7, 16, 1, 40), ("wide40b", 40, 22, 16, 1, 40), ("wide70", 70, 5, 24, 1, 70),
             ("full40", 40, 2, 96, 30, 40)]
    for name, nf, seed, M, t0, t1 in cases:
        d = synth.ba_inputs(nf, seed=seed, M=M)
        want = fo.ba(d["poses"], d["patches"], d["intrinsics"], d["target"], d["weight"], np.array([1e-4]),
                     d["ii"], d["jj"], d["kk"], d["M"], t0, t1, 2)
        row = {"case": name, "t0": t0, "t1": t1, "edges": len(d["kk"])}
        for tag, eff in (("ours", False), ("ours_eff", True), ("ref", False), ("ref_eff", True)):
            P = dev(d["poses"])[None].clone(); Z = dev(d["patches"])[None].clone()
            a = (dev(d["intrinsics"])[None], dev(d["target"])[None], dev(d["weight"])[None],
                 torch.tensor([1e-4], device="cuda"), dev(d["ii"]), dev(d["jj"]), dev(d["kk"]))
            if tag.startswith("ref"):
                if ref is None:
                    continue
                ref.forward(P, Z, *a, d["M"], t0, t1, 2, eff)
            else:
                pkg.fastba.BA(P, Z, *a, t0, t1, d["M"], 2, eff)
            torch.cuda.synchronize()
            row[tag + "_poses"] = rel(P[0].cpu().numpy(), want[0])
            row[tag + "_depths"] = rel(Z[0].cpu().numpy()[:, 2], want[1][:, 2])
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
