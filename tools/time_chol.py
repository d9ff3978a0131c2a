"""Blocked Cholesky solve alone (development tool): ms per dpvo_chol_solve at several n, error against float64,
and per-kernel CUPTI records at the largest size.   python tools/time_chol.py [n ...]"""
import importlib, json, os, sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from trace_global_ba import kernel_profile, time_call


def main():
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    mode = 1 if "--fp32" in sys.argv else 0          # bulk trailing update: 0 split-TF32 tensor cores, 1 fp32 FMA
    pkg.cuda_ba.chol_bulk_mode(mode)
    sizes = [int(a) for a in args] or [384, 1000, 2310, 4890, 5994]
    rows = []
    for n in sizes:
        g = torch.Generator(device="cuda").manual_seed(n)
        M = torch.randn(n, n + 8, device="cuda", generator=g, dtype=torch.float64)
        S64 = M @ M.t() / (n + 8) + 0.5 * torch.eye(n, device="cuda", dtype=torch.float64)
        y64 = torch.randn(n, device="cuda", generator=g, dtype=torch.float64)
        S, y = S64.float(), y64.float()
        x, info = pkg.cuda_ba.cholesky_solve(S, y)
        want = torch.linalg.solve(S.double(), y.double())
        err = float((x.double() - want).norm() / want.norm())
        ms, reps = time_call(lambda: pkg.cuda_ba.cholesky_solve(S, y))
        row = {"n": n, "bulk": "fp32" if mode else "split_tf32", "info": int(info.item()), "rel_err_vs_float64": err, "ms": ms, "reps_ms": reps}
        if n == sizes[-1]:
            row.update(kernel_profile(lambda: pkg.cuda_ba.cholesky_solve(S, y)))
        rows.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "time_chol.json"), "w") as f:
        json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
