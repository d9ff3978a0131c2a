"""Blocked Cholesky solve vs numpy on random SPD systems (development tool)."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
pkg = importlib.import_module("deep-visual-odometry-slam_b200")
sizes = [int(a) for a in sys.argv[1:]] or [17, 64, 65, 128, 192, 234, 384, 1000]
for n in sizes:
    rng = np.random.default_rng(n)
    M = rng.standard_normal((n, n + 8))
    S = (M @ M.T / (n + 8) + 0.5 * np.eye(n)).astype(np.float32)
    y = rng.standard_normal(n).astype(np.float32)
    x, info = pkg.cuda_ba.cholesky_solve(torch.from_numpy(S).cuda(), torch.from_numpy(y).cuda())
    torch.cuda.synchronize()
    want = np.linalg.solve(S.astype(np.float64), y.astype(np.float64))
    err = np.linalg.norm(x.cpu().numpy() - want) / np.linalg.norm(want)
    print(n, "info", int(info.item()), "rel err", err, flush=True)
