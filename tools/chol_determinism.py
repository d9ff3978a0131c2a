"""Run-to-run determinism of the blocked Cholesky solve (development tool): the same system solved repeatedly must give
bit-identical results in both bulk arithmetics (every matrix element has one writer per launch)."""
import importlib, json, os, sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    for n in [int(a) for a in sys.argv[1:]] or [2310, 4890]:
        g = torch.Generator(device="cuda").manual_seed(n)
        M = torch.randn(n, n + 8, device="cuda", generator=g, dtype=torch.float64)
        S = (M @ M.t() / (n + 8) + 0.5 * torch.eye(n, device="cuda", dtype=torch.float64)).float()
        y = torch.randn(n, device="cuda", generator=g)
        want = torch.linalg.solve(S.double(), y.double())
        for mode in (1, 0):
            pkg.cuda_ba.chol_bulk_mode(mode)
            xs, errs, infos = [], [], []
            for _ in range(8):
                x, info = pkg.cuda_ba.cholesky_solve(S, y)
                xs.append(x.clone()); infos.append(int(info.item()))
                errs.append(float((x.double() - want).norm() / want.norm()))
            same = all(torch.equal(xs[0], x) for x in xs[1:])
            print(json.dumps({"n": n, "mode": mode, "bit_identical": same, "info": infos, "errs": [f"{e:.2e}" for e in errs]}), flush=True)
    pkg.cuda_ba.chol_bulk_mode(0)


if __name__ == "__main__":
    main()
