"""Where a chol_panel_kernel launch spends its cycles (development tool; needs tools/build_prof.sh's library).

    DPVO_B200_LIB=deep-visual-odometry-slam_b200/libdpvo_b200_prof.so python tools/chol_stage_profile.py [n ...]

Thread 0 of CTA 1 (a CTA that does every stage) adds the cycles between consecutive marks; printed per launch."""
import ctypes, importlib, json, os, sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NAMES = {1: "tile loads", 2: "previous panel's update of the diagonal tile (the row tile's runs under the factorisation)", 3: "barrier + trailing update before a factor32",
         4: "factor32 #1", 5: "invert32 #1", 6: "trip-1 barrier wait", 8: "factor32 #2", 9: "invert32 #2",
         10: "trip-2 barrier wait", 12: "X10 = -X11 Y", 13: "L21 product + store"}


def main():
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    lib = importlib.import_module("deep-visual-odometry-slam_b200._lib").lib
    buf = (ctypes.c_longlong * 32)()
    out = []
    for n in [int(a) for a in sys.argv[1:]] or [1000, 4890]:
        g = torch.Generator(device="cuda").manual_seed(n)
        M = torch.randn(n, n + 8, device="cuda", generator=g)
        S = M @ M.t() / (n + 8) + 0.5 * torch.eye(n, device="cuda")
        y = torch.randn(n, device="cuda", generator=g)
        for _ in range(3):
            pkg.cuda_ba.cholesky_solve(S, y)
        lib.dpvo_chol_profile_read(buf, 1)
        reps = 5
        for _ in range(reps):
            pkg.cuda_ba.cholesky_solve(S, y)
        lib.dpvo_chol_profile_read(buf, 1)
        launches = max(int(buf[31]), 1)
        row = {"n": n, "panel_launches_sampled": launches,
               "cycles_per_launch": {NAMES.get(i, str(i)): round(buf[i] / launches) for i in range(1, 14) if buf[i]}}
        row["total_cycles_per_launch"] = sum(row["cycles_per_launch"].values())
        out.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "chol_stage_profile.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
