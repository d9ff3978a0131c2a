"""Timing of the correlation backward (training path, BASELINE config 4 shape: 15-frame clip, 80
patches/frame, ~18k edges, AMP fp16 features) for this repo and, with --ref, the reference's
extension on identical inputs (development / measurement tool).
    python tools/time_corr_bwd.py [--ref] [--dtype half|float]
"""
import argparse, importlib, json, os, sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", action="store_true")
    ap.add_argument("--dtype", default="half")
    ap.add_argument("--edges", type=int, default=18000)
    a = ap.parse_args()
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
    d = synth.corr_inputs(a.edges, seed=2, mem=15, n_patches=15 * 80, oob_frac=0.02)
    dt = torch.float16 if a.dtype == "half" else torch.float32
    dev = lambda x: torch.from_numpy(x).cuda()
    g = dev(d["gmap"]).to(dt); c = dev(d["coords"]); us = dev(d["us"]); vs = dev(d["vs"])
    rows = []
    for lvl, (name, scale) in enumerate((("fmap1", 1.0), ("fmap2", 0.25))):
        f = dev(d[name]).to(dt)
        cc = (c * scale).contiguous()
        grad = torch.randn(1, a.edges, 7, 7, 3, 3, device="cuda", dtype=dt)

        def timeit(fn, iters=10):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            ts = []
            for _ in range(iters):
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            return float(np.median(ts))

        pkg.cuda_corr.ORDER_EDGES = False
        ms_append = timeit(lambda: pkg.cuda_corr.backward(g, f, cc, us, vs, grad, 3))
        pkg.cuda_corr.ORDER_EDGES = True
        row = {"op": "corr_backward", "level": lvl, "dtype": a.dtype, "edges": a.edges,
               "ms_b200": timeit(lambda: pkg.cuda_corr.backward(g, f, cc, us, vs, grad, 3)),
               "ms_b200_append_order": ms_append,
               "ms_b200_forward": timeit(lambda: pkg.cuda_corr.forward(g, f, cc, us, vs, 3))}
        if a.ref:
            from oracle.ref_ext import load_ref
            ref = load_ref("cuda_corr")
            if ref is not None:
                row["ms_reference_ext"] = timeit(lambda: ref.backward(g, f, cc, us, vs, grad, 3), iters=5)
                row["ms_reference_ext_forward"] = timeit(lambda: ref.forward(g, f, cc, us, vs, 3), iters=5)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
