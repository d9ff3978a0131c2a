"""Kernel-level timing of the fused 2-level correlation lookup (development tool).

    python tools/time_corr.py [--edges 45312] [--iters 20] [--ref]

Prints one JSON line per configuration: CUDA-event time, algorithmic bytes (SURVEY §8d,
computed exactly from the coords) and achieved GB/s.  --ref also times the reference's
own extension (oracle/_ref/cuda_corr.so) on the same inputs.
"""
import argparse
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def algorithmic_bytes(*a, **k):
    import importlib
    return importlib.import_module("deep-visual-odometry-slam_b200.synth").corr_algorithmic_bytes(*a, **k)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--edges", type=int, default=0, help="0 = DPVO steady-state graph (45,312)")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--ref", action="store_true")
    a = ap.parse_args()
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
    edges = synth.dpvo_edges(40) if a.edges == 0 else None
    d = synth.corr_inputs(a.edges, seed=1, edges=edges, oob_frac=0.02)
    E = len(d["us"])
    nbytes = algorithmic_bytes(d["coords"], [(120, 160), (30, 40)], [1.0, 0.25])
    dev = lambda x: torch.from_numpy(x).cuda()
    g = dev(d["gmap"]); f1n = dev(d["fmap1"]); f2n = dev(d["fmap2"])
    f1 = f1n.permute(0, 1, 3, 4, 2).contiguous().permute(0, 1, 4, 2, 3)
    f2 = f2n.permute(0, 1, 3, 4, 2).contiguous().permute(0, 1, 4, 2, 3)
    c = dev(d["coords"]); us = dev(d["us"]); vs = dev(d["vs"])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def timeit(fn, iters):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(iters):
            flush.zero_()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts)), float(np.min(ts))

    cc = pkg.cuda_corr
    outs = {}
    for label, flag in (("b200 fused 2-level, append order", False), ("b200 fused 2-level, locality order (sort included)", True)):
        cc.ORDER_FORWARD = flag
        outs[flag] = pkg.altcorr.corr_pyramid(g, (f1, f2), c, us, vs, 3)
        med, mn = timeit(lambda: pkg.altcorr.corr_pyramid(g, (f1, f2), c, us, vs, 3), a.iters)
        print(json.dumps({"impl": label, "edges": E, "ms_median": med, "ms_min": mn,
                          "alg_bytes": nbytes, "alg_GBs": nbytes / med / 1e6, "bytes_per_edge": nbytes / E}))
    print(json.dumps({"ordered_equals_unordered_bitwise": bool(torch.equal(outs[True], outs[False]))}))
    med, mn = timeit(lambda: cc.edge_order(c, vs, f1.shape[1], 120, 160, 32), a.iters)
    print(json.dumps({"impl": "edge_order alone (3 launches)", "ms_median": med, "ms_min": mn}))
    for tile in (16, 32, 64):
        perm = cc.edge_order(c, vs, f1.shape[1], 120, 160, tile)
        real = cc.edge_order
        cc.edge_order = lambda *a_, **k_: perm
        cc.ORDER_FORWARD = True
        med, mn = timeit(lambda: pkg.altcorr.corr_pyramid(g, (f1, f2), c, us, vs, 3), a.iters)
        cc.edge_order = real
        print(json.dumps({"impl": f"lookup kernel alone, precomputed locality order, tile {tile}", "ms_median": med, "ms_min": mn,
                          "alg_GBs": nbytes / med / 1e6}))
    if a.ref:
        from oracle.ref_ext import load_ref
        ref = load_ref("cuda_corr")
        if ref is not None:
            def ref_fn():
                c1, = ref.forward(g, f1n, c, us, vs, 3)
                c2, = ref.forward(g, f2n, c / 4, us, vs, 3)
                return torch.stack([c1, c2], -1).view(1, E, -1)
            med, mn = timeit(ref_fn, max(3, a.iters // 4))
            print(json.dumps({"impl": "reference ext (sm_100 rebuild)", "edges": E, "ms_median": med, "ms_min": mn,
                              "alg_GBs": nbytes / med / 1e6}))


if __name__ == "__main__":
    main()
