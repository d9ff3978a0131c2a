"""Replays the scripted frame sequence of tests/golden/make_golden_edges.py through an implementation of the
edge bookkeeping (the numpy oracle on CPU, the device edge list on the GPU) and compares every frame's state
with what the reference's own methods produced (tests/golden/ref_cpu_edges.npz) — bit for bit."""
import importlib.util
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def golden_module():
    spec = importlib.util.spec_from_file_location("make_golden_edges", os.path.join(HERE, "golden", "make_golden_edges.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def golden():
    return np.load(os.path.join(HERE, "golden", "ref_cpu_edges.npz"))


def replay(impl, name, seed, loop_closure):
    """impl: forw(n) back(n) count() set_rows(net, target, weight) remove_frame(k) shift(k, n)
    remove_old(n, window, loop_closure, optimization_window) snapshot() -> dict, final() -> dict."""
    m, g = golden_module(), golden()
    cfg, seq = m.CFG, m.sequence(seed, loop_closure)
    M, n = cfg["PATCHES_PER_FRAME"], 0
    for t in range(m.FRAMES):
        n += 1
        impl.forw(n)
        impl.back(n)
        if n >= 8:
            impl.set_rows(*seq["rows"](impl.count()))
            mm = seq["motion"]() + seq["motion"]()                    # keyframe(), dpvo.py:264-266
            if mm / 2 < cfg["KEYFRAME_THRESH"]:
                k = n - cfg["KEYFRAME_INDEX"]
                impl.remove_frame(k)
                impl.shift(k, n)
                n -= 1
            impl.remove_old(n, cfg["REMOVAL_WINDOW"], loop_closure, cfg["OPTIMIZATION_WINDOW"])
        snap = impl.snapshot()
        assert n == int(g[f"{name}_t{t}_n"])
        for key in ("ii", "jj", "kk", "ii_inac", "jj_inac", "kk_inac"):
            want = g[f"{name}_t{t}_{key}"]
            assert snap[key].shape == want.shape and np.array_equal(snap[key], want), (name, t, key)
    fin = impl.final()
    for key, got in fin.items():
        want = g[f"{name}_final_{key}"]
        assert got.shape == want.shape and got.dtype == want.dtype, (key, got.shape, want.shape, got.dtype, want.dtype)
        assert np.array_equal(got.view(np.uint8), want.view(np.uint8)), (name, "final", key)
    return len(fin)
