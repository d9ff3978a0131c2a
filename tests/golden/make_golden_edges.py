"""Golden vectors for the patch-graph edge bookkeeping (N4) from the UNMODIFIED reference code on CPU.

    python tests/golden/make_golden_edges.py     (needs /root/reference)

dpvo/dpvo.py cannot be imported without a GPU (it builds `SE3.Identity(device="cuda")` at import), so the
method definitions `DPVO.append_factors`, `remove_factors`, `keyframe`, `__edges_forw`, `__edges_back`
(dpvo.py:215-239, 262-310, 362-375) are taken from the file's syntax tree at generation time, compiled
as they stand into a class of the same name and run on CPU tensors.  Substitutions, all outside the code
under test: `torch.arange/zeros(..., device="cuda")` lose the device argument; `SE3` is a placeholder (the
relative pose keyframe() stores in `pg.delta` is not edge bookkeeping); `motionmag` returns the scripted
values of `sequence()` so that some frames are dropped.  `reduce_edges` / `_format`
(loop_closure/optim_utils.py:22-64) are compiled the same way with `nb.njit` as the identity decorator.

Nothing of the reference is stored in this repository: only the outputs (tests/golden/ref_cpu_edges.npz).
"""
import ast
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/methods/dpvo/dpvo"

# small patch graph: every code path of keyframe() is reached within 40 frames
CFG = dict(BUFFER_SIZE=64, PATCHES_PER_FRAME=4, PATCH_LIFETIME=5, REMOVAL_WINDOW=8, OPTIMIZATION_WINDOW=5,
           KEYFRAME_INDEX=4, KEYFRAME_THRESH=12.5, CLASSIC_LOOP_CLOSURE=False)
DIM, MEM, PMEM, FRAMES = 8, 12, 16, 40
FM1, FM2, GM = (6, 5, 4), (6, 2, 2), (6, 3, 3)      # per-slot shapes of the fmap1 / fmap2 / gmap rings


def sequence(seed, loop_closure):
    """Scripted per-frame data shared by the golden script and the tests: the motion magnitudes keyframe()
    sees and the network outputs update() would have written (net fp16, target, weight), drawn on demand."""
    rng = np.random.default_rng(seed)
    return dict(rng=rng, loop_closure=bool(loop_closure),
                motion=lambda: float(rng.choice([3.0, 30.0], p=[0.45, 0.55])),
                rows=lambda e: (rng.standard_normal((e, DIM)).astype(np.float16),
                                rng.standard_normal((e, 2)).astype(np.float32),
                                rng.random((e, 2)).astype(np.float32)))


def initial_buffers(seed):
    """Frame-indexed state keyframe() shifts (dpvo.py:287-297); content is arbitrary but distinct per slot."""
    rng = np.random.default_rng(1000 + seed)
    N, M = CFG["BUFFER_SIZE"], CFG["PATCHES_PER_FRAME"]
    return dict(tstamps_=np.arange(N, dtype=np.int64) * 3,
                colors_=rng.integers(0, 255, (N, M, 3)).astype(np.uint8),
                poses_=rng.standard_normal((N, 7)).astype(np.float32),
                patches_=rng.standard_normal((N, M, 3, 3, 3)).astype(np.float32),
                intrinsics_=rng.standard_normal((N, 4)).astype(np.float32),
                imap_=rng.standard_normal((PMEM, M, DIM)).astype(np.float16),
                gmap_=rng.standard_normal((PMEM, M) + GM).astype(np.float16),
                fmap1_=rng.standard_normal((1, MEM) + FM1).astype(np.float16),
                fmap2_=rng.standard_normal((1, MEM) + FM2).astype(np.float16))


REDUCE_CASES = [dict(seed=0, n=600, frames=120, max_num_edges=1000, nms=1),
                dict(seed=1, n=300, frames=90, max_num_edges=5, nms=2),
                dict(seed=2, n=50, frames=40, max_num_edges=1000, nms=0),
                dict(seed=3, n=0, frames=10, max_num_edges=10, nms=1)]


def reduce_case(c):
    rng = np.random.default_rng(50 + c["seed"])
    n, f = c["n"], c["frames"]
    ii = rng.integers(0, f, n).astype(np.int64)
    jj = rng.integers(0, f, n).astype(np.int64)
    mag = (rng.random(n) * 64).astype(np.float32)
    if n:
        mag[rng.integers(0, n, max(n // 20, 1))] = 2000.0           # the reference treats >= 1000 as infinite
    return mag, ii, jj


# ------------------------------------------------------------------ generation (needs /root/reference)
def _reference_class():
    import torch

    class _Torch:
        """torch with the `device="cuda"` of arange / zeros dropped."""
        def __getattr__(self, name):
            f = getattr(torch, name)
            if name in ("arange", "zeros", "ones", "empty"):
                return lambda *a, **k: f(*a, **{x: y for x, y in k.items() if x != "device"})
            return f

    class _SE3:
        def __init__(self, data): self.data = data
        def __mul__(self, other): return self
        def inv(self): return self

    tree = ast.parse(open(os.path.join(REF, "dpvo.py")).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "DPVO")
    want = {"append_factors", "remove_factors", "keyframe", "__edges_forw", "__edges_back"}
    cls.body = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in want]
    assert {n.name for n in cls.body} == want
    cls.bases, cls.decorator_list = [], []
    utree = ast.parse(open(os.path.join(REF, "utils.py")).read())
    fm = next(n for n in utree.body if isinstance(n, ast.FunctionDef) and n.name == "flatmeshgrid")
    mod = ast.Module(body=[fm, cls], type_ignores=[])
    ns = {"torch": _Torch(), "SE3": _SE3}
    exec(compile(mod, "reference:dpvo.py", "exec"), ns)
    return ns["DPVO"]


def _reference_reduce():
    tree = ast.parse(open(os.path.join(REF, "loop_closure", "optim_utils.py")).read())
    fns = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("_format", "reduce_edges")]
    assert len(fns) == 2
    nb = types.SimpleNamespace(njit=lambda **k: (lambda f: f), bool_=np.bool_)
    ns = {"np": np, "nb": nb}
    exec(compile(ast.Module(body=fns, type_ignores=[]), "reference:optim_utils.py", "exec"), ns)
    return ns["reduce_edges"]


def run_reference(seed, loop_closure):
    import torch
    DPVO = _reference_class()
    seq = sequence(seed, loop_closure)
    cfg = types.SimpleNamespace(LOOP_CLOSURE=seq["loop_closure"], **CFG)
    N, M = CFG["BUFFER_SIZE"], CFG["PATCHES_PER_FRAME"]
    buf = {k: (v if k == "tstamps_" else torch.from_numpy(v)) for k, v in initial_buffers(seed).items()}
    e = lambda *s, dt=torch.long: torch.zeros(*s, dtype=dt)
    pg = types.SimpleNamespace(ii=e(0), jj=e(0), kk=e(0), net=e(1, 0, DIM, dt=torch.half),
                               weight=e(1, 0, 2, dt=torch.float), target=e(1, 0, 2, dt=torch.float),
                               ii_inac=e(0), jj_inac=e(0), kk_inac=e(0), weight_inac=e(1, 0, 2, dt=torch.float),
                               target_inac=e(1, 0, 2, dt=torch.float), delta={},
                               index_=torch.arange(N)[:, None].repeat(1, M).contiguous(),
                               **{k: buf[k] for k in ("tstamps_", "colors_", "poses_", "patches_", "intrinsics_")})
    d = DPVO.__new__(DPVO)
    d.cfg, d.pg, d.M, d.DIM, d.mem, d.pmem = cfg, pg, M, DIM, MEM, PMEM
    d.kwargs = {"device": "cpu", "dtype": torch.half}
    d.n = d.m = 0
    d.ix = pg.index_.view(-1)
    d.imap_, d.gmap_, d.fmap1_, d.fmap2_ = buf["imap_"], buf["gmap_"], buf["fmap1_"], buf["fmap2_"]
    d.motionmag = lambda i, j: seq["motion"]()
    out = {}
    for t in range(FRAMES):
        d.n += 1
        d.m += M
        d.append_factors(*d._DPVO__edges_forw())
        d.append_factors(*d._DPVO__edges_back())
        if d.n >= 8:                                            # update() then keyframe(), dpvo.py:467-469
            net, target, weight = seq["rows"](pg.ii.numel())
            pg.net, pg.target, pg.weight = (torch.from_numpy(x)[None] for x in (net, target, weight))
            d.keyframe()
        for k in ("ii", "jj", "kk", "ii_inac", "jj_inac", "kk_inac"):
            out[f"t{t}_{k}"] = getattr(pg, k).numpy().copy()
        out[f"t{t}_n"] = np.int64(d.n)
    for k in ("net", "target", "weight", "target_inac", "weight_inac"):
        out[f"final_{k}"] = getattr(pg, k)[0].numpy().copy()
    for k in ("tstamps_", "colors_", "poses_", "patches_", "intrinsics_"):
        v = getattr(pg, k)
        out[f"final_{k}"] = (v if isinstance(v, np.ndarray) else v.numpy()).copy()
    for k in ("imap_", "gmap_", "fmap1_", "fmap2_"):
        out[f"final_{k}"] = getattr(d, k).numpy().copy()
    return out


def main():
    out = {}
    for name, (seed, lc) in dict(plain=(0, False), loop=(1, True)).items():
        for k, v in run_reference(seed, lc).items():
            out[f"{name}_{k}"] = v
    reduce_edges = _reference_reduce()
    for i, c in enumerate(REDUCE_CASES):
        mag, ii, jj = reduce_case(c)
        out[f"reduce{i}"] = np.asarray(reduce_edges(mag, ii, jj, c["max_num_edges"], c["nms"]), dtype=np.int64).reshape(-1, 2)
    p = os.path.join(HERE, "ref_cpu_edges.npz")
    np.savez_compressed(p, **out)
    print(p, os.path.getsize(p), "bytes;", "frames kept:", out["plain_t%d_n" % (FRAMES - 1)], out["loop_t%d_n" % (FRAMES - 1)],
          "edges:", out["plain_t%d_ii" % (FRAMES - 1)].size, "inactive:", out["plain_t%d_ii_inac" % (FRAMES - 1)].size,
          "reduce:", [out[f"reduce{i}"].shape[0] for i in range(len(REDUCE_CASES))])


if __name__ == "__main__":
    sys.exit(main())
