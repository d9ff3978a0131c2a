"""GPU parity for cuda_ba.solve_system (classic loop-closure PGO step, fastba/ba.cpp:120-180) through the
C ABI (dpvo_pgo_solve): against the reference extension's own outputs (tests/golden/ref_cpu_pgo.npz,
made on CPU by tests/golden/make_golden_pgo.py).  The reference solves in double and returns float32;
the device path (float64 normal equations, float32 blocked Cholesky, float64-residual refinement)
must agree to 1e-4 relative (norm-wise) with the golden."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from conftest import load_pkg

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(__file__)


def _cases():
    spec = importlib.util.spec_from_file_location("make_golden_pgo", os.path.join(HERE, "golden", "make_golden_pgo.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


@pytest.mark.parametrize("name", ["small", "chain40", "free_all", "chain300"])
def test_solve_system_vs_reference(name):
    pkg = load_pkg()
    m = _cases()
    g = np.load(os.path.join(HERE, "golden", "ref_cpu_pgo.npz"))
    n, loops, freen, seed = m.CASES[name]
    Ji, Jj, ii, jj, res = m.pgo_case(n, loops, seed)
    dev = lambda x: torch.from_numpy(x).cuda()
    delta, info = pkg.cuda_ba.solve_system(dev(Ji), dev(Jj), dev(ii), dev(jj), dev(res), m.EP, m.LM, freen,
                                           return_info=True)
    assert int(info.item()) == 0
    assert delta.shape == g[name].shape and delta.dtype == torch.float32
    got = delta.cpu().numpy()
    assert _rel(got, g[name]) < 1e-4, (name, _rel(got, g[name]))
    if freen >= 0:
        assert (got[freen:] == 0).all()                  # fixed poses do not move
    # the plain list return of the reference signature
    out = pkg.cuda_ba.solve_system(dev(Ji), dev(Jj), dev(ii), dev(jj), dev(res), m.EP, m.LM, freen)
    assert isinstance(out, list) and len(out) == 1 and torch.equal(out[0], delta)


def test_solve_system_refinement_matters_and_errors():
    pkg = load_pkg()
    m = _cases()
    g = np.load(os.path.join(HERE, "golden", "ref_cpu_pgo.npz"))
    n, loops, freen, seed = m.CASES["chain300"]
    Ji, Jj, ii, jj, res = m.pgo_case(n, loops, seed)
    dev = lambda x: torch.from_numpy(x).cuda()
    e = []
    for it in (0, 2):
        d, = pkg.cuda_ba.solve_system(dev(Ji), dev(Jj), dev(ii), dev(jj), dev(res), m.EP, m.LM, freen, refinements=it)
        e.append(_rel(d.cpu().numpy(), g["chain300"]))
    assert e[1] <= e[0] and e[1] < 1e-4
    # self edge: the reference exit(1)s; here info = 2 and a zero step
    jj_bad = jj.copy(); jj_bad[3] = ii[3]
    d, info = pkg.cuda_ba.solve_system(dev(Ji), dev(Jj), dev(ii), dev(jj_bad), dev(res), m.EP, m.LM, freen,
                                       return_info=True)
    assert int(info.item()) == 2 and float(d.abs().max()) == 0.0
    with pytest.raises(RuntimeError):
        pkg.cuda_ba.solve_system(torch.from_numpy(Ji), torch.from_numpy(Jj), torch.from_numpy(ii), torch.from_numpy(jj),
                                 torch.from_numpy(res), m.EP, m.LM, freen)
    # timing-free sanity at a larger size: 1,000 poses, all free but the last 5
    Ji, Jj, ii, jj, res = m.pgo_case(1000, 300, 9)
    d, info = pkg.cuda_ba.solve_system(dev(Ji), dev(Jj), dev(ii), dev(jj), dev(res), m.EP, m.LM, 995, return_info=True)
    assert int(info.item()) == 0 and torch.isfinite(d).all() and (d[995:] == 0).all()
