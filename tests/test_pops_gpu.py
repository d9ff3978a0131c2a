"""GPU parity tests for the fused projective transform (projective_ops.transform / flow_mag /
point_cloud) through Python surface -> ctypes -> C ABI -> csrc/transform.cu, against
  (a) tests/golden/ref_cpu_pops.npz — the reference's own projective_ops.py on its lietorch CPU backend,
  (b) the float64 restatement (oracle/pops_oracle.py) on the full-size DPVO graph,
  (c) this repo's lietorch kernels composed the way the reference composes them.
Tolerance: fp32 results within 1e-5 relative (norm-wise); validity masks bit-exact away from the
Z = 0.2 threshold.
"""
import os

import numpy as np
import pytest
import torch

from conftest import load_pkg
from oracle import pops_oracle as po

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_cpu_pops.npz")
TOL = 1e-5


def _rel(a, b):
    a = np.asarray(a.cpu() if torch.is_tensor(a) else a, np.float64)
    b = np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def _dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def _args(g):
    return (_dev(g["poses"])[None], _dev(g["patches"])[None], _dev(g["intrinsics"])[None],
            _dev(g["ii"]), _dev(g["jj"]), _dev(g["kk"]))


def test_transform_matches_reference_cpu_golden():
    pops = load_pkg("projective_ops")
    g = np.load(GOLD)
    a = _args(g)
    c = pops.transform(*a)
    assert c.shape == g["coords"].shape and _rel(c, g["coords"]) < TOL
    cd = pops.transform(*a, depth=True)
    assert cd.shape == g["coords_depth"].shape and _rel(cd, g["coords_depth"]) < TOL
    c2, v = pops.transform(*a, valid=True)
    assert torch.equal(c2, c) and np.array_equal(v.cpu().numpy(), g["valid"])
    assert _rel(pops.transform(*a, tonly=True), g["coords_tonly"]) < TOL
    c3, jv, (Ji, Jj, Jz) = pops.transform(*a, jacobian=True)
    assert torch.equal(c3, c) and np.array_equal(jv.cpu().numpy(), g["jvalid"])
    assert Ji.shape == g["Ji"].shape and Jz.shape == g["Jz"].shape
    assert _rel(Ji, g["Ji"]) < TOL and _rel(Jj, g["Jj"]) < TOL and _rel(Jz, g["Jz"]) < TOL


def test_flow_mag_and_point_cloud_match_reference_cpu_golden():
    pops = load_pkg("projective_ops")
    g = np.load(GOLD)
    a = _args(g)
    f, fv = pops.flow_mag(*a, beta=0.5)
    assert f.shape == g["flow"].shape and fv.dtype == torch.bool
    assert _rel(f, g["flow"]) < 2e-4      # differences of nearly equal float32 coordinates
    assert np.array_equal(fv.cpu().numpy(), g["flow_valid"])
    pts = pops.point_cloud(a[0], a[1], a[2], _dev(g["ix"]))
    assert pts.shape == g["points"].shape and _rel(pts, g["points"]) < TOL


def test_motionmag_pair_matches_two_masked_flow_mag_means():
    """DPVO.keyframe's motionmag(i, j) + motionmag(j, i) (dpvo.py:257-264, 266-272) in one launch: against the reference
    formulation (boolean-mask selection, flow_mag of the selected edges, mean) built from pops.flow_mag — itself pinned
    to the reference's projective_ops.py above — and against the float64 oracle.  1e-6 relative (fp32 sums of ~100
    terms in another order); NaN for an empty direction; bit-identical on repetition (it drives a decision)."""
    pops = load_pkg("projective_ops"); synth = load_pkg("synth")
    d = synth.ba_inputs(40, seed=4)
    a = (_dev(d["poses"])[None], _dev(d["patches"])[None], _dev(d["intrinsics"])[None],
         _dev(d["ii"]), _dev(d["jj"]), _dev(d["kk"]))
    ii, jj, kk = a[3], a[4], a[5]
    pairs = torch.stack([ii, jj], 1).unique(dim=0).cpu().numpy()
    both = [(int(i), int(j)) for i, j in pairs if i < j and ((pairs[:, 0] == j) & (pairs[:, 1] == i)).any()]
    assert len(both) >= 3
    for i, j in both[:2] + both[-2:]:
        want = 0.0
        want64 = 0.0
        for x, y in ((i, j), (j, i)):
            k = torch.nonzero((ii == x) & (jj == y)).view(-1)
            f, _ = pops.flow_mag(a[0], a[1], a[2], ii[k], jj[k], kk[k], beta=0.5)
            want += float(f.double().mean())
            kc = k.cpu().numpy()
            f64, _ = po.flow_mag(d["poses"], d["patches"], d["intrinsics"], d["ii"][kc], d["jj"][kc], d["kk"][kc], beta=0.5)
            want64 += float(np.mean(f64))
        got = pops.motionmag_pair(*a, i, j, beta=0.5)
        assert got.dim() == 0 and abs(float(got) - want) <= 1e-6 * abs(want)
        assert abs(float(got) - want64) <= 2e-4 * abs(want64)      # flow_mag's own bar against float64 (above)
        assert torch.equal(got, pops.motionmag_pair(*a, i, j, beta=0.5))
        assert float(pops.motionmag_pair(*a, j, i, beta=0.5)) == float(got)
    # one direction only / no edge at all: NaN like torch's mean of an empty selection; empty edge list too
    only = [(int(i), int(j)) for i, j in pairs if not ((pairs[:, 0] == j) & (pairs[:, 1] == i)).any()]
    if only:
        assert np.isnan(float(pops.motionmag_pair(*a, *only[0], beta=0.5)))
    assert np.isnan(float(pops.motionmag_pair(*a, 10 ** 6, 3, beta=0.5)))
    e = torch.zeros(0, dtype=torch.int64, device="cuda")
    assert np.isnan(float(pops.motionmag_pair(a[0], a[1], a[2], e, e, e, 0, 1)))
    out = torch.zeros(1, device="cuda")
    r = pops.motionmag_pair(*a, *both[0], out=out)
    assert r.data_ptr() == out.data_ptr() and float(out[0]) == float(r)


def test_transform_full_size_vs_oracle_and_lietorch_composition():
    """45,312-edge DPVO graph: against float64, and against Inv/Mul/Act4 of this repo's lietorch
    kernels composed like projective_ops.py:57-68."""
    pops = load_pkg("projective_ops"); lb = load_pkg("lietorch_backends"); synth = load_pkg("synth")
    d = synth.ba_inputs(40, seed=4)
    a = (_dev(d["poses"])[None], _dev(d["patches"])[None], _dev(d["intrinsics"])[None],
         _dev(d["ii"]), _dev(d["jj"]), _dev(d["kk"]))
    got, val = pops.transform(*a, valid=True)
    want, wval = po.transform(d["poses"], d["patches"], d["intrinsics"], d["ii"], d["jj"], d["kk"], valid=True)
    assert _rel(got, want) < TOL
    Z = None
    # composition with the lietorch kernels
    G = lb.mul(3, a[0][0, a[4]].contiguous(), lb.inv(3, a[0][0, a[3]].contiguous()))
    K = a[2][0]
    pt = a[1][0, a[5]]                                                     # [E,3,P,P]
    fx, fy, cx, cy = (K[a[3], k][:, None, None] for k in range(4))
    X0 = torch.stack([(pt[:, 0] - cx) / fx, (pt[:, 1] - cy) / fy, torch.ones_like(pt[:, 2]), pt[:, 2]], -1)
    E, P = X0.shape[0], X0.shape[1]
    X1 = lb.act4(3, G[:, None, None, :].expand(E, P, P, 7).reshape(-1, 7).contiguous(), X0.reshape(-1, 4).contiguous())
    X1 = X1.view(E, P, P, 4)
    fxj, fyj, cxj, cyj = (K[a[4], k][:, None, None] for k in range(4))
    dd = 1.0 / X1[..., 2].clamp(min=0.1)
    comp = torch.stack([fxj * (dd * X1[..., 0]) + cxj, fyj * (dd * X1[..., 1]) + cyj], -1)[None]
    assert _rel(got, comp.cpu().numpy()) < 2e-6
    Z = X1[..., 2]
    away = (Z - 0.2).abs() > 1e-4
    assert torch.equal(val[0][away], (Z > 0.2).float()[away])
    assert np.mean(val.cpu().numpy() == wval) > 0.9999


def test_transform_edge_cases():
    pops = load_pkg("projective_ops")
    g = np.load(GOLD)
    a = _args(g)
    e = torch.zeros(0, dtype=torch.int64, device="cuda")
    assert pops.transform(a[0], a[1], a[2], e, e, e).shape == (1, 0, 3, 3, 2)
    bad = a[3].clone(); bad[5] = 10 ** 6                       # out-of-range frame: zeros, nothing read
    c = pops.transform(a[0], a[1], a[2], bad, a[4], a[5])
    assert float(c[0, 5].abs().max()) == 0.0 and torch.equal(c[0, 6:], pops.transform(*a)[0, 6:])
    with pytest.raises(RuntimeError):
        pops.transform(a[0].cpu(), a[1], a[2], a[3], a[4], a[5])
    with pytest.raises(RuntimeError):
        pops.transform(a[0], a[1], a[2], a[3].int(), a[4], a[5])
    with pytest.raises(RuntimeError):
        with torch.enable_grad():
            pops.transform(a[0].clone().requires_grad_(True), a[1], a[2], a[3], a[4], a[5])
