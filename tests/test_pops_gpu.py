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


@pytest.mark.parametrize("loop_closure", [False, True])
def test_keyframe_plan_counts_match_the_reference_rules(loop_closure):
    """`keyframe_plan` = motionmag_pair + a dry run of keyframe()'s edge edits (dpvo.py:272-310) under both outcomes:
    the counts / minima must equal the reference's boolean-mask formulation on a random edge list that exercises every
    rule (edges touching the removed frame, renumbering, the removal window, the loop-closure exemption).  Integers:
    exact; word 0 is bit-identical to motionmag_pair."""
    pops = load_pkg("projective_ops"); synth = load_pkg("synth")
    d = synth.ba_inputs(48, seed=9)
    M = int(d["M"])
    poses, patches, intr = _dev(d["poses"])[None], _dev(d["patches"])[None], _dev(d["intrinsics"])[None]
    N = poses.shape[1]
    assert patches.shape[1] == N * M
    ix = torch.arange(N, device="cuda").repeat_interleave(M)
    g = torch.Generator().manual_seed(5)
    E = 30000
    ii = torch.randint(0, 48, (E,), generator=g)
    jj = torch.randint(0, 48, (E,), generator=g)
    kk = ii * M + torch.randint(0, M, (E,), generator=g)
    ii, jj, kk = ii.cuda(), jj.cuda(), kk.cuda()
    n, k, RW, OW = 48, 44, 22, 12
    i, j = k - 1, k + 1
    out = pops.keyframe_plan(poses, patches, intr, ii, jj, kk, ix, i, j, k, n, M, RW, loop_closure, OW, beta=0.5)
    h = out.cpu()
    assert h.dtype == torch.int32 and h.numel() == 8
    mm = pops.motionmag_pair(poses, patches, intr, ii, jj, kk, i, j, beta=0.5)
    assert h[:1].view(torch.float32)[0].item() == float(mm) and np.isfinite(float(mm))

    def rule(src, a, b, nn):
        rem = src < nn - RW
        if loop_closure:
            rem = rem & ~(((b - a) > 30) & (b > nn - OW))
        return rem

    rem_a = rule(ix[kk], ii, jj, n)
    assert int(h[1]) == int(rem_a.sum()) and int(h[2]) == int(ii[~rem_a].min())
    drop = (ii == k) | (jj == k)
    i2 = ii - (ii > k).long(); j2 = jj - (jj > k).long(); k2 = kk - M * (ii > k).long()
    rem_b = rule(ix[k2], i2, j2, n - 1) & ~drop
    keep_b = ~drop & ~rem_b
    assert int(h[3]) == int(drop.sum()) and int(h[4]) == int(rem_b.sum()) and int(h[5]) == int(i2[keep_b].min())
    assert 0 < int(h[1]) < E and 0 < int(h[3]) < E and 0 < int(h[4]) < E
    # ... and the oracle's restatement, pinned on CPU to the reference-run sequence (tests/test_oracle_cpu.py)
    from oracle import edges_oracle as eo
    want = eo.keyframe_plan_counts(ix.cpu().numpy(), M, ii.cpu().numpy(), jj.cpu().numpy(), kk.cpu().numpy(), k, n, RW,
                                   loop_closure, OW)
    assert [int(v) for v in h[1:6]] == [pops.PLAN_NONE if v is None else v for v in want]
    if loop_closure:                                                             # the exemption is exercised
        assert int(((ix[kk] < n - RW) & ((jj - ii) > 30) & (jj > n - OW)).sum()) > 0
    # words 6 / 7 echo the edge store's device count and error flag when the caller passes them
    assert h[6:8].tolist() == [0, 0]
    lc_t = torch.tensor([E], dtype=torch.int64, device="cuda"); inf_t = torch.tensor([2], dtype=torch.int32, device="cuda")
    h2 = pops.keyframe_plan(poses, patches, intr, ii, jj, kk, ix, i, j, k, n, M, RW, loop_closure, OW, live_count=lc_t,
                            info=inf_t).cpu()
    assert h2[6:8].tolist() == [E, 2] and torch.equal(h2[:6], h[:6])
    # nothing left alive: the minima read "none"
    e0 = torch.zeros(0, dtype=torch.int64, device="cuda")
    h0 = pops.keyframe_plan(poses, patches, intr, e0, e0, e0, ix, i, j, k, n, M, RW, loop_closure, OW).cpu()
    assert h0[1:6].tolist() == [0, pops.PLAN_NONE, 0, 0, pops.PLAN_NONE] and np.isnan(h0[:1].view(torch.float32)[0].item())
    # repeatable (it drives a decision and the host's edge counts)
    assert torch.equal(out[:6], pops.keyframe_plan(poses, patches, intr, ii, jj, kk, ix, i, j, k, n, M, RW, loop_closure, OW)[:6])


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
