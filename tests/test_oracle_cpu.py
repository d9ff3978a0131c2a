"""CPU suite, part 1: the oracle (oracle/*.py) against the golden vectors produced by the reference's
own CPU code (tests/golden/ref_cpu_*.npz, generator: tests/golden/make_golden_cpu.py) and against
brute-force restatements on tiny inputs.  No GPU, no /root/reference needed."""
import os

import numpy as np
import pytest

from conftest import load_pkg
from oracle import altcorr_oracle as oc
from oracle import fastba_oracle as fo
from oracle import lietorch_oracle as lo

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _relmax(a, b):
    return np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / max(np.abs(b).max(), 1e-12)


# ------------------------------------------------------------------ lietorch SE3 vs reference CPU backend
@pytest.fixture(scope="module")
def se3():
    return np.load(os.path.join(GOLD, "ref_cpu_se3.npz"))


def test_se3_forward_ops_match_reference_cpu(se3):
    g = se3
    X, Y = g["X"], g["Y"]
    tol = 2e-6   # the reference computes in float32
    assert _relmax(lo.exp(g["xi"]), X) < tol
    assert _relmax(lo.log(X), g["log"]) < tol
    assert _relmax(lo.inv(X), g["inv"]) < tol
    assert _relmax(lo.mul(X, Y), g["mul"]) < tol
    assert _relmax(lo.adj(X, g["a6"]), g["adj"]) < tol
    assert _relmax(lo.adjT(X, g["a6"]), g["adjT"]) < tol
    assert _relmax(lo.act(X, g["p3"]), g["act"]) < tol
    assert _relmax(lo.act4(X, g["p4"]), g["act4"]) < tol
    assert _relmax(lo.matrix(X), g["matrix"]) < tol
    assert _relmax(lo.projector(X), g["projector"]) < tol


def test_se3_backward_ops_match_reference_cpu(se3):
    g = se3
    X, Y = g["X"], g["Y"]
    tol = 2e-6
    assert _relmax(lo.inv_backward(g["gX"], X), g["inv_bwd"]) < tol
    a, b = lo.mul_backward(g["gX"], X, Y)
    assert _relmax(a, g["mul_bwd_X"]) < tol and _relmax(b, g["mul_bwd_Y"]) < tol
    a, b = lo.act_backward(g["g3"], X, g["p3"])
    assert _relmax(a, g["act_bwd_X"]) < tol and _relmax(b, g["act_bwd_p"]) < tol
    a, b = lo.act4_backward(g["g4"], X, g["p4"])
    assert _relmax(a, g["act4_bwd_X"]) < tol and _relmax(b, g["act4_bwd_p"]) < tol
    a, b = lo.adj_backward(g["g6"], X, g["a6"])
    assert _relmax(a, g["adj_bwd_X"]) < tol and _relmax(b, g["adj_bwd_a"]) < tol
    a, b = lo.adjT_backward(g["g6"], X, g["a6"])
    assert _relmax(a, g["adjT_bwd_X"]) < tol and _relmax(b, g["adjT_bwd_a"]) < tol


def test_se3_q_dependent_ops_match_reference_cpu(se3):
    """exp/log backward and Jinv go through se3.h:144-173 (Q).  Its closed-form coefficients cancel
    catastrophically in float32 just above the series threshold (theta ~ 1e-6..1e-3), so the reference's
    own float32 values are noise there: those samples are excluded, everything else must match."""
    g = se3
    th = np.linalg.norm(g["xi"][:, 3:], axis=1)
    ok = (th < lo.EPS) | (th > 1e-2)
    assert ok.sum() >= 40
    X = g["X"]
    tol = 5e-6
    assert _relmax(lo.jinv(X, g["a6"])[ok], g["Jinv"][ok]) < tol
    assert _relmax(lo.exp_backward(g["gX"], g["xi"])[ok], g["expm_bwd"][ok]) < tol
    assert _relmax(lo.log_backward(g["g6"], X)[ok], g["logm_bwd"][ok]) < tol


def test_se3_identities():
    """The reference's own property tests (lietorch/run_tests.py:16-120), in real float64."""
    rng = np.random.default_rng(0)
    a = rng.normal(0, 0.7, (64, 6))
    X = lo.exp(a)
    assert np.abs(lo.log(X) - a).max() < 1e-9
    I = lo.mul(X, lo.inv(X))
    assert np.abs(I - np.array([0, 0, 0, 0, 0, 0, 1.0])).max() < 1e-12
    b = rng.normal(0, 0.5, (64, 6))
    lhs = lo.mul(X, lo.exp(b))
    rhs = lo.mul(lo.exp(lo.adj(X, b)), X)
    assert np.abs(lhs - rhs).max() < 1e-9
    p = rng.normal(0, 1, (64, 3))
    T = lo.matrix(X)
    assert np.abs(lo.act(X, p) - (T[:, :3, :3] @ p[..., None])[..., 0] - T[:, :3, 3]).max() < 1e-12
    # left Jacobian: finite differences of Exp
    eps = 1e-6
    J = lo.left_jacobian(a)
    for k in range(6):
        d = np.zeros(6); d[k] = eps
        num = lo.log(lo.mul(lo.exp(a + d), lo.inv(lo.exp(a)))) / eps
        assert np.abs(num - J[:, :, k]).max() < 1e-4
    assert np.abs(lo.left_jacobian(a) @ lo.left_jacobian_inverse(a) - np.eye(6)).max() < 1e-9


# ------------------------------------------------------------------ reprojection / Jacobians vs projective_ops
@pytest.fixture(scope="module")
def tr():
    return np.load(os.path.join(GOLD, "ref_cpu_transform.npz"))


def test_reproject_matches_reference_transform(tr):
    want = tr["coords"]                                   # [1,E,3,3,2] (projective_ops.py:53-68)
    got = fo.reproject(tr["poses"], tr["patches"], tr["intrinsics"], tr["ii"], tr["jj"], tr["kk"])
    got = np.transpose(got, (0, 1, 3, 4, 2))
    assert got.shape == want.shape
    assert _relmax(got, want) < 2e-6


def test_ba_jacobians_match_reference_transform(tr):
    """fastba's analytic Ji/Jj/Jz (ba_cuda.cu:276-333) equal projective_ops' (projective_ops.py:71-108)
    wherever the latter's clamps are inactive (Z > 0.2)."""
    r, w, Ji, Jj, Jz = fo.edge_terms(tr["poses"].astype(np.float64), tr["patches"].astype(np.float64),
                                     tr["intrinsics"].astype(np.float64), tr["target"].astype(np.float64),
                                     tr["weight"].astype(np.float64), tr["ii"], tr["jj"], tr["kk"])
    v = tr["valid"][0] > 0.5
    assert v.mean() > 0.9
    # fastba keeps Ji = +Adj^T(Jj) and applies the minus sign when accumulating (ba_cuda.cu:337-369);
    # projective_ops returns the signed derivative Ji = -adjT(Jj) (projective_ops.py:104)
    assert _relmax(-Ji[v], tr["Ji"][0][v]) < 1e-5
    assert _relmax(Jj[v], tr["Jj"][0][v]) < 1e-5
    assert _relmax(Jz[v], tr["Jz"][0][v][..., 0]) < 1e-5


def test_ba_oracle_jacobians_finite_differences():
    """r = target - proj;  d(proj)/d(xi_j) = Jj, d(proj)/d(xi_i) = Ji for left perturbations
    G <- Exp(xi) G (the retraction of ba_cuda.cu:178-206), d(proj)/d(depth) = Jz."""
    synth = load_pkg("synth")
    d = synth.ba_inputs(9, seed=4, M=6)
    P = d["poses"].astype(np.float64); Z = d["patches"].astype(np.float64)
    K = d["intrinsics"].astype(np.float64)
    n = len(P)
    tw = (d["target"].astype(np.float64), d["weight"].astype(np.float64))
    ii, jj, kk = d["ii"], d["jj"], d["kk"]
    P2 = np.concatenate([P, P])          # frames n..2n-1: the copy that gets perturbed
    K2 = np.concatenate([K, K])
    r0, w, Ji, Jj, Jz = fo.edge_terms(P2, Z, K2, *tw, ii, jj + n, kk)
    eps = 1e-6
    Z1 = Z.copy(); Z1[:, 2] += eps
    r1 = fo.edge_terms(P2, Z1, K2, *tw, ii, jj + n, kk)[0]
    assert np.abs(-(r1 - r0) / eps - Jz).max() < 1e-4 * max(np.abs(Jz).max(), 1)

    def perturbed(k):
        dxi = np.zeros(6); dxi[k] = eps
        Q = P2.copy()
        for f in range(n, 2 * n):
            Q[f, :3], Q[f, 3:] = fo._retr_se3(dxi, P2[f, :3], P2[f, 3:])
        return Q
    for k in range(6):
        Q = perturbed(k)
        rj = fo.edge_terms(Q, Z, K2, *tw, ii, jj + n, kk)[0]          # role j perturbed
        assert np.abs(-(rj - r0) / eps - Jj[:, :, k]).max() < 1e-3 * max(np.abs(Jj).max(), 1)
        ri = fo.edge_terms(Q, Z, K2, *tw, ii + n, jj, kk)[0]          # role i perturbed
        assert np.abs(-(ri - r0) / eps + Ji[:, :, k]).max() < 1e-3 * max(np.abs(Ji).max(), 1)   # d/dxi_i = -Ji


# ------------------------------------------------------------------ correlation oracle vs brute force
def test_corr_oracle_vs_python_loops():
    rng = np.random.default_rng(1)
    B, N1, N2, C, P, H, W, M, R = 1, 3, 2, 5, 3, 9, 11, 4, 1
    f1 = rng.standard_normal((B, N1, C, P, P)).astype(np.float32)
    f2 = rng.standard_normal((B, N2, C, H, W)).astype(np.float32)
    coords = np.stack([rng.uniform(-2, W + 1, (B, M, P, P)), rng.uniform(-2, H + 1, (B, M, P, P))], 2).astype(np.float32)
    us = rng.integers(0, N1, M); vs = rng.integers(0, N2, M)
    got = oc.corr_forward(f1, f2, coords, us, vs, R)
    D = 2 * R + 2
    for m in range(M):
        for i0 in range(P):
            for j0 in range(P):
                x, y = coords[0, m, 0, i0, j0], coords[0, m, 1, i0, j0]
                fx, fy = int(np.floor(x)), int(np.floor(y))
                raw = np.zeros((D, D), np.float32)
                for a in range(D):
                    for b in range(D):
                        i1, j1 = fy + a - R, fx + b - R
                        if 0 <= i1 < H and 0 <= j1 < W:
                            raw[a, b] = np.float32(np.dot(f1[0, us[m], :, i0, j0].astype(np.float64),
                                                          f2[0, vs[m], :, i1, j1].astype(np.float64)))
                dx, dy = np.float32(x - np.floor(x)), np.float32(y - np.floor(y))
                for a in range(D - 1):
                    for b in range(D - 1):
                        v = (1 - dx) * (1 - dy) * raw[a, b] + dx * (1 - dy) * raw[a, b + 1] + \
                            (1 - dx) * dy * raw[a + 1, b] + dx * dy * raw[a + 1, b + 1]
                        assert abs(got[0, m, b, a, i0, j0] - v) < 1e-4 * max(1, abs(v))   # [x-off, y-off] order


def test_neighbors_oracle_vs_brute_force():
    rng = np.random.default_rng(2)
    kk = rng.integers(0, 6, 60); jj = rng.integers(0, 5, 60)
    ix, jx = fo.neighbors(kk, jj)
    for e in range(60):
        grp = [f for f in range(60) if kk[f] == kk[e]]
        grp.sort(key=lambda f: (jj[f], f))      # std::stable_sort by jj (ba.cpp:82-85)
        p = grp.index(e)
        assert ix[e] == (grp[p - 1] if p > 0 else -1)
        assert jx[e] == (grp[p + 1] if p + 1 < len(grp) else -1)


def test_ba_oracle_lowers_cost_and_respects_fixed_poses():
    synth = load_pkg("synth")
    d = synth.ba_inputs(12, seed=5, M=16)
    t0, t1 = 2, 12
    P1, Z1 = fo.ba(d["poses"], d["patches"], d["intrinsics"], d["target"], d["weight"], np.array([1e-4]),
                   d["ii"], d["jj"], d["kk"], d["M"], t0, t1, 2)

    def cost(po, pa):
        r, w, *_ = fo.edge_terms(np.asarray(po, np.float64), np.asarray(pa, np.float64), d["intrinsics"].astype(np.float64),
                                 d["target"].astype(np.float64), d["weight"].astype(np.float64), d["ii"], d["jj"], d["kk"])
        return float((w * r * r).sum())
    assert cost(P1, Z1) < 0.5 * cost(d["poses"], d["patches"])
    assert np.array_equal(P1[:t0], d["poses"][:t0].astype(np.float64))
    assert np.array_equal(Z1[:, :2], d["patches"][:, :2].astype(np.float64))


def test_synthetic_graph_matches_survey_sizes():
    synth = load_pkg("synth")
    ii, jj, kk = synth.dpvo_edges(40)
    assert len(kk) == 45312 and len(np.unique(kk)) == 22 * 96
    assert np.array_equal(ii, kk // 96)
    assert (np.abs(ii - jj) <= 13).all()


def test_torch_cpu_correlation_matches_numpy_restatement(synth):
    """oracle/altcorr_torch.py (the multi-threaded CPU baseline bench.py times) against the numpy
    restatement: fp32 vs float64 accumulation differ by at most a final-rounding flip."""
    from oracle import altcorr_oracle as oc, altcorr_torch as ot
    d = synth.corr_inputs(300, seed=11, mem=3, n_patches=32, oob_frac=0.1)
    want = oc.corr_pyramid(d["gmap"], [d["fmap1"], d["fmap2"]], [1.0, 0.25], d["coords"], d["us"], d["vs"], 3)
    got = ot.corr_pyramid(ot.prepare(d["gmap"], [d["fmap1"], d["fmap2"]]), [1.0, 0.25], d["coords"], d["us"], d["vs"], 3).numpy()
    assert got.shape == want.shape and got.dtype == want.dtype
    a, b = got.astype(np.float64), want.astype(np.float64)
    assert np.linalg.norm(a - b) / np.linalg.norm(b) < 1e-4
    assert (a == b).mean() > 0.99


# ------------------------------------------------------------------ projective_ops vs reference CPU
def test_projective_ops_restatement_matches_reference_cpu():
    from oracle import pops_oracle as po
    g = np.load(os.path.join(GOLD, "ref_cpu_pops.npz"))
    a = (g["poses"], g["patches"], g["intrinsics"], g["ii"], g["jj"], g["kk"])
    tol = 5e-6
    rel = lambda x, y: np.linalg.norm(np.asarray(x, np.float64) - y) / max(np.linalg.norm(y), 1e-30)
    assert rel(po.transform(*a), g["coords"]) < tol
    assert rel(po.transform(*a, depth=True), g["coords_depth"]) < tol
    c, v = po.transform(*a, valid=True)
    assert np.array_equal(v, g["valid"])
    assert rel(po.transform(*a, tonly=True), g["coords_tonly"]) < tol
    c, jv, (Ji, Jj, Jz) = po.transform(*a, jacobian=True)
    assert np.array_equal(jv, g["jvalid"])
    assert rel(Ji, g["Ji"]) < tol and rel(Jj, g["Jj"]) < tol and rel(Jz, g["Jz"]) < tol
    f, fv = po.flow_mag(*a, beta=0.5)
    assert rel(f, g["flow"]) < 2e-4          # differences of nearly equal float32 coordinates
    assert np.array_equal(fv, g["flow_valid"])
    assert rel(po.point_cloud(g["poses"], g["patches"], g["intrinsics"], g["ix"]), g["points"]) < tol
    # the older fixture (uniform intrinsics, jacobian=True) too
    t = np.load(os.path.join(GOLD, "ref_cpu_transform.npz"))
    c, jv, (Ji, Jj, Jz) = po.transform(t["poses"], t["patches"], t["intrinsics"], t["ii"], t["jj"], t["kk"], jacobian=True)
    assert rel(c, t["coords"]) < tol and rel(Ji, t["Ji"]) < tol and rel(Jj, t["Jj"]) < tol and rel(Jz, t["Jz"]) < tol


def test_pure_torch_ba_restatement_matches_reference_cpu():
    """oracle/ba_torch_oracle.py (dpvo/ba.py::BA, the CPU baseline bench.py times) against the reference's
    own ba.BA output (ref_cpu_ba.npz).  Both are float32: state to 1e-4, update to 5e-3."""
    from oracle import ba_torch_oracle as bt
    t = np.load(os.path.join(GOLD, "ref_cpu_transform.npz")); g = np.load(os.path.join(GOLD, "ref_cpu_ba.npz"))
    P, Z = bt.BA(t["poses"], t["patches"], t["intrinsics"], t["target"], t["weight"], 1e-4, t["ii"], t["jj"], t["kk"],
                 list(g["bounds"]), ep=float(g["ep"]), fixedp=int(g["fixedp"]))
    rel = lambda a, b: np.linalg.norm(a.astype(np.float64) - b) / np.linalg.norm(b)
    wantP = g["poses_out"].reshape(-1, 7); wantZ = g["patches_out"].reshape(-1, 3, 3, 3)
    assert rel(P, wantP) < 1e-4 and rel(Z[:, 2], wantZ[:, 2]) < 3e-4
    assert rel(P - t["poses"], wantP - t["poses"]) < 5e-3
    assert np.array_equal(P[:1], t["poses"][:1]) and np.array_equal(Z[:, :2], t["patches"][:, :2])


def test_update_oracle_vs_reference():
    """Update operator (N1): the functional restatement against the reference nn.Module's own float32
    output (tests/golden/make_golden_update.py).  Parameters come from the same seeded generator."""
    import importlib
    import torch
    from oracle import update_oracle as uo
    synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_cpu_update.npz"))
    sd = uo.state_dict_tensors(synth.update_state_dict(int(g["seed_w"])))
    d = synth.update_inputs(n_frames=int(g["n_frames"]), M=int(g["M"]), seed=int(g["seed_x"]))
    ii, jj, kk = (torch.from_numpy(d[k]) for k in ("ii", "jj", "kk"))
    imap = torch.from_numpy(d["imap"])
    net, delta, weight = uo.forward(sd, torch.from_numpy(d["net"]), imap[kk % imap.shape[0]],
                                    torch.from_numpy(d["corr"]), ii, jj, kk)
    assert net.shape == g["net"].shape
    np.testing.assert_allclose(net.numpy(), g["net"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(delta.numpy(), g["delta"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(weight.numpy(), g["weight"], rtol=1e-4, atol=1e-5)


def _pgo_cases():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_pgo", os.path.join(os.path.dirname(__file__), "golden",
                                                                                   "make_golden_pgo.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_pgo_oracle_vs_reference():
    """cuda_ba.solve_system: dense float64 restatement against the reference extension's own outputs."""
    from oracle import pgo_oracle
    m = _pgo_cases()
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_cpu_pgo.npz"))
    for name, (n, loops, freen, seed) in m.CASES.items():
        if n > 100:
            continue                                     # dense numpy J^T J of the 300-pose case: GPU test only
        Ji, Jj, ii, jj, res = m.pgo_case(n, loops, seed)
        got = pgo_oracle.solve_system(Ji, Jj, ii, jj, res, m.EP, m.LM, freen)
        np.testing.assert_allclose(got, g[name], rtol=2e-4, atol=2e-6, err_msg=name)


# ---------------------------------------------------------------- N4: patch-graph edge bookkeeping
class _OracleEdges:
    def __init__(self, m, seed):
        from oracle import edges_oracle as eo
        self.eo, self.m = eo, m
        N, M = m.CFG["BUFFER_SIZE"], m.CFG["PATCHES_PER_FRAME"]
        self.e = eo.Edges(np.repeat(np.arange(N), M), M, m.DIM)
        self.buf = m.initial_buffers(seed)
        self.r = m.CFG["PATCH_LIFETIME"]

    def forw(self, n): self.e.append_factors(*self.eo.edges_forw(n, self.e.M, self.r))
    def back(self, n): self.e.append_factors(*self.eo.edges_back(n, self.e.M, self.r))
    def count(self): return self.e.ii.size
    def set_rows(self, net, target, weight): self.e.net, self.e.target, self.e.weight = net, target, weight
    def remove_frame(self, k): self.e.remove_frame(k)
    def remove_old(self, *a): self.e.remove_old(*a)

    def shift(self, k, n):
        b, m = self.buf, self.m
        for key in ("tstamps_", "colors_", "poses_", "patches_", "intrinsics_"):
            self.eo.frames_shift(b[key], k, n)
        self.eo.frames_shift(b["imap_"], k, n, m.PMEM)
        self.eo.frames_shift(b["gmap_"], k, n, m.PMEM)
        self.eo.frames_shift(b["fmap1_"][0], k, n, m.MEM)
        self.eo.frames_shift(b["fmap2_"][0], k, n, m.MEM)

    def snapshot(self):
        return {k: getattr(self.e, k) for k in ("ii", "jj", "kk", "ii_inac", "jj_inac", "kk_inac")}

    def final(self):
        out = {k: getattr(self.e, k) for k in ("net", "target", "weight", "target_inac", "weight_inac")}
        out.update(self.buf)
        return out


@pytest.mark.parametrize("name,seed,lc", [("plain", 0, False), ("loop", 1, True)])
def test_edges_oracle_vs_reference(name, seed, lc):
    """oracle/edges_oracle.py against the reference's own append_factors / remove_factors / keyframe run on CPU."""
    import edges_replay
    m = edges_replay.golden_module()
    assert edges_replay.replay(_OracleEdges(m, seed), name, seed, lc) == 14


class _PlanCheckingEdges(_OracleEdges):
    """The replayed oracle, with `edges_oracle.keyframe_plan_counts` evaluated BEFORE every keyframe() edit and compared
    with what the (reference-pinned) edits then do: the plan's counts for the outcome that is taken must be exact."""

    def __init__(self, m, seed):
        super().__init__(m, seed)
        self.pending, self.checked = None, [0, 0]

    def _plan(self, k, n, lc, ow):
        c = self.m.CFG
        return self.eo.keyframe_plan_counts(self.e.ix, self.e.M, self.e.ii, self.e.jj, self.e.kk, k, n,
                                            c["REMOVAL_WINDOW"], lc, ow)

    def remove_frame(self, k):
        n = k + self.m.CFG["KEYFRAME_INDEX"]
        # the loop-closure flag is not known here: both plans are kept, remove_old picks the one that applies
        self.pending = (k, n, self.e.ii.size)
        self.before = (self.e.ii.copy(), self.e.jj.copy(), self.e.kk.copy())
        super().remove_frame(k)

    def remove_old(self, n, window, lc, ow):
        if self.pending is not None:
            k, n0, size0 = self.pending
            assert n0 - 1 == n
            ii, jj, kk = self.before
            _, _, dropped, moved, mn = self.eo.keyframe_plan_counts(self.e.ix, self.e.M, ii, jj, kk, k, n0, window, lc, ow)
            assert size0 - self.e.ii.size == dropped
            self.pending = None
            self.checked[1] += 1
        else:
            moved, mn, _, _, _ = self._plan(n - self.m.CFG["KEYFRAME_INDEX"], n, lc, ow)
            self.checked[0] += 1
        size, inac = self.e.ii.size, self.e.ii_inac.size
        super().remove_old(n, window, lc, ow)
        assert size - self.e.ii.size == moved and self.e.ii_inac.size - inac == moved
        assert (int(self.e.ii.min()) if self.e.ii.size else None) == mn


@pytest.mark.parametrize("name,seed,lc", [("plain", 0, False), ("loop", 1, True)])
def test_keyframe_plan_oracle_vs_reference_sequence(name, seed, lc):
    """`edges_oracle.keyframe_plan_counts` (the expected output of the product's one-read keyframe plan) on the scripted
    sequence whose every state is the reference's own (tests/golden/ref_cpu_edges.npz): both outcomes occur."""
    import edges_replay
    m = edges_replay.golden_module()
    impl = _PlanCheckingEdges(m, seed)
    edges_replay.replay(impl, name, seed, lc)
    assert impl.checked[0] > 0 and impl.checked[1] > 0, impl.checked


def test_reduce_edges_oracle_vs_reference():
    import edges_replay
    from oracle import edges_oracle as eo
    m, g = edges_replay.golden_module(), edges_replay.golden()
    for i, c in enumerate(m.REDUCE_CASES):
        mag, ii, jj = m.reduce_case(c)
        got = eo.reduce_edges(mag, ii, jj, c["max_num_edges"], c["nms"])
        assert np.array_equal(got, g[f"reduce{i}"]), i


def test_viewer_oracle_pose_to_matrix_vs_lietorch_oracle():
    """N5: oracle/viewer_oracle.pose_to_matrix (unpinned restatement) agrees with the pinned lietorch oracle:
    the 4x4 matrix of the inverse pose."""
    from oracle import viewer_oracle as vo, lietorch_oracle as lo
    rng = np.random.default_rng(11)
    q = rng.standard_normal((64, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    p = np.concatenate([rng.standard_normal((64, 3)), q], 1).astype(np.float32)
    m = vo.pose_to_matrix(p)
    assert m.shape == (64, 4, 4) and m.dtype == np.float32
    assert np.abs(m - lo.matrix(lo.inv(p.astype(np.float64)))).max() < 1e-5
    # identical frames under identity poses confirm every inner pixel once per other frame
    N, ht, wd = 3, 6, 7
    poses = np.zeros((N, 7), np.float32); poses[:, 6] = 1
    pts, col, cnt, _, inner = vo.iproj(0, N, 5.0, np.zeros((N, 3, ht, wd), np.uint8), poses,
                                       np.full((N, ht, wd), 0.3, np.float32), np.tile(np.float32([[1, 1, 0.4, 0.3]]), (N, 1)))
    assert np.all(cnt[inner] == N - 1) and np.all(cnt[~inner] == 0)


def _encoder_golden():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_encoder", os.path.join(os.path.dirname(__file__), "golden",
                                                                                       "make_golden_encoder.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m, np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_cpu_encoder.npz"))


@pytest.mark.parametrize("name", ["fnet", "inet"])
def test_encoder_oracle_vs_reference(name):
    """N3: oracle/encoder_oracle.py against the reference BasicEncoder4 module's own float32 outputs."""
    from oracle import encoder_oracle as eo
    from conftest import load_pkg
    synth = load_pkg("synth")
    m, g = _encoder_golden()
    c = m.CASES[name]
    got = eo.basic_encoder4(synth.encoder_state_dict(c["out_dim"], c["seed"]),
                            synth.encoder_images(m.N_IMG, m.H, m.W, m.IMG_SEED), c["norm_fn"])[0]
    assert got.shape == g[name].shape
    assert np.abs(got - g[name]).max() < 2e-5 * max(1.0, np.abs(g[name]).max())


def test_sparse_ba_oracle_equals_dense_ba_oracle():
    """oracle/fastba_oracle.ba_sparse (scipy.sparse float64, used at global-BA sizes) against the dense float64
    restatement `ba()` (itself pinned to the reference extension on the GPU box) on a window both can hold."""
    import importlib
    from oracle import fastba_oracle as fo
    synth = importlib.import_module("deep-visual-odometry-slam_b200.synth") if False else None
    import importlib.util, os
    spec = importlib.util.spec_from_file_location("_synth", os.path.join(os.path.dirname(os.path.dirname(__file__)),
                                                  "deep-visual-odometry-slam_b200", "synth.py"))
    synth = importlib.util.module_from_spec(spec); spec.loader.exec_module(synth)
    d = synth.ba_inputs(18, seed=5, M=8, n_buf=24)
    args = (d["poses"], d["patches"], d["intrinsics"], d["target"], d["weight"], 1e-4, d["ii"], d["jj"], d["kk"], 8, 3, 18, 2)
    p0, z0, u0 = fo.ba(*args, return_updates=True)
    p1, z1, u1 = fo.ba_sparse(*args)
    assert np.abs(p0 - p1).max() < 1e-9 and np.abs(z0 - z1).max() < 1e-9
    for (dx0, dz0, k0), (dx1, dz1, k1) in zip(u0, u1):
        assert np.array_equal(k0, k1)
        assert np.abs(dx0 - dx1).max() < 1e-9 * max(1.0, np.abs(dx0).max())
        assert np.abs(dz0 - dz1).max() < 1e-9 * max(1.0, np.abs(dz0).max())
