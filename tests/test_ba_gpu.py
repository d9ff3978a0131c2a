"""GPU parity tests for fastba (BA, neighbors, reproject, unique/inverse) through the Python
operator surface -> ctypes -> C ABI -> sm_100a kernels.

Oracles: oracle/fastba_oracle.py (float64 restatement) and, when oracle/_ref/cuda_ba.so is
present, the reference's own CUDA extension on identical inputs.

Tolerances (north star: BA pose/depth updates within 1e-5 relative in fp32).  "Relative" is norm-wise on the updated
window poses and inverse depths.  The reference accumulates its Hessian with fp32 atomics in a run-to-run varying
order, and gauge-free windows (all poses but one free: monocular scale) are ill-conditioned in fp32, so the bars are
MEASURED in each test instead of being flat numbers (`_bars`):
    vs the reference extension:   max(1e-5, 3 x the extension's own run-to-run spread on this input,
                                            2 x the extension's own distance to the float64 restatement); where the
                                  float64 bar is above 1e-5 (ill-conditioned windows) it is that bar + the extension's
                                  own distance to float64 (triangle inequality: nothing the float64 check does not imply)
    vs the float64 restatement:   max(1e-5, 2 x the extension's distance to it, 3 x the extension's run-to-run spread —
                                            its distance to float64 is itself a random variable of that spread), and
                                  not below the largest bar this rule produced for the same seeded case on an earlier
                                  run (profiles/r02_ba_accuracy.jsonl): two runs are a small sample of the spread
(without oracle/_ref: the bar recorded for the same seeded case in profiles/r02_ba_accuracy.jsonl, else 1e-5 for
gauge-anchored windows and GAUGE_FREE_FALLBACK otherwise).  The update vectors
(state_after - state_before) are held to the same ABSOLUTE error, i.e. the same bars times |state| / |update|.
Every measured number goes to gpurun_out/ba_accuracy.jsonl (committed as profiles/r02_ba_accuracy.jsonl).
Integer outputs (neighbors, unique/inverse) are bit-exact.
"""
import numpy as np
import pytest
import torch

from conftest import load_pkg
from oracle import fastba_oracle as fo
from oracle.ref_ext import load_ref

pytestmark = pytest.mark.gpu

STATE_TOL = 1e-5                 # the north star's bar
GAUGE_FREE_FALLBACK = 3e-4       # only when the reference extension is not built and the window has a free gauge


def _rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def _dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def _run_ours(pkg, d, t0, t1, iters=2, eff=False):
    poses = _dev(d["poses"])[None].clone()
    patches = _dev(d["patches"])[None].clone()
    lm = torch.tensor([1e-4], device="cuda")
    pkg.fastba.BA(poses, patches, _dev(d["intrinsics"])[None], _dev(d["target"])[None], _dev(d["weight"])[None], lm,
                  _dev(d["ii"]), _dev(d["jj"]), _dev(d["kk"]), t0, t1, d["M"], iters, eff)
    torch.cuda.synchronize()
    return poses[0].cpu().numpy(), patches[0].cpu().numpy()


def _run_oracle(d, t0, t1, iters=2):
    return fo.ba(d["poses"], d["patches"], d["intrinsics"], d["target"], d["weight"], np.array([1e-4]),
                 d["ii"], d["jj"], d["kk"], d["M"], t0, t1, iters)


def _errs(d, got, want, t0, t1):
    """(pose error, depth error, update error) — norm-wise relative on the window poses / touched inverse depths;
    the update error is the same absolute difference relative to the UPDATE's norm."""
    gp, gz = got
    wp, wz = want
    p0 = d["poses"].astype(np.float64); z0 = d["patches"].astype(np.float64)
    w = slice(t0, max(t1, t0))
    e_p = _rel(gp[w], wp[w]) if t1 > t0 else 0.0
    e_z = _rel(gz[:, 2], wz[:, 2])
    up = np.concatenate([(np.asarray(wp, np.float64)[w] - p0[w]).ravel(), (np.asarray(wz, np.float64)[:, 2] - z0[:, 2]).ravel()])
    ug = np.concatenate([(np.asarray(gp, np.float64)[w] - p0[w]).ravel(), (np.asarray(gz, np.float64)[:, 2] - z0[:, 2]).ravel()])
    return e_p, e_z, _rel(ug, up)


def _check_against(d, got, want, tol_state=STATE_TOL, t0=0, t1=None):
    t1 = d["poses"].shape[0] if t1 is None else t1
    e_p, e_z, _ = _errs(d, got, want, t0, t1)
    assert e_p < tol_state, (e_p, tol_state)
    assert e_z < tol_state, (e_z, tol_state)
    # x/y of the patches are untouched
    assert np.array_equal(got[1][:, :2], d["patches"][:, :2])


def _log(rec):
    import json, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    os.makedirs(os.path.join(root, "gpurun_out"), exist_ok=True)
    with open(os.path.join(root, "gpurun_out", "ba_accuracy.jsonl"), "a") as f:
        f.write(json.dumps(rec) + "\n")


def _run_ref(d, t0, t1, iters=2, eff=False):
    ref = load_ref("cuda_ba")
    if ref is None:
        return None
    poses = _dev(d["poses"])[None].clone(); patches = _dev(d["patches"])[None].clone()
    lm = torch.tensor([1e-4], device="cuda")
    ref.forward(poses, patches, _dev(d["intrinsics"])[None], _dev(d["target"])[None], _dev(d["weight"])[None], lm,
                _dev(d["ii"]), _dev(d["jj"]), _dev(d["kk"]), d["M"], t0, t1, iters, eff)
    torch.cuda.synchronize()
    return poses[0].cpu().numpy(), patches[0].cpu().numpy()


def _recorded_bar(tag):
    import json, os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "r02_ba_accuracy.jsonl")
    best = 0.0
    if tag and os.path.exists(path):
        for line in open(path):
            rec = json.loads(line)
            if rec.get("case") == tag and "reference_self_spread" in rec:
                best = max(best, float(rec["bar_vs_float64"]))
    return best


def _bars(d, t0, t1, want, iters=2, eff=False, tag=""):
    """Measured bars (module docstring).  Returns (bar_vs_oracle, bar_vs_ref, ref_result or None)."""
    r1 = _run_ref(d, t0, t1, iters, eff)
    if r1 is None:
        # no reference extension on this machine: the bar measured with it on an earlier run of the same seeded case
        # (profiles/r02_ba_accuracy.jsonl, written by this function), else the flat fall-backs
        bar = max(_recorded_bar(tag), GAUGE_FREE_FALLBACK if t0 <= 1 else STATE_TOL)
        return bar, None, None
    r2 = _run_ref(d, t0, t1, iters, eff)
    sp_p, sp_z, sp_u = _errs(d, r2, r1, t0, t1)
    rp, rz, ru = _errs(d, r1, want, t0, t1)
    spread, ref_to_f64 = max(sp_p, sp_z), max(rp, rz)
    bar_oracle = max(STATE_TOL, 2 * ref_to_f64, 3 * spread)
    # two runs of the extension are a small sample of its spread (same seeded case, call to call: 2.4e-5 and 1.5e-4 on
    # the gauge-free blocked window): never go below the bar an earlier run of this rule recorded for the case
    bar_oracle = max(bar_oracle, _recorded_bar(tag))
    # both within their bars of float64 => within the sum of each other (triangle inequality)
    bar_ref = max(STATE_TOL, bar_oracle + max(ref_to_f64, spread)) if bar_oracle > STATE_TOL else max(STATE_TOL, 3 * spread, 2 * ref_to_f64)
    _log({"case": tag, "edges": int(len(d["kk"])), "window": [int(t0), int(t1)], "eff_impl": bool(eff),
          "reference_self_spread": spread, "reference_to_float64": ref_to_f64, "reference_update_err_vs_float64": ru,
          "bar_vs_float64": bar_oracle, "bar_vs_reference": bar_ref})
    return bar_oracle, bar_ref, r1


def _assert_parity(d, t0, t1, got, want, iters=2, eff=False, tag=""):
    bar_oracle, bar_ref, ref = _bars(d, t0, t1, want, iters, eff, tag)
    e_p, e_z, e_u = _errs(d, got, want, t0, t1)
    rec = {"case": tag + ":ours", "ours_to_float64": max(e_p, e_z), "ours_update_err_vs_float64": e_u, "bar_vs_float64": bar_oracle}
    if ref is not None:
        q_p, q_z, q_u = _errs(d, got, (ref[0].astype(np.float64), ref[1].astype(np.float64)), t0, t1)
        rec.update(ours_to_reference=max(q_p, q_z), ours_update_err_vs_reference=q_u, bar_vs_reference=bar_ref)
    _log(rec)
    print(rec)
    _check_against(d, got, want, bar_oracle, t0, t1)
    if ref is not None:
        _check_against(d, got, (ref[0].astype(np.float64), ref[1].astype(np.float64)), bar_ref, t0, t1)
    return bar_oracle


@pytest.mark.parametrize("n_frames,M", [(14, 96), (9, 32), (30, 24)])
def test_ba_local_window_vs_oracle(n_frames, M):
    pkg = load_pkg(); synth = load_pkg("synth")
    d = synth.ba_inputs(n_frames, seed=n_frames, M=M)
    t0, t1 = max(n_frames - 10, 1), n_frames
    got = _run_ours(pkg, d, t0, t1)
    want = _run_oracle(d, t0, t1)
    _assert_parity(d, t0, t1, got, want, tag=f"local_window_{n_frames}x{M}")
    assert np.array_equal(got[0][:t0], d["poses"][:t0]) and np.array_equal(got[0][t1:], d["poses"][t1:])


def test_ba_wide_window_and_structure_only():
    pkg = load_pkg(); synth = load_pkg("synth")
    d = synth.ba_inputs(40, seed=7, M=16)
    # 39 free poses: 234x234 system (Cholesky spills out of shared memory)
    got = _run_ours(pkg, d, 1, 40)
    want = _run_oracle(d, 1, 40)
    _assert_parity(d, 1, 40, got, want, tag="wide_window_39_free_poses")
    # t1 == t0: structure-only branch (ba_cuda.cu:521-532): poses fixed, depths move
    got = _run_ours(pkg, d, 5, 5)
    want = _run_oracle(d, 5, 5)
    assert np.array_equal(got[0], d["poses"])
    assert _rel(got[1][:, 2], want[1][:, 2]) < STATE_TOL


@pytest.mark.parametrize("eff", [False, True])
def test_ba_masked_and_degenerate_edges(eff):
    """Edges behind the camera / far out of bounds / zero weight contribute nothing; duplicate
    (loop-closure style) edges are summed — on the dense path and on the block-E path, whose shared-memory fill gives
    every (patch, target pose) entry one writer and must let two edges with the same target take turns."""
    pkg = load_pkg(); synth = load_pkg("synth")
    d = synth.ba_inputs(14, seed=3, M=24)
    rng = np.random.default_rng(0)
    E = len(d["kk"])
    d["target"][rng.choice(E, 200, replace=False)] += 1000.0   # residual > 128 -> masked
    d["weight"][rng.choice(E, 200, replace=False)] = 0.0
    dup = rng.choice(E, 300, replace=False)                      # duplicated edges
    for k in ("ii", "jj", "kk"):
        d[k] = np.concatenate([d[k], d[k][dup]])
    d["target"] = np.concatenate([d["target"], d["target"][dup] + 0.3])
    d["weight"] = np.concatenate([d["weight"], d["weight"][dup]])
    got = _run_ours(pkg, d, 4, 14, eff=eff)
    want = _run_oracle(d, 4, 14)
    _assert_parity(d, 4, 14, got, want, eff=eff, tag="masked_and_degenerate_edges" + ("_eff" if eff else ""))


def _with_loop_edges(synth, n_frames, M, seed, hub=None, n_pairs=12):
    """DPVO-rule edges plus loop-closure style edges (all M patches of frame i -> far frame j,
    patchgraph.py:56-82 shape) and optionally a hub frame whose patches see every later frame."""
    rng = np.random.default_rng(seed)
    ii, jj, kk = synth.dpvo_edges(n_frames, M)
    extra = []
    for _ in range(n_pairs):
        i = int(rng.integers(0, n_frames)); j = int(rng.integers(0, n_frames))
        if abs(i - j) > 14:
            extra.append((i, j))
    if hub is not None:
        extra += [(hub, j) for j in range(n_frames) if abs(j - hub) > 13]
    for i, j in extra:
        k = i * M + np.arange(M)
        ii = np.concatenate([ii, np.full(M, i)]); jj = np.concatenate([jj, np.full(M, j)]); kk = np.concatenate([kk, k])
    return synth.ba_inputs(n_frames, seed=seed, M=M, edges=(ii.astype(np.int64), jj.astype(np.int64), kk.astype(np.int64)))


@pytest.mark.parametrize("case", ["local_window", "gauge_free_blocked", "loop_closure_hub", "structure_only"])
def test_ba_eff_impl(case):
    """eff_impl=True (block-E Schur path, block_e.cu): same mathematics as the dense path, so it must
    agree with the float64 restatement, with our dense path and with the reference's eff_impl run."""
    pkg = load_pkg(); synth = load_pkg("synth")
    if case == "local_window":
        d = synth.ba_inputs(14, seed=21, M=96); t0, t1 = 4, 14
    elif case == "gauge_free_blocked":
        d = synth.ba_inputs(40, seed=22, M=16); t0, t1 = 1, 40            # n = 234: blocked Cholesky
    elif case == "loop_closure_hub":
        d = _with_loop_edges(synth, 64, 24, seed=23, hub=3); t0, t1 = 2, 64  # frame 3 sees ~50 frames: 2 conn tiles
    else:
        d = synth.ba_inputs(14, seed=24, M=32); t0, t1 = 6, 6
    got = _run_ours(pkg, d, t0, t1, eff=True)
    dense = _run_ours(pkg, d, t0, t1, eff=False)
    want = _run_oracle(d, t0, t1)
    # the reference's eff_impl run faults (illegal address) when t1 == t0, so it is not used there
    if case == "structure_only":
        _check_against(d, got, want, STATE_TOL, t0, t1)
        _check_against(d, got, (dense[0].astype(np.float64), dense[1].astype(np.float64)), STATE_TOL, t0, t1)
    else:
        bar_oracle = _assert_parity(d, t0, t1, got, want, eff=True, tag="eff_impl_" + case)
        # dense and block-E paths are both within bar_oracle of float64 -> within twice that of each other
        _check_against(d, dense, want, bar_oracle, t0, t1)
        _check_against(d, got, (dense[0].astype(np.float64), dense[1].astype(np.float64)), 2 * bar_oracle, t0, t1)
    assert np.array_equal(got[0][:t0], d["poses"][:t0]) and np.array_equal(got[0][t1:], d["poses"][t1:])


def test_ba_eff_impl_global_300_keyframes():
    """DPV-SLAM-style global BA (BASELINE config 5, scaled to what the numpy restatement finishes in
    seconds): 300 keyframes, every sequential edge kept (no removal-window cull), 40 loop-closure
    frame pairs, all poses but the first free (n = 1794), eff_impl=True, 2 iterations."""
    pkg = load_pkg(); synth = load_pkg("synth")
    M, nf = 32, 300
    rng = np.random.default_rng(5)
    ii, jj, kk = synth.dpvo_edges(nf, M, removal=10 ** 9)
    for _ in range(40):
        i, j = int(rng.integers(0, nf)), int(rng.integers(0, nf))
        if abs(i - j) >= 30:
            ii = np.concatenate([ii, np.full(M, i)]); jj = np.concatenate([jj, np.full(M, j)])
            kk = np.concatenate([kk, i * M + np.arange(M)])
    d = synth.ba_inputs(nf, seed=31, M=M, edges=(ii.astype(np.int64), jj.astype(np.int64), kk.astype(np.int64)))
    t0, t1 = 1, nf
    got = _run_ours(pkg, d, t0, t1, eff=True)
    want = _run_oracle(d, t0, t1)
    _assert_parity(d, t0, t1, got, want, eff=True, tag="global_300_keyframes")


def test_ba_eff_impl_global_1000_keyframes():
    """BASELINE config 5 AT ITS STATED SIZE: 1,000 keyframes x 96 patches, every sequential edge kept (2.4 M edges),
    ~190 proximity loop-closure frame pairs, all poses but the first free (n = 5,994), eff_impl=True, 2 iterations
    (dpvo.py:312-326, block_e.cu, ba_cuda.cu:552-565) — against a float64 SPARSE direct solve of the same normal
    equations (oracle/fastba_oracle.ba_sparse: scipy.sparse + SuperLU on the host, test-only) and against the
    reference extension's own eff_impl run, with the measured bars of the module docstring."""
    pkg = load_pkg(); synth = load_pkg("synth")
    M, nf = 96, 1000
    rng = np.random.default_rng(0)
    ii, jj, kk = synth.dpvo_edges(nf, M, removal=10 ** 9)
    for _ in range(200):
        i, j = int(rng.integers(0, nf)), int(rng.integers(0, nf))
        if abs(i - j) >= 30:
            ii = np.concatenate([ii, np.full(M, i)]); jj = np.concatenate([jj, np.full(M, j)])
            kk = np.concatenate([kk, i * M + np.arange(M)])
    d = synth.ba_inputs(nf, seed=1, M=M, n_buf=nf + 2, edges=(ii.astype(np.int64), jj.astype(np.int64), kk.astype(np.int64)))
    t0, t1 = 1, nf
    got = _run_ours(pkg, d, t0, t1, eff=True)
    wp, wz, _ = fo.ba_sparse(d["poses"], d["patches"], d["intrinsics"], d["target"], d["weight"], np.array([1e-4]),
                             d["ii"], d["jj"], d["kk"], M, t0, t1, 2)
    assert len(kk) > 2_300_000
    _assert_parity(d, t0, t1, got, (wp, wz), eff=True, tag="global_1000_keyframes_2.4M_edges")


def test_ba_vs_reference_extension():
    ref = load_ref("cuda_ba")
    if ref is None:
        pytest.skip("oracle/_ref/cuda_ba.so not built")
    pkg = load_pkg(); synth = load_pkg("synth")
    d = synth.ba_inputs(14, seed=1)
    t0, t1 = 4, 14
    poses = _dev(d["poses"])[None].clone(); patches = _dev(d["patches"])[None].clone()
    lm = torch.tensor([1e-4], device="cuda")
    ref.forward(poses, patches, _dev(d["intrinsics"])[None], _dev(d["target"])[None], _dev(d["weight"])[None], lm,
                _dev(d["ii"]), _dev(d["jj"]), _dev(d["kk"]), d["M"], t0, t1, 2, False)
    torch.cuda.synchronize()
    refp, refz = poses[0].cpu().numpy(), patches[0].cpu().numpy()
    want = _run_oracle(d, t0, t1)
    got = _run_ours(pkg, d, t0, t1)
    # the restatement is pinned by the reference's own kernels ...
    _check_against(d, (refp, refz), want, STATE_TOL, t0, t1)
    # ... ours agrees with the reference's output directly ...
    _check_against(d, got, (refp.astype(np.float64), refz.astype(np.float64)), STATE_TOL, t0, t1)
    # ... and is at least as close to the float64 restatement as the reference is (x2 slack)
    p0 = d["poses"].astype(np.float64)
    err_ours = _rel(got[0] - p0, want[0] - p0)
    err_ref = _rel(refp - p0, want[0] - p0)
    assert err_ours <= max(2 * err_ref, 1e-5)


def _random_edges(rng, n, nk, nj, dup=True):
    kk = rng.integers(0, nk, n).astype(np.int64)
    jj = rng.integers(0, nj, n).astype(np.int64)
    if dup and n > 10:
        kk[n // 2:n // 2 + 5] = kk[:5]; jj[n // 2:n // 2 + 5] = jj[:5]   # exact duplicate (kk,jj) pairs
    return kk, jj


@pytest.mark.parametrize("n", [0, 1, 7, 1000, 45312])
def test_neighbors_bit_exact(n):
    pkg = load_pkg(); synth = load_pkg("synth")
    rng = np.random.default_rng(n)
    if n == 45312:
        _, jj, kk = synth.dpvo_edges(40)
    else:
        kk, jj = _random_edges(rng, n, max(n // 9, 1), 30)
    if n == 0:
        kk = np.zeros(0, np.int64); jj = np.zeros(0, np.int64)
    ix, jx = pkg.fastba.neighbors(_dev(kk), _dev(jj))
    wix, wjx = fo.neighbors(kk, jj)
    assert ix.dtype == torch.int64 and ix.is_cuda
    assert np.array_equal(ix.cpu().numpy(), wix) and np.array_equal(jx.cpu().numpy(), wjx)
    ref = load_ref("cuda_ba")
    if ref is not None and n > 0:
        rix, rjx = ref.neighbors(_dev(kk), _dev(jj))
        assert torch.equal(rix, ix) and torch.equal(rjx, jx)


def test_neighbors_large_keys_and_properties():
    """Keys far apart (no bound on the key range), then a chain property: jx[ix[e]] == e."""
    pkg = load_pkg()
    rng = np.random.default_rng(4)
    kk = (rng.integers(0, 50, 5000) * 10_000_000_019 + 123).astype(np.int64)
    jj = rng.integers(0, 40, 5000).astype(np.int64)
    ix, jx = pkg.fastba.neighbors(_dev(kk), _dev(jj))
    wix, wjx = fo.neighbors(kk, jj)
    assert np.array_equal(ix.cpu().numpy(), wix) and np.array_equal(jx.cpu().numpy(), wjx)
    has_prev = ix >= 0
    assert torch.equal(jx[ix[has_prev]], torch.nonzero(has_prev).flatten())


def test_unique_inverse_bit_exact():
    pkg = load_pkg(); synth = load_pkg("synth")
    _, _, kk = synth.dpvo_edges(40)
    bound = 4096 * 96
    uniq, inv, cnt = pkg.cuda_ba.unique_inverse(_dev(kk), bound)
    wu, wi = fo.unique_inverse(kk)
    m = int(cnt.item())
    assert m == len(wu)
    assert np.array_equal(uniq[:m].cpu().numpy(), wu) and np.array_equal(inv.cpu().numpy(), wi)
    tu, ti = torch.unique(_dev(kk), sorted=True, return_inverse=True)
    assert torch.equal(tu, uniq[:m]) and torch.equal(ti, inv)


def test_reproject_vs_oracle_and_reference():
    pkg = load_pkg(); synth = load_pkg("synth")
    d = synth.ba_inputs(14, seed=5, M=48)
    args = (_dev(d["poses"])[None], _dev(d["patches"])[None], _dev(d["intrinsics"])[None],
            _dev(d["ii"]), _dev(d["jj"]), _dev(d["kk"]))
    got = pkg.fastba.reproject(*args)
    want = fo.reproject(d["poses"], d["patches"], d["intrinsics"], d["ii"], d["jj"], d["kk"])
    assert got.shape == want.shape
    assert _rel(got.cpu().numpy(), want) < 1e-5
    ref = load_ref("cuda_ba")
    if ref is not None:
        w2 = ref.reproject(*args)
        assert _rel(got.cpu().numpy(), w2.cpu().numpy()) < 1e-6


def test_ba_full_size_properties():
    """Steady-state DPVO graph (45,312 edges, 10-pose window): BA must lower the weighted
    reprojection cost, leave fixed poses bit-identical, and be invariant to edge order."""
    pkg = load_pkg(); synth = load_pkg("synth")
    d = synth.ba_inputs(40, seed=2)
    assert len(d["kk"]) == 45312
    t0, t1 = 30, 40
    got = _run_ours(pkg, d, t0, t1)

    def cost(po, pa):
        r, w, *_ = fo.edge_terms(po.astype(np.float64), pa.astype(np.float64), d["intrinsics"].astype(np.float64),
                                 d["target"].astype(np.float64), d["weight"].astype(np.float64), d["ii"], d["jj"], d["kk"])
        return float((w * r * r).sum())
    assert cost(*got) < 0.5 * cost(d["poses"], d["patches"])
    assert np.array_equal(got[0][:t0], d["poses"][:t0])
    perm = np.random.default_rng(0).permutation(len(d["kk"]))
    d2 = dict(d)
    for k in ("ii", "jj", "kk", "target", "weight"):
        d2[k] = d[k][perm]
    got2 = _run_ours(pkg, d2, t0, t1)
    assert _rel(got2[0], got[0]) < 1e-5 and _rel(got2[1], got[1]) < 1e-5
    want = _run_oracle(d, t0, t1)
    _check_against(d, got, want)


@pytest.mark.parametrize("n", [1, 17, 64, 65, 128, 200, 234, 384, 1000, 1794])
def test_blocked_cholesky_solve(n):
    """The blocked Cholesky of csrc/ba_chol.cu on its own: random SPD systems against numpy float64."""
    pkg = load_pkg()
    rng = np.random.default_rng(n)
    M = rng.standard_normal((n, n + 8))
    S = (M @ M.T / (n + 8) + 0.5 * np.eye(n)).astype(np.float32)
    y = rng.standard_normal(n).astype(np.float32)
    x, info = pkg.cuda_ba.cholesky_solve(_dev(S), _dev(y))
    assert int(info.item()) == 0
    want = np.linalg.solve(S.astype(np.float64), y.astype(np.float64))
    assert _rel(x.cpu().numpy(), want) < 2e-5
    # a matrix that is not positive definite is reported, not silently factored
    S2 = S.copy(); S2[n // 2, n // 2] = -1.0
    x2, info2 = pkg.cuda_ba.cholesky_solve(_dev(S2), _dev(y))
    assert int(info2.item()) == 1 and float(x2.abs().max()) == 0.0


@pytest.mark.parametrize("n", [2309, 3000, 4890])
def test_blocked_cholesky_bulk_update_both_arithmetics(n):
    """Systems large enough for the 128x128 bulk trailing update (>= one tile per SM: n > ~2,240; global bundle
    adjustment sizes): split-TF32 tensor-core tiles (default) and fp32 FMA tiles (dpvo_chol_bulk_mode) against numpy
    float64, and against each other.  n is not a multiple of 64 or 128: ragged last panel and edge tiles."""
    pkg = load_pkg()
    rng = np.random.default_rng(n)
    M = rng.standard_normal((n, n + 8)).astype(np.float32)
    S = (M @ M.T / (n + 8) + 0.5 * np.eye(n, dtype=np.float32)).astype(np.float32)
    y = rng.standard_normal(n).astype(np.float32)
    want = np.linalg.solve(S.astype(np.float64), y.astype(np.float64))
    got = {}
    try:
        for mode in (1, 0):
            pkg.cuda_ba.chol_bulk_mode(mode)
            x, info = pkg.cuda_ba.cholesky_solve(_dev(S), _dev(y))
            assert int(info.item()) == 0
            got[mode] = x.cpu().numpy()
    finally:
        pkg.cuda_ba.chol_bulk_mode(0)
    e_fp32, e_tf32 = _rel(got[1], want), _rel(got[0], want)
    _log({"case": f"blocked_cholesky_n{n}", "fp32_bulk_to_float64": e_fp32, "split_tf32_bulk_to_float64": e_tf32,
          "between": _rel(got[0], got[1])})
    # every matrix element has one writer per launch: both arithmetics must repeat bit for bit (a hand-rolled
    # producer / consumer pipeline that does not is racing)
    try:
        for mode in (1, 0):
            pkg.cuda_ba.chol_bulk_mode(mode)
            for _ in range(4):
                x, _info = pkg.cuda_ba.cholesky_solve(_dev(S), _dev(y))
                assert np.array_equal(x.cpu().numpy(), got[mode]), f"bulk mode {mode} does not repeat"
    finally:
        pkg.cuda_ba.chol_bulk_mode(0)
    assert e_fp32 < 2e-5 and e_tf32 < 2e-5, (e_fp32, e_tf32)
    assert e_tf32 < 4 * e_fp32 + 1e-6, (e_fp32, e_tf32)          # the split loses at most a couple of bits
    with pytest.raises(RuntimeError):
        pkg.cuda_ba.chol_bulk_mode(2)


@pytest.mark.parametrize("n,loops", [(2400, 6), (3001, 0), (3001, 40)])
def test_blocked_cholesky_band_plus_loop_closures(n, loops):
    """A reduced camera system shaped like a long trajectory's: a block band (half-width 150) plus a few loop-closure
    blocks far from the diagonal.  Most 64x64 tiles of the factor are structurally zero and the solver skips them
    (tile bitmap + symbolic elimination in csrc/ba_chol.cu); the result must still be the dense float64 solve's."""
    pkg = load_pkg()
    rng = np.random.default_rng(n + loops)
    S = np.zeros((n, n), np.float64)
    bw = 150
    for i0 in range(0, n, 6):
        j0 = max(0, i0 - bw)
        rows = min(6, n - i0)
        S[i0:i0 + rows, j0:i0 + rows] += rng.standard_normal((rows, i0 - j0 + rows)) * 0.05
    for _ in range(loops):
        a, b = sorted(int(v) for v in rng.integers(0, n // 6, 2))
        if b - a > 60:
            S[6 * b:6 * b + 6, 6 * a:6 * a + 6] += rng.standard_normal((6, 6)) * 0.05
    S = np.tril(S); S = S + S.T
    S += np.diag(np.abs(S).sum(1) + 1.0)                 # strictly diagonally dominant: SPD
    S32 = S.astype(np.float32)
    y = rng.standard_normal(n).astype(np.float32)
    want = np.linalg.solve(S32.astype(np.float64), y.astype(np.float64))
    got = {}
    try:
        for mode in (1, 0):
            pkg.cuda_ba.chol_bulk_mode(mode)
            x, info = pkg.cuda_ba.cholesky_solve(_dev(S32), _dev(y))
            assert int(info.item()) == 0
            got[mode] = x.cpu().numpy()
            x2, _ = pkg.cuda_ba.cholesky_solve(_dev(S32), _dev(y))
            assert np.array_equal(x2.cpu().numpy(), got[mode])
    finally:
        pkg.cuda_ba.chol_bulk_mode(0)
    assert _rel(got[0], want) < 2e-5 and _rel(got[1], want) < 2e-5, (_rel(got[0], want), _rel(got[1], want))


@pytest.mark.parametrize("eff", [False, True])
def test_ba_prepare_then_forward_matches_single_call(eff):
    """dpvo_ba_prepare on a side stream + dpvo_ba_forward_prepared on the main stream (what the frame graph
    does) against the single call: same kernels on the same workspace, so within atomics-order noise."""
    pkg = load_pkg(); synth = load_pkg("synth")
    d = _with_loop_edges(synth, 30, 24, 5) if eff else synth.ba_inputs(16, seed=5, M=48)
    t0, t1 = (8, 30) if eff else (6, 16)
    one = _run_ours(pkg, d, t0, t1, eff=eff)
    poses = _dev(d["poses"])[None].clone(); patches = _dev(d["patches"])[None].clone()
    ii, jj, kk = _dev(d["ii"]), _dev(d["jj"]), _dev(d["kk"])
    lm = torch.tensor([1e-4], device="cuda")
    side = torch.cuda.Stream()
    done = torch.cuda.Event()
    with torch.cuda.stream(side):
        ws = pkg.fastba.BA_prepare(poses, patches, ii, jj, kk, t0, t1, d["M"], eff_impl=eff)
        done.record(side)
    torch.cuda.current_stream().wait_event(done)
    for _ in range(2):                                   # the prepared workspace is consumed: prepare again per call
        poses.copy_(_dev(d["poses"])[None]); patches.copy_(_dev(d["patches"])[None])
        pkg.fastba.BA(poses, patches, _dev(d["intrinsics"])[None], _dev(d["target"])[None], _dev(d["weight"])[None], lm,
                      ii, jj, kk, t0, t1, d["M"], 2, eff, prepared=ws)
        torch.cuda.synchronize()
        got = poses[0].cpu().numpy(), patches[0].cpu().numpy()
        assert _rel(got[0], one[0]) < 2e-5 and _rel(got[1][:, 2], one[1][:, 2]) < 2e-5
        ws = pkg.fastba.BA_prepare(poses, patches, ii, jj, kk, t0, t1, d["M"], eff_impl=eff, workspace=ws)
