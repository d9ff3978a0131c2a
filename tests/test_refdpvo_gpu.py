"""Row X1 / (b) / A18: the reference's OWN `dpvo.dpvo.DPVO` class (dpvo/dpvo.py, untouched copy installed by
oracle/build_ref.py under oracle/_ref/site) runs a synthetic 480x640 sequence twice — once on the reference's CUDA
extensions (oracle/_ref/*.so), once on this repo's drop-in `cuda_corr` / `cuda_ba` / `lietorch_backends` — with identical
seeds, weights, images and teacher (tools/refdpvo.py).  Each arm is its own process because the reference binds the
extension modules at import time.

Asserted (north star): per-frame edge bookkeeping ii/jj/kk BIT-EXACT (sha1 per frame + full lists on event frames) over
120 frames that include the 8-frame initialisation with 12 updates (dpvo.py:461-466), keyframe removals
(dpvo.py:266-310), proximity loop-closure edges (patchgraph.py:56-82) and global bundle adjustments with eff_impl=True
(dpvo.py:312-326); poses of the window (in DPVO's own scale normalisation, patchgraph.py:84-90 — monocular scale is a
gauge freedom of the sliding window and drifts freely between global BAs) within max(POSE_FLOOR, 3 x the reference
arm's OWN run-to-run spread) at every frame: the reference accumulates its Hessian with float atomics in a varying
order, so two runs of the reference itself do not repeat; that spread is measured here by running the reference arm
twice.  The two arms differ by fp16 rounding in the lookup and fp32 summation order in BA; the teacher keeps the
trajectory contracting, so the difference stays at that level instead of growing chaotically."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import refdpvo  # noqa: E402

pytestmark = pytest.mark.gpu

FRAMES = 120
POSE_FLOOR = 2e-4        # absolute, on scale-normalised poses (trajectory extent ~1) / unit quaternions


def _run(backend, out, update="reference", frames=FRAMES, pipeline="reference"):
    cmd = [sys.executable, os.path.join(ROOT, "tools", "refdpvo.py"), "--backend", backend, "--mode", "parity",
           "--frames", str(frames), "--update", update, "--pipeline", pipeline, "--out", out]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stderr[-3000:]
    summary = json.loads(r.stdout.strip().splitlines()[-1])
    return summary, dict(np.load(out)), r.stderr


@pytest.fixture(scope="module")
def arms(tmp_path_factory):
    if not (refdpvo.have_reference_install() and refdpvo.have_reference_ext()):
        pytest.skip("oracle/_ref (reference extensions + python package) not built")
    d = tmp_path_factory.mktemp("refdpvo")
    ref = _run("ref", str(d / "ref.npz"))
    ref2 = _run("ref", str(d / "ref2.npz"))
    ours = _run("dropin", str(d / "dropin.npz"))
    return ref, ours, ref2


def _pose_diff(a, b):
    return np.abs(a["poses_tail"] - b["poses_tail"]).reshape(len(a["n"]), -1).max(1)


def _compare(ref, ours, tag, ref2=None, floor=POSE_FLOOR):
    (sr, a, _), (so, b, err) = ref, ours
    assert "dropin" in so["modules"]["cuda_corr"] and "_ref" in sr["modules"]["cuda_corr"], (sr["modules"], so["modules"])
    assert "Warning BA failed" not in err, "the drop-in cuda_ba raised inside DPVO.update (dpvo.py:355-356 swallows it)"
    # the sequence exercises everything the state machine has
    assert sr["removed"] >= 3 and sr["gba"] >= 3 and sr["loop_added"] >= 1, sr
    first_bad = None
    for k in ("n", "m", "E", "Einac", "removed", "gba", "loop_added", "h_ii", "h_jj", "h_kk"):
        ne = np.nonzero(a[k] != b[k])[0]
        if len(ne):
            first_bad = (k, int(ne[0]))
            break
    dpose = _pose_diff(a, b)
    spread = _pose_diff(a, ref2[1]) if ref2 is not None and np.array_equal(a["n"], ref2[1]["n"]) else np.zeros(1)
    tol = max(floor, 3.0 * float(spread.max()))
    rec = {"what": tag, "frames": int(len(a["n"])), "final_n": int(a["final_n"]), "removed": int(a["removed"].sum()),
           "global_ba": int(a["gba"].sum()), "loop_edge_batches": int(a["loop_added"].sum()),
           "edges_last": int(a["E"][-1]), "inactive_last": int(a["Einac"][-1]),
           "first_bookkeeping_mismatch": first_bad, "pose_absdiff_max": float(dpose.max()),
           "reference_self_spread_max": float(spread.max()), "pose_tolerance": tol,
           "pose_absdiff_by_frame": [float(f"{x:.3e}") for x in dpose[::10]]}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "refdpvo_parity.jsonl"), "a") as f:
        f.write(json.dumps(rec) + "\n")
    assert first_bad is None, rec
    for key in a:
        if key[:3] in ("ii_", "jj_", "kk_"):
            assert np.array_equal(a[key], b[key]), key
    assert dpose.max() < tol, rec
    assert np.array_equal(a["tstamps"], b["tstamps"])
    sa, sb = a["final_depths"].mean(), b["final_depths"].mean()
    d_depth = np.abs(a["final_depths"] / sa - b["final_depths"] / sb).max()
    assert d_depth < max(5e-3, 10 * tol), d_depth


def test_reference_dpvo_class_on_dropins_matches_reference_extensions(arms):
    _compare(arms[0], arms[1], "dpvo.DPVO unmodified: drop-in cuda_corr/cuda_ba/lietorch_backends vs reference extensions",
             ref2=arms[2])


def test_reference_dpvo_class_with_b200_update_operator(arms, tmp_path):
    """Same sequence with `network.update` bound to the tcgen05 Update operator."""
    ours = _run("dropin", str(tmp_path / "dropin_upd.npz"), update="b200")
    # the operator keeps its residual stream in fp16 where autocast keeps fp32 (update.py docstring): its flow / weight
    # outputs differ from the torch module's at the 1e-2 level (bounded in tests/test_update_gpu.py), which the teacher's
    # ALPHA * tanh(.) passes into the BA targets -> a looser floor on the poses; the bookkeeping must still be bit-exact
    _compare(arms[0], ours, "dpvo.DPVO unmodified + dpvo_b200_update.Update vs reference extensions + torch Update",
             ref2=arms[2], floor=5e-3)


def test_b200_dpvo_class_matches_reference_class(arms, tmp_path):
    """This repo's OWN `DPVO` class (deep-visual-odometry-slam_b200/dpvo.py: the reference's API and state machine on
    the B200 operators and the device-resident buffer layout — channels-last rings, fixed-capacity device edge list,
    fused reprojection / lookup / Update operator / BA / encoders) against the reference class on the reference
    extensions, same sequence, seeds, weights and teacher: bit-identical ii / jj / kk at every frame, poses within the
    same measured bar (looser floor: its encoders and Update operator are the fp16 kernels, see the test above)."""
    ours = _run("dropin", str(tmp_path / "b200_pipeline.npz"), pipeline="b200")
    _compare(arms[0], ours, "deep-visual-odometry-slam_b200.dpvo.DPVO vs reference dpvo.DPVO on reference extensions",
             ref2=arms[2], floor=5e-3)
