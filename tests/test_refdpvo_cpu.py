"""Host logic of the reference-DPVO harness (tools/refdpvo.py) without a GPU: the reference's own `dpvo.dpvo.DPVO` class
(oracle/_ref/site, installed by oracle/build_ref.py from /root/reference) is driven for a few frames with every "cuda"
device request redirected to the CPU and the oracle restatements standing in for cuda_corr / cuda_ba (the reference's
lietorch_backends.so supplies its CPU backend).  Checks the pieces the GPU parity tests rely on: the two stubs, the
synthetic sequence, the teacher (ground-truth reprojection keeps BA converging), the per-frame log, and the DPVO
bookkeeping rules the B200 `DPVO` class mirrors (edge counts per frame during initialisation)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import refdpvo  # noqa: E402


def test_synthetic_sequence_is_a_closed_loop_with_unit_quaternions():
    seq = refdpvo.SynthSequence(200, seed=3)
    assert seq.gt.shape == (200, 7)
    assert np.allclose(np.linalg.norm(seq.gt[:, 3:], axis=1), 1.0, atol=1e-12)
    # the camera comes back to where it started (loop closure material) and has slow stretches (keyframe removal)
    d = np.linalg.norm(seq.centres - seq.centres[0], axis=1)
    assert d.max() > 4.0 and d[60:200].min() < 0.5
    step = np.linalg.norm(np.diff(seq.centres, axis=0), axis=1)
    assert step.min() < 0.1 < 0.25 < step.max()
    img = seq.image(17)
    assert img.shape == (3, 480, 640) and img.dtype == np.uint8 and img.std() > 10


def test_config_is_default_yaml_over_config_py():
    if not refdpvo.have_reference_install():
        pytest.skip("oracle/_ref/site not installed")
    cfg = refdpvo.make_cfg()
    assert (cfg.PATCHES_PER_FRAME, cfg.REMOVAL_WINDOW, cfg.OPTIMIZATION_WINDOW, cfg.PATCH_LIFETIME) == (96, 22, 10, 13)
    assert cfg.LOOP_CLOSURE is True and cfg.MAX_EDGE_AGE == 1000 and cfg.BUFFER_SIZE == 4096
    assert refdpvo.make_cfg(LOOP_CLOSURE=False).LOOP_CLOSURE is False


@pytest.mark.timeout(600)
def test_reference_dpvo_class_dry_run_on_cpu(tmp_path):
    if not (refdpvo.have_reference_install() and os.path.exists(os.path.join(refdpvo.REF_DIR, "lietorch_backends.so"))):
        pytest.skip("oracle/_ref (reference python package + lietorch CPU backend) not built")
    out = str(tmp_path / "dry.npz")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "refdpvo.py"), "--cpu-dryrun", "--frames", "11",
                        "--patches", "4", "--out", out], capture_output=True, text=True, timeout=580)
    assert r.returncode == 0, r.stderr[-2000:]
    s = json.loads(r.stdout.strip().splitlines()[-1])
    d = np.load(out)
    M = 4
    # frames 0..7 initialise (n = 1..8); edge count after frame with n frames = M * n^2 while n <= PATCH_LIFETIME
    assert list(d["n"][:8]) == list(range(1, 9))
    assert list(d["E"][:8]) == [M * n * n for n in range(1, 9)]
    assert s["final_n"] >= 9 and "Warning BA failed" not in r.stdout
    # teacher-forced BA: the estimated camera moves along the ground-truth direction (scale is free)
    p = d["final_poses"]
    assert np.isfinite(p).all() and np.linalg.norm(p[-1, :3]) > 10 * np.linalg.norm(p[1, :3]) * 0.05
