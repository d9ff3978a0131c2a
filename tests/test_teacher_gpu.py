"""The sequence harness' teacher (tools/refdpvo.py) as one launch (tools/teacher_kernel.cu) against its torch float64
formulation on the same inputs.  Harness code, not product: both arms of the sequence comparison run the same teacher;
this only shows that the one-launch form computes the same targets (float64 geometry, then float32: 1e-5 px absolute on
targets of tens of pixels; validity identical away from the Z = 0.2 threshold)."""
import os
import sys
import types

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import refdpvo  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_teacher_kernel_matches_torch_formulation(dtype):
    if refdpvo._TEACHER_LIB is None:
        pytest.fail("tools/libteacher.so is missing: run __graft_entry__.build()")
    frames, n, M, E = 40, 22, 96, 6000
    seq = refdpvo.SynthSequence(frames, seed=5)
    g = torch.Generator().manual_seed(3)
    pg = types.SimpleNamespace()
    pg.tstamps_ = np.zeros(64, dtype=np.int64)
    pg.tstamps_[:n + 2] = np.sort(np.random.RandomState(1).choice(frames, n + 2, replace=False))
    pt = torch.zeros(64, M, 3, 3, 3)
    pt[:, :, 0] = torch.rand(64, M, 1, 1, generator=g) * 640 / 4
    pt[:, :, 1] = torch.rand(64, M, 1, 1, generator=g) * 480 / 4
    pt[:, :, 2] = torch.rand(64, M, 1, 1, generator=g) + 0.5
    pg.patches_ = pt.cuda()
    pg.intrinsics_ = (torch.as_tensor(seq.intrinsics, dtype=torch.float32) / 4).repeat(64, 1).cuda()
    slam = types.SimpleNamespace(pg=pg, n=n)
    ii = torch.randint(0, n + 1, (E,), generator=g).cuda()
    jj = torch.randint(0, n + 1, (E,), generator=g).cuda()
    kk = (ii.cpu() * M + torch.randint(0, M, (E,), generator=g)).cuda()
    slam._last_coords = (torch.rand(1, E, 2, 3, 3, generator=g) * 100).cuda()
    delta = (torch.randn(1, E, 2, generator=g) * 2).cuda().to(dtype)
    weight = torch.rand(1, E, 2, generator=g).cuda().to(dtype)
    inner = lambda net, *a, **k: (net, (delta, weight, None))
    teacher = refdpvo.make_teacher(inner, slam, seq, 0.25)
    net = torch.zeros(1, E, 4).cuda()
    _, (d1, w1, _) = teacher(net, None, None, None, ii, jj, kk)
    lib = refdpvo._TEACHER_LIB
    try:
        refdpvo._TEACHER_LIB = None
        _, (d0, w0, _) = teacher(net, None, None, None, ii, jj, kk)
        gt, ok = teacher.gt_centres(ii, jj, kk)
    finally:
        refdpvo._TEACHER_LIB = lib
    assert d1.shape == d0.shape == (1, E, 2) and d1.dtype == torch.float32 and w1.dtype == torch.float32
    assert 0.2 < float(ok.float().mean()) and float((d0 != 0).float().mean()) > 0.2      # the case exercises both branches
    same = (w1 != 0).any(-1) == (w0 != 0).any(-1)
    assert float(same.float().mean()) > 0.9995                                            # Z = 0.2 ties only
    s = same[0]
    assert float((d1[0, s] - d0[0, s]).abs().max()) < 1e-3 * max(1.0, float(d0.abs().max()) * 1e-2)
    assert float((d1[0, s] - d0[0, s]).abs().median()) < 1e-5
    assert torch.equal(w1[0, s], w0[0, s])
