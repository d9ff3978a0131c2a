"""`dpvo_chol_solve_noalloc` (include/dpvo_b200.h): the blocked Cholesky solve with its scratch in a file-scope device
array of the library instead of a caller workspace — no allocation inside the call.  Called through the C-ABI with
ctypes; outputs compared with the existing `dpvo_chol_solve` entry point on the same inputs (same kernels, no atomics:
the two must agree to the last bit; asserted at the 2e-5 relative bar of test_blocked_cholesky_solve, with the bit-exact
check on top) on two calls in a row that share the static array (the second problem is smaller and would read the first
one's left-over bytes if anything that must be rebuilt per call were not).  File name keeps it last in pytest's order."""
import ctypes

import numpy as np
import pytest
import torch

from conftest import load_pkg

pytestmark = pytest.mark.gpu

c_p, c_i32, c_i64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64


def _system(n, seed):
    rng = np.random.default_rng(seed)
    M = rng.standard_normal((n, n + 8))
    S = (M @ M.T / (n + 8) + 0.5 * np.eye(n)).astype(np.float32)
    y = rng.standard_normal(n).astype(np.float32)
    return torch.from_numpy(S).cuda(), torch.from_numpy(y).cuda()


def _rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def _existing(lib, S, y, stream):
    n = y.numel()
    lib.dpvo_chol_workspace_bytes.restype = c_i64
    nbytes = int(lib.dpvo_chol_workspace_bytes(c_i32(n)))
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    x = torch.zeros(n, dtype=torch.float32, device="cuda")
    info = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    rc = lib.dpvo_chol_solve(c_p(S.data_ptr()), c_p(y.data_ptr()), c_i32(n), c_p(x.data_ptr()), c_p(info.data_ptr()),
                             c_p(ws.data_ptr()), c_i64(nbytes), stream)
    assert rc == 0, lib.dpvo_last_error()
    return x, info


def test_chol_solve_noalloc_matches_workspace_entry_point_twice_in_a_row():
    lib = load_pkg()._lib.lib
    lib.dpvo_last_error.restype = ctypes.c_char_p
    lib.dpvo_chol_solve_noalloc_max_n.restype = c_i32
    max_n = int(lib.dpvo_chol_solve_noalloc_max_n())
    assert max_n == 768
    torch.cuda.set_device(0)
    stream = c_p(torch.cuda.current_stream().cuda_stream)
    # first call: the largest size (fills the whole static array); second: a smaller, different system, then one with
    # a non-positive pivot (the failure flag lives in the static array too and must not stick), then a good one again
    cases = [(max_n, 1, False), (234, 2, False), (130, 3, True), (65, 4, False)]
    for n, seed, poison in cases:
        S, y = _system(n, seed)
        if poison:
            S[n // 2, n // 2] = -1.0
        # outputs are the caller's, allocated BEFORE the call; the allocator statistics show the call itself allocated
        # nothing through torch, and the entry point has no workspace argument to allocate into (free device memory is
        # not compared: the driver loads a module's kernels lazily on their first launch)
        x = torch.full((n,), 7.0, dtype=torch.float32, device="cuda")
        info = torch.full((1,), -1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]
        rc = lib.dpvo_chol_solve_noalloc(c_p(S.data_ptr()), c_p(y.data_ptr()), c_i32(n), c_p(x.data_ptr()),
                                         c_p(info.data_ptr()), stream)
        torch.cuda.synchronize()
        assert rc == 0, lib.dpvo_last_error()
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == before
        xe, infoe = _existing(lib, S, y, stream)
        torch.cuda.synchronize()
        assert int(info.item()) == int(infoe.item()) == (1 if poison else 0)
        if poison:
            assert float(x.abs().max()) == 0.0 and float(xe.abs().max()) == 0.0
            continue
        a, b = x.cpu().numpy(), xe.cpu().numpy()
        assert np.isfinite(a).all()
        assert _rel(a, b) < 2e-5, (n, _rel(a, b))
        assert np.array_equal(a, b), "same kernels, no atomics: expected bit-identical results"
        want = np.linalg.solve(S.cpu().numpy().astype(np.float64), y.cpu().numpy().astype(np.float64))
        assert _rel(a, want) < 2e-5


def test_chol_solve_noalloc_rejects_sizes_beyond_the_static_scratch():
    lib = load_pkg()._lib.lib
    lib.dpvo_last_error.restype = ctypes.c_char_p
    lib.dpvo_chol_solve_noalloc_max_n.restype = c_i32
    n = int(lib.dpvo_chol_solve_noalloc_max_n()) + 1
    S, y = _system(n, 5)
    x = torch.zeros(n, dtype=torch.float32, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.dpvo_chol_solve_noalloc(c_p(S.data_ptr()), c_p(y.data_ptr()), c_i32(n), c_p(x.data_ptr()),
                                     c_p(info.data_ptr()), c_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.dpvo_last_error().decode()
    assert "exceeds the static scratch" in msg and "768" in msg, msg
    # n = 0 is a successful no-op like the workspace entry point
    assert lib.dpvo_chol_solve_noalloc(c_p(0), c_p(0), c_i32(0), c_p(0), c_p(0), c_p(0)) == 0
