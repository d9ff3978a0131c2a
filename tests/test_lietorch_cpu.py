"""lietorch_backends on CPU tensors (`dpvo_lie_host`: the device templates compiled for the host — the reference's own
device dispatch, lietorch/include/dispatch.h:12-43), all four groups, float32 and float64, against the golden vectors
of the reference's CPU backend (tests/golden/ref_cpu_lie_groups.npz) and, when oracle/_ref/lietorch_backends.so is
present (build container), against that backend directly on larger seeded batches.  No GPU needed."""
import numpy as np
import pytest
import torch

from conftest import load_pkg
from lie_groups_common import DIMS, JAC_OPS, all_ops, check_group
from oracle.ref_ext import load_ref


@pytest.mark.parametrize("G", [1, 2, 3, 4])
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.float64, 1e-10)])
def test_host_path_matches_reference_cpu_golden(G, dtype, tol):
    lb = load_pkg("lietorch_backends")
    check_group(lb, G, dtype, "cpu", tol)


@pytest.mark.parametrize("G", [1, 2, 3, 4])
def test_host_path_matches_reference_cpu_backend_large_batch(G):
    ref = load_ref("lietorch_backends")
    if ref is None:
        pytest.skip("oracle/_ref/lietorch_backends.so not built")
    lb = load_pkg("lietorch_backends")
    K, N = DIMS[G]
    g = torch.Generator().manual_seed(100 + G)
    n = 3000
    a = torch.randn(n, K, generator=g, dtype=torch.float64) * 0.6
    if G in (2, 4):
        a[:, K - 1] *= 0.4
    d = dict(a=a, X=ref.expm(G, a), Y=ref.expm(G, 0.5 * torch.randn(n, K, generator=g, dtype=torch.float64)),
             p3=torch.randn(n, 3, generator=g, dtype=torch.float64),
             p4=torch.randn(n, 4, generator=g, dtype=torch.float64), aK=torch.randn(n, K, generator=g, dtype=torch.float64),
             gN=torch.randn(n, N, generator=g, dtype=torch.float64), gK=torch.randn(n, K, generator=g, dtype=torch.float64),
             g3=torch.randn(n, 3, generator=g, dtype=torch.float64), g4=torch.randn(n, 4, generator=g, dtype=torch.float64))
    want = all_ops(ref, G, d)
    got = all_ops(lb, G, d)
    rot = slice(0, 3) if G in (1, 2) else slice(3, 6)
    ok = (a[:, rot].norm(dim=1) > 1e-2).numpy()
    for k in want:
        w, h = want[k].numpy(), got[k].numpy()
        if k in JAC_OPS:
            w, h = w[ok], h[ok]
        err = np.linalg.norm(h - w) / max(np.linalg.norm(w), 1e-30)
        assert err < 1e-10, (G, k, err)


def test_host_path_errors_and_empty():
    lb = load_pkg("lietorch_backends")
    assert lb.inv(4, torch.empty(0, 8)).shape == (0, 8)
    with pytest.raises(RuntimeError):
        lb.inv(5, torch.zeros(2, 7))
    with pytest.raises(RuntimeError):
        lb.inv(3, torch.zeros(2, 8))
    with pytest.raises(RuntimeError):
        lb.mul(3, torch.zeros(2, 7), torch.zeros(3, 7))
