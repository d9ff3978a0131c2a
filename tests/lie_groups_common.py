"""Shared by tests/test_lietorch_cpu.py (host path, runs without a GPU) and tests/test_lietorch_gpu.py (device path):
every lietorch_backends operator on the golden inputs of tests/golden/ref_cpu_lie_groups.npz (produced by the
reference's own CPU backend, make_golden_lie_groups.py)."""
import os

import numpy as np
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_cpu_lie_groups.npz")
DIMS = {1: (3, 4), 2: (4, 5), 3: (6, 7), 4: (7, 8)}
INPUTS = ("a", "X", "Y", "p3", "p4", "aK", "gN", "gK", "g3", "g4")
# operators that go through the SO3 left Jacobian / Q, where this implementation deliberately uses the series below
# theta = 1e-2 (lie.cuh SeriesTh) and the reference's closed forms are rounding noise in float32 just above 1e-6
JAC_OPS = ("Jinv", "expm_bwd", "logm_bwd")


def all_ops(lb, G, d):
    X, Y = d["X"], d["Y"]
    out = dict(exp=lb.expm(G, d["a"]), log=lb.logm(G, X), inv=lb.inv(G, X), mul=lb.mul(G, X, Y), adj=lb.adj(G, X, d["aK"]),
               adjT=lb.adjT(G, X, d["aK"]), act=lb.act(G, X, d["p3"]), act4=lb.act4(G, X, d["p4"]),
               matrix=lb.as_matrix(G, X), projector=lb.projector(G, X), Jinv=lb.Jinv(G, X, d["aK"]))
    out["expm_bwd"] = lb.expm_backward(G, d["gN"], d["a"])[0]
    out["logm_bwd"] = lb.logm_backward(G, d["gK"], X)[0]
    out["inv_bwd"] = lb.inv_backward(G, d["gN"], X)[0]
    out["mul_bwd_X"], out["mul_bwd_Y"] = lb.mul_backward(G, d["gN"], X, Y)
    out["act_bwd_X"], out["act_bwd_p"] = lb.act_backward(G, d["g3"], X, d["p3"])
    out["act4_bwd_X"], out["act4_bwd_p"] = lb.act4_backward(G, d["g4"], X, d["p4"])
    out["adj_bwd_X"], out["adj_bwd_a"] = lb.adj_backward(G, d["gK"], X, d["aK"])
    out["adjT_bwd_X"], out["adjT_bwd_a"] = lb.adjT_backward(G, d["gK"], X, d["aK"])
    return out


def check_group(lb, G, dtype, device, tol):
    g = np.load(GOLD)
    gold = {k[len(f"g{G}_"):]: g[k] for k in g.files if k.startswith(f"g{G}_")}
    d = {k: torch.from_numpy(gold[k]).to(dtype).to(device).contiguous() for k in INPUTS}
    got = all_ops(lb, G, d)
    rot = slice(0, 3) if G in (1, 2) else slice(3, 6)
    th = np.linalg.norm(gold["a"][:, rot], axis=1)
    ok = (th < 1e-6) | (th > 1e-2)
    worst = {}
    for k, v in got.items():
        want = gold["X"] if k == "exp" else gold[k]
        have = v.detach().cpu().double().numpy()
        assert have.shape == want.shape, (k, have.shape, want.shape)
        assert v.dtype == dtype and v.device.type == torch.device(device).type
        if k in JAC_OPS:
            have, want = have[ok], want[ok]
        err = np.linalg.norm(have - want) / max(np.linalg.norm(want), 1e-30)
        worst[k] = err
        assert err < tol, (G, k, err)
        scale = np.abs(want).reshape(want.shape[0], -1).max(1).reshape((-1,) + (1,) * (want.ndim - 1)) + 1.0
        assert np.all(np.abs(have - want) <= 8 * tol * scale), (G, k)
    return worst
