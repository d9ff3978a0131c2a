"""GPU parity tests for the Update operator (N1): tensor-core edge-MLP layers (csrc/update_gemm.cu),
row kernels (csrc/update_agg.cu) and the whole operator (update.py), through the C ABI.

Oracle: oracle/update_oracle.py (float32 restatement of net.py:69-88, pinned on CPU against the
reference module's own output in tests/golden/ref_cpu_update.npz).

Tolerances.  The reference runs this operator under torch.autocast(fp16) (dpvo.py:332): fp16 GEMM
inputs / outputs, fp32 LayerNorm and residual stream.  The fused path stores fp16 activations with
fp32 accumulation and statistics, so both sit at fp16 rounding distance from the float32 oracle:
  * single layer: |err| <= 2e-3 * max|want| + 1e-3 (one fp16 rounding of the output);
  * whole operator: norm-wise relative error of net / delta / weight <= 1e-2 AND not more than
    3x the error of the autocast graph itself on the same inputs.
"""
import importlib
import os

import numpy as np
import pytest
import torch

from conftest import load_pkg
from oracle import update_oracle as uo

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a = a.double(); b = b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _ref_layer(A, W, b, gate=None, res1=None, res2=None, idx=None, ln=None, act=0):
    v = A.float() @ W.float().t() + b
    if gate is not None: v = v * gate.float()
    if res1 is not None: v = v + res1.float()
    if res2 is not None: v = v + (res2[idx] if idx is not None else res2).float()
    if ln is not None: v = torch.nn.functional.layer_norm(v, (384,), ln[0], ln[1], ln[2])
    if act == 1: v = v.relu()
    if act == 2: v = v.sigmoid()
    return v


@pytest.mark.parametrize("M,K,ld", [(1, 384, 384), (128, 64, 64), (300, 384, 384), (5000, 384, 392),
                                    (1000, 882, 896), (20000, 882, 888)])
def test_linear384_vs_torch(M, K, ld):
    pkg = load_pkg()
    ops = importlib.import_module(pkg.__name__ + ".update_ops")
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(M + K)
    r = lambda *s: torch.randn(*s, device=dev, generator=g)
    Abuf = torch.full((M, ld), float("nan"), dtype=torch.float16, device=dev)   # padding must never be read
    Abuf[:, :K] = r(M, K).half()
    A = Abuf[:, :K]
    Wp = torch.full((384, ld), float("nan"), dtype=torch.float16, device=dev)
    Wp[:, :K] = (r(384, K) / K ** 0.5).half()
    W = Wp[:, :K]
    b = r(384)
    gate = torch.rand(M, 384, device=dev, generator=g).half()
    res1 = r(M, 384).half()
    table = r(77, 384).half()
    idx = torch.randint(0, 77, (M,), device=dev, generator=g)
    gam, bet = r(384), r(384)
    cases = [dict(), dict(act=1), dict(act=2), dict(res1=res1), dict(gate=gate, res1=res1),
             dict(res1=res1, res2=table, idx=idx, ln=(gam, bet, 1e-3)), dict(ln=(gam, bet, 1e-3), act=1),
             dict(gate=gate, res1=res1, ln=(gam, bet, 1e-3))]
    for kw in cases:
        want = _ref_layer(A, W, b, **kw)
        got = ops.linear384(A, W, b, **kw).float()
        err = (got - want).abs().max().item()
        assert err <= 2e-3 * want.abs().max().item() + 1e-3, (sorted(kw), err)


def test_linear384_argument_errors():
    pkg = load_pkg()
    ops = importlib.import_module(pkg.__name__ + ".update_ops")
    A = torch.zeros(4, 384, dtype=torch.float16, device="cuda")
    W = torch.zeros(384, 384, dtype=torch.float16, device="cuda")
    with pytest.raises(RuntimeError):
        ops.linear384(A.float(), W)
    with pytest.raises(RuntimeError):
        ops.linear384(A, W[:100])
    with pytest.raises(RuntimeError):
        ops.linear384(A.cpu(), W.cpu())
    with pytest.raises(RuntimeError):
        ops.linear384(torch.zeros(4, 882, dtype=torch.float16, device="cuda"),
                      torch.zeros(384, 882, dtype=torch.float16, device="cuda"))      # 1764-byte rows: not TMA-able
    assert ops.linear384(A[:0], W).shape == (0, 384)


def _operator_case(pkg, n_frames, M, seed_w, seed_x):
    synth = importlib.import_module(pkg.__name__ + ".synth")
    sd_np = synth.update_state_dict(seed_w)
    d = synth.update_inputs(n_frames=n_frames, M=M, seed=seed_x)
    return sd_np, d


def _run_ours(pkg, sd_np, d, fused_ctx):
    upd_mod = importlib.import_module(pkg.__name__ + ".update")
    op = upd_mod.Update.from_state_dict({k: torch.from_numpy(v) for k, v in sd_np.items()}, device="cuda")
    ii, jj, kk = (torch.from_numpy(d[k]).cuda() for k in ("ii", "jj", "kk"))
    net = torch.from_numpy(d["net"]).cuda().half()[None]
    corr = torch.from_numpy(d["corr"]).cuda().half()[None]
    imap = torch.from_numpy(d["imap"]).cuda().half()
    tidx = kk % imap.shape[0]
    inp = (imap, tidx) if fused_ctx else imap[tidx][None]
    out, (delta, weight, none) = op(net, inp, corr, None, ii, jj, kk)
    assert none is None and out.shape == (1, len(d["ii"]), 384) and out.dtype == torch.float16
    assert delta.shape == weight.shape == (1, len(d["ii"]), 2)
    return out[0].float(), delta[0], weight[0]


def _run_oracle(sd_np, d, autocast):
    sd = uo.state_dict_tensors(sd_np, device="cuda")
    ii, jj, kk = (torch.from_numpy(d[k]).cuda() for k in ("ii", "jj", "kk"))
    # same fp16-rounded inputs as the fused path
    net = torch.from_numpy(d["net"]).cuda().half().float()
    corr = torch.from_numpy(d["corr"]).cuda().half().float()
    imap = torch.from_numpy(d["imap"]).cuda().half().float()
    with torch.no_grad():
        o = uo.forward(sd, net, imap[kk % imap.shape[0]], corr, ii, jj, kk, autocast=autocast)
    return tuple(t.float() for t in o)


@pytest.mark.parametrize("fused_ctx", [False, True])
def test_update_vs_reference_golden(fused_ctx):
    """Small graph: the reference nn.Module's own float32 output (CPU golden)."""
    pkg = load_pkg()
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_cpu_update.npz"))
    sd_np, d = _operator_case(pkg, int(g["n_frames"]), int(g["M"]), int(g["seed_w"]), int(g["seed_x"]))
    net, delta, weight = _run_ours(pkg, sd_np, d, fused_ctx)
    a_net, a_delta, a_weight = _run_oracle(sd_np, d, autocast=True)
    for name, got, auto in (("net", net, a_net), ("delta", delta, a_delta), ("weight", weight, a_weight)):
        want = torch.from_numpy(g[name]).cuda()
        e_ours, e_auto = _rel(got, want), _rel(auto, want)
        assert e_ours <= 1e-2, (name, e_ours, e_auto)
        assert e_ours <= 3 * e_auto + 1e-4, (name, e_ours, e_auto)


@pytest.mark.parametrize("n_frames,M", [(2, 1), (3, 4), (5, 13), (14, 16), (22, 96)])
def test_update_vs_oracle_sizes(n_frames, M):
    """Up to the default sliding window (22 frames x 96 patches): float32 oracle on the same device."""
    pkg = load_pkg()
    sd_np, d = _operator_case(pkg, n_frames, M, seed_w=11, seed_x=n_frames)
    net, delta, weight = _run_ours(pkg, sd_np, d, True)
    w_net, w_delta, w_weight = _run_oracle(sd_np, d, autocast=False)
    a_net, a_delta, a_weight = _run_oracle(sd_np, d, autocast=True)
    for name, got, want, auto in (("net", net, w_net, a_net), ("delta", delta, w_delta, a_delta),
                                  ("weight", weight, w_weight, a_weight)):
        e_ours, e_auto = _rel(got, want), _rel(auto, want)
        assert torch.isfinite(got).all()
        assert e_ours <= 1e-2, (name, e_ours, e_auto)
        assert e_ours <= 3 * e_auto + 1e-4, (name, e_ours, e_auto)


def test_softagg_groups_match_torch():
    """Grouped softmax / sum alone, ragged groups including singletons, large logits (running max)."""
    pkg = load_pkg()
    lib = pkg._lib
    L = lib.lib
    L.dpvo_softagg_workspace_bytes.restype = lib.c_i64
    g = torch.Generator(device="cuda").manual_seed(5)
    E, G = 5000, 700
    ids = torch.randint(0, G - 50, (E,), device="cuda", generator=g)           # ids >= G-50 stay empty
    ids[:3] = torch.tensor([G - 1, G - 1, G - 7], device="cuda")               # plus a pair and a singleton
    f = torch.randn(E, 384, device="cuda", generator=g).half()
    gm = (20 * torch.randn(E, 384, device="cuda", generator=g)).half()
    y = torch.zeros(G, 384, dtype=torch.float16, device="cuda")
    ws = torch.empty(int(L.dpvo_softagg_workspace_bytes(lib.c_i64(E), lib.c_i64(G))), dtype=torch.uint8, device="cuda")
    lib.check(L.dpvo_softagg_f16(lib.ptr(f), lib.ptr(gm), lib.ptr(ids), lib.c_i64(E), lib.c_i64(G), lib.ptr(y),
                                 lib.c_i32(0), lib.ptr(ws), lib.c_i64(ws.numel()), lib.stream_of(f)), "softagg")
    idx = ids.view(-1, 1).expand(E, 384)
    gmax = torch.full((G, 384), float("-inf"), device="cuda").scatter_reduce(0, idx, gm.float(), "amax", include_self=False)
    ex = torch.exp(gm.float() - gmax[ids])
    den = torch.zeros(G, 384, device="cuda").scatter_add_(0, idx, ex)
    want = torch.zeros(G, 384, device="cuda").scatter_add_(0, idx, f.float() * ex / (den[ids] + 1e-12))
    assert (y.float() - want).abs().max().item() <= 2e-3 * want.abs().max().item() + 1e-3
    empty = torch.ones(G, dtype=torch.bool, device="cuda"); empty[ids] = False
    assert empty.any() and (y[empty] == 0).all()


def test_update_iteration_matches_staged_oracle():
    """The full DPVO.update body (hotpath.UpdateIteration.step_full): its Update stage against the float32
    oracle on the correlation features the same run produced, and target = centre + delta feeding BA."""
    pkg = load_pkg()
    hotpath = importlib.import_module(pkg.__name__ + ".hotpath")
    synth = importlib.import_module(pkg.__name__ + ".synth")
    it = hotpath.UpdateIteration("cuda:0", n_frames=12, seed=2, M=16)
    net_in = it.net.clone()
    poses_before = it.poses0.clone()
    delta, weight = it.step_full()
    torch.cuda.synchronize()
    poses_after = it.poses.clone()
    # replay the stages by hand for the oracle
    it.poses.copy_(it.poses0); it.patches.copy_(it.patches0)
    coords = it.reproject()
    corr = it.corr(coords)                                               # contiguous [1,E,882] variant of the same lookup
    sd = uo.state_dict_tensors(synth.update_state_dict(2 + 3), device="cuda")
    with torch.no_grad():
        w_net, w_delta, w_weight = uo.forward(sd, net_in[0].float(), it.imap.float()[it.ctx_index], corr[0].float(),
                                              it.ii, it.jj, it.kk)
    assert _rel(delta[0], w_delta) <= 1e-2 and _rel(weight[0], w_weight) <= 1e-2
    assert _rel(it.net[0].float(), w_net) <= 1e-2
    assert torch.isfinite(poses_after).all() and not torch.equal(poses_after, poses_before)   # BA moved the window


def test_update_sparse_indices_and_contiguous_corr():
    """Frame / patch indices far from zero (long sequence: absolute indices), a contiguous [1,E,882] corr tensor
    (repacked to 16-byte rows inside) and bounds read from the index tensors (n_frames / n_patches not given)."""
    pkg = load_pkg()
    sd_np, d = _operator_case(pkg, 6, 8, seed_w=3, seed_x=9)
    d = dict(d)
    d["ii"] = d["ii"] + 900; d["jj"] = d["jj"] + 900; d["kk"] = d["kk"] + 900 * 8
    net, delta, weight = _run_ours(pkg, sd_np, d, False)
    w_net, w_delta, w_weight = _run_oracle(sd_np, d, autocast=False)
    assert _rel(net, w_net) <= 1e-2 and _rel(delta, w_delta) <= 1e-2 and _rel(weight, w_weight) <= 1e-2


def test_update_empty_edge_list():
    pkg = load_pkg()
    upd_mod = importlib.import_module(pkg.__name__ + ".update")
    synth = importlib.import_module(pkg.__name__ + ".synth")
    op = upd_mod.Update.from_state_dict({k: torch.from_numpy(v) for k, v in synth.update_state_dict(0).items()},
                                        device="cuda")
    z = torch.zeros(0, dtype=torch.int64, device="cuda")
    net, (delta, weight, _) = op(torch.zeros(1, 0, 384, device="cuda").half(), torch.zeros(1, 0, 384, device="cuda").half(),
                                 torch.zeros(1, 0, 882, device="cuda").half(), None, z, z, z)
    assert net.shape == (1, 0, 384) and delta.shape == (1, 0, 2) and weight.shape == (1, 0, 2)


def test_update_graph_cache_replays():
    """Opt-in CUDA-graph cache (update.Update.use_graphs): with the operands in persistent buffers the second call with
    the same signature captures, later calls replay; results must equal the eager run (to the operator's own run-to-run
    spread: its grouped sums and head dot products use float atomics), follow NEW contents of the same buffers, and a
    different edge count must fall back to an eager run."""
    pkg = load_pkg()
    upd_mod = importlib.import_module(pkg.__name__ + ".update")
    sd_np, d = _operator_case(pkg, 22, 32, seed_w=3, seed_x=4)
    op = upd_mod.Update.from_state_dict({k: torch.from_numpy(v) for k, v in sd_np.items()}, device="cuda")
    ii, jj, kk = (torch.from_numpy(d[k]).cuda() for k in ("ii", "jj", "kk"))
    E = ii.numel()
    assert E >= op.graph_min_edges
    net = torch.from_numpy(d["net"]).cuda().half()[None].contiguous()
    corr_buf = torch.zeros(1, E, 896, dtype=torch.float16, device="cuda")
    corr_buf[:, :, :882] = torch.from_numpy(d["corr"]).cuda().half()
    corr = corr_buf[:, :, :882]
    imap = torch.from_numpy(d["imap"]).cuda().half()
    tidx = (kk % imap.shape[0]).contiguous()
    nf, npat = int(max(ii.max(), jj.max())) + 1, int(kk.max()) + 1
    call = lambda: op(net, (imap, tidx), corr, None, ii, jj, kk, n_frames=nf, n_patches=npat)
    clone = lambda o: (o[0].clone(), o[1][0].clone(), o[1][1].clone())
    eager = clone(call())
    op.use_graphs = True
    first = clone(call())                       # first sighting: eager
    assert not getattr(op, "_graphs", {})
    second = clone(call())                      # second sighting: capture + replay
    assert len(op._graphs) == 1
    third = clone(call())                       # replay
    same = lambda a, b: _rel(a.float(), b.float()) < 1e-3
    for got in (first, second, third):
        for a, b in zip(got, eager):
            assert same(a, b)
    # new contents in the same buffers are picked up by the replay
    net.mul_(0.5)
    corr_buf.mul_(-1.0)
    replayed = clone(call())
    op.use_graphs = False
    want = clone(call())
    for a, b in zip(replayed, want):
        assert same(a, b)
    assert not same(replayed[0], eager[0])
    # a different edge count is a different signature: eager, cache untouched
    op.use_graphs = True
    sub = slice(0, E - 64)
    out = op(net[:, sub].contiguous(), (imap, tidx[sub].contiguous()), corr_buf[:, sub, :882], None, ii[sub].contiguous(),
             jj[sub].contiguous(), kk[sub].contiguous(), n_frames=nf, n_patches=npat)
    assert out[0].shape[1] == E - 64 and len(op._graphs) == 1


def test_update_one_call_path_matches_launch_by_launch_path():
    """csrc/update_forward.cu (the whole operator behind one C call) against update.py's launch-by-launch path: same
    kernels, same order, same operands -> equal to the operator's run-to-run spread (float atomics in the grouped sums)."""
    pkg = load_pkg()
    upd_mod = importlib.import_module(pkg.__name__ + ".update")
    for n_frames, M, fused in ((22, 96, True), (5, 13, False), (14, 16, True)):
        sd_np, d = _operator_case(pkg, n_frames, M, seed_w=5, seed_x=n_frames)
        op = upd_mod.Update.from_state_dict({k: torch.from_numpy(v) for k, v in sd_np.items()}, device="cuda")
        ii, jj, kk = (torch.from_numpy(d[k]).cuda() for k in ("ii", "jj", "kk"))
        net = torch.from_numpy(d["net"]).cuda().half()[None]
        corr = torch.from_numpy(d["corr"]).cuda().half()[None]
        imap = torch.from_numpy(d["imap"]).cuda().half()
        tidx = kk % imap.shape[0]
        inp = (imap, tidx) if fused else imap[tidx][None]
        assert not op.launch_by_launch
        a = op(net, inp, corr, None, ii, jj, kk)
        op.launch_by_launch = True
        b = op(net, inp, corr, None, ii, jj, kk)
        b2 = op(net, inp, corr, None, ii, jj, kk)
        spread = max(_rel(b2[0].float(), b[0].float()), _rel(b2[1][0], b[1][0]), _rel(b2[1][1], b[1][1]))
        for x, y in ((a[0], b[0]), (a[1][0], b[1][0]), (a[1][1], b[1][1])):
            assert _rel(x.float(), y.float()) <= max(1e-4, 3 * spread), (n_frames, M, _rel(x.float(), y.float()), spread)
