"""CPU suite, part 3: the N > 1 host logic under a real 2-process gloo group:
sequence sharding without a collective, barrier + max-over-ranks timing as bench.py uses it, and the
data-parallel flat-buffer gradient all-reduce with the folded found-inf flag."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import load_pkg, ROOT


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _model(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3))


def _worker(rank, world, port, q):
    import sys
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    par = load_pkg("parallel")
    r, w = par.init(backend="gloo")
    assert (r, w) == (rank, world) and par.world_size() == world
    out = {}
    # --- inference: shard by independent sequence, no collective; time = max over ranks
    out["shard"] = par.shard_sequences(5, rank, world)
    par.barrier()
    out["max"] = par.max_over_ranks(10.0 + rank)
    out["sum"] = par.sum_over_ranks(1.0 + rank)
    # --- training: same weights everywhere, different clip per rank
    model = _model(0)
    red = par.FlatGradAllReducer(model.parameters())
    x = torch.randn(4, 7, generator=torch.Generator().manual_seed(100 + rank))
    red.zero_grad()
    model(x).square().sum().backward()
    local = red.flat[:red.numel].clone()
    inf = red.allreduce()
    # plain numpy through the queue: torch tensors travel by fd passing, which races with this worker's exit
    out["local"] = local.numpy().copy()
    out["reduced"] = red.flat[:red.numel].clone().numpy().copy()
    out["grad_is_view"] = all(p.grad.data_ptr() >= red.flat.data_ptr() for p in model.parameters())
    out["inf"] = bool(inf)
    # --- overflow on ONE rank must make EVERY rank skip
    red.zero_grad()
    model(x).square().sum().backward()
    if rank == 1:
        red.flat[3] = float("inf")
    out["inf2"] = bool(red.allreduce())
    out["finite_after"] = bool(torch.isfinite(red.flat).all())
    q.put((rank, out))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_gloo_group():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=240) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # sharding: disjoint, complete, balanced
    assert sorted(res[0]["shard"] + res[1]["shard"]) == list(range(5))
    assert abs(len(res[0]["shard"]) - len(res[1]["shard"])) <= 1
    # timing reductions
    assert res[0]["max"] == res[1]["max"] == 11.0
    assert res[0]["sum"] == res[1]["sum"] == 3.0
    # gradient all-reduce == mean of the per-rank gradients, identical on both ranks, in place
    mean = 0.5 * (res[0]["local"] + res[1]["local"])
    for r in range(world):
        assert np.allclose(res[r]["reduced"], mean, rtol=1e-6, atol=1e-7)
        assert res[r]["grad_is_view"] and not res[r]["inf"]
        assert res[r]["inf2"] and res[r]["finite_after"]
    assert np.array_equal(res[0]["reduced"], res[1]["reduced"])


def test_shard_sequences_properties():
    par = load_pkg("parallel")
    for n in (0, 1, 7, 8, 1000):
        for world in (1, 2, 4, 8):
            parts = [par.shard_sequences(n, r, world) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1
    with pytest.raises(ValueError):
        par.shard_sequences(4, 2, 2)


def test_single_process_reducer_is_local():
    par = load_pkg("parallel")
    model = _model(1)
    red = par.FlatGradAllReducer(model.parameters())
    model(torch.ones(2, 7)).sum().backward()
    before = red.flat[:red.numel].clone()
    assert not bool(red.allreduce())
    assert torch.equal(before, red.flat[:red.numel])
    assert red.nbytes() == 4 * (sum(p.numel() for p in model.parameters()) + 1)


def test_reducer_survives_optimizer_zero_grad_set_to_none():
    """The reference's loop calls `optimizer.zero_grad()` (train.py:295), which on torch >= 2 sets `.grad = None`;
    backward then allocates gradients outside the flat buffer.  allreduce() must pick them up (copy + re-point)
    instead of reducing a stale buffer."""
    par = load_pkg("parallel")
    model = _model(3)
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    red = par.FlatGradAllReducer(model.parameters())
    x = torch.ones(2, 7)
    for step in range(3):
        opt.zero_grad()                                   # set_to_none=True
        assert all(p.grad is None for p in model.parameters())
        (model(x) * (step + 1)).square().sum().backward()
        want = torch.cat([p.grad.reshape(-1).clone() for p in model.parameters()])
        assert not bool(red.allreduce(sanitize=False))
        assert torch.equal(red.flat[:red.numel], want)
        for p, sl in zip(red.params, red.slices):
            assert p.grad.data_ptr() == sl.data_ptr()     # re-pointed into the buffer
        opt.step()
    assert red.rebinds == 3 * len(red.params)
    # with the reducer's own zero_grad nothing is copied
    red.zero_grad()
    model(x).sum().backward()
    n = red.rebinds
    red.allreduce()
    assert red.rebinds == n
    # sanitize=False keeps a non-finite gradient visible to GradScaler.unscale_
    red.flat[0] = float("nan")
    assert bool(red.allreduce(sanitize=False)) and not bool(torch.isfinite(red.flat[:red.numel]).all())
