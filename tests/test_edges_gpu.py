"""GPU parity for the device-side patch-graph edge bookkeeping (N4) through the C ABI (dpvo_edges_*,
dpvo_frames_shift): the scripted 40-frame sequences of tests/golden/ref_cpu_edges.npz (produced by the
reference's own append_factors / remove_factors / keyframe bodies on CPU) must be reproduced bit for bit,
frame by frame; large masks are checked against the numpy oracle; capacity / index errors raise."""
import numpy as np
import pytest
import torch

import edges_replay
from conftest import load_pkg

pytestmark = pytest.mark.gpu


class _DeviceEdges:
    def __init__(self, m, seed, capacity=1024):
        self.pg = load_pkg("patchgraph")
        self.m = m
        N, M = m.CFG["BUFFER_SIZE"], m.CFG["PATCHES_PER_FRAME"]
        ix = torch.arange(N, device="cuda")[:, None].repeat(1, M).reshape(-1).contiguous()
        self.e = self.pg.PatchGraphEdges(ix, M, capacity, net_dim=m.DIM)
        self.buf = {k: torch.from_numpy(v).cuda() for k, v in m.initial_buffers(seed).items()}
        self.r = m.CFG["PATCH_LIFETIME"]

    def forw(self, n): self.e.edges_forw(n, self.r)
    def back(self, n): self.e.edges_back(n, self.r)
    def count(self): return self.e.active.count()

    def set_rows(self, net, target, weight):
        a, E = self.e.active, net.shape[0]
        a.net_[:E] = torch.from_numpy(net).cuda()
        a.target_[:E] = torch.from_numpy(target).cuda()
        a.weight_[:E] = torch.from_numpy(weight).cuda()

    def remove_frame(self, k): self.e.remove_frame(k)
    def remove_old(self, *a): self.e.remove_old(*a)

    def shift(self, k, n):
        b, m, fs = self.buf, self.m, self.pg.frames_shift
        for key in ("tstamps_", "colors_", "poses_", "patches_", "intrinsics_"):
            fs(b[key], k, n)
        fs(b["imap_"], k, n, m.PMEM)
        fs(b["gmap_"], k, n, m.PMEM)
        fs(b["fmap1_"][0], k, n, m.MEM)
        fs(b["fmap2_"][0], k, n, m.MEM)

    def snapshot(self):
        a, i = self.e.active, self.e.inactive
        na, ni = a.count(), i.count()
        self.e.check_info()
        return dict(ii=a.ii_[:na].cpu().numpy(), jj=a.jj_[:na].cpu().numpy(), kk=a.kk_[:na].cpu().numpy(),
                    ii_inac=i.ii_[:ni].cpu().numpy(), jj_inac=i.jj_[:ni].cpu().numpy(), kk_inac=i.kk_[:ni].cpu().numpy())

    def final(self):
        a, i = self.e.active, self.e.inactive
        na, ni = a.count(), i.count()
        out = dict(net=a.net_[:na], target=a.target_[:na], weight=a.weight_[:na], target_inac=i.target_[:ni],
                   weight_inac=i.weight_[:ni])
        out.update(self.buf)
        return {k: v.cpu().numpy() for k, v in out.items()}


@pytest.mark.parametrize("name,seed,lc", [("plain", 0, False), ("loop", 1, True)])
def test_sequence_matches_reference_bit_for_bit(name, seed, lc):
    m = edges_replay.golden_module()
    assert edges_replay.replay(_DeviceEdges(m, seed), name, seed, lc) == 14


@pytest.mark.parametrize("E,dim,store", [(45312, 384, True), (45312, 384, False), (5000, 8, True), (1025, 2, True),
                                         (1024, 0, False), (1, 8, True)])
def test_remove_factors_large_masks_vs_oracle(E, dim, store):
    """Multi-chunk compaction at the default window's edge count: order kept, rows moved with their edges."""
    from oracle import edges_oracle as eo
    pg = load_pkg("patchgraph")
    rng = np.random.default_rng(E + dim)
    M, frames = 96, 40
    ix = np.repeat(np.arange(frames), M)
    kk = rng.integers(0, frames * M, E)
    jj = rng.integers(0, frames, E)
    o = eo.Edges(ix, M, dim)
    o.append_factors(kk, jj)
    o.net = rng.standard_normal((E, dim)).astype(np.float16)
    o.target = rng.standard_normal((E, 2)).astype(np.float32)
    o.weight = rng.random((E, 2)).astype(np.float32)
    d = pg.PatchGraphEdges(torch.from_numpy(ix).cuda(), M, E + 7, net_dim=dim)
    d.append_factors(torch.from_numpy(kk).cuda(), torch.from_numpy(jj).cuda())
    a = d.active
    if dim:
        a.net_[:E] = torch.from_numpy(o.net).cuda()
    a.target_[:E] = torch.from_numpy(o.target).cuda()
    a.weight_[:E] = torch.from_numpy(o.weight).cuda()
    for rnd, p in enumerate((0.3, 0.0, 0.9, 1.0)):
        n = o.ii.size
        mask = rng.random(n) < p
        o.remove_factors(mask, store)
        d.remove_factors(torch.from_numpy(mask).cuda(), store)
        na = a.count()
        d.check_info()
        assert na == o.ii.size, rnd
        for key in ("ii", "jj", "kk", "target", "weight"):
            assert np.array_equal(getattr(a, key + "_")[:na].cpu().numpy(), getattr(o, key)), (rnd, key)
        if dim:
            assert np.array_equal(a.net_[:na].cpu().numpy().view(np.uint16), o.net.view(np.uint16)), rnd
        ni = d.inactive.count()
        assert ni == o.ii_inac.size
        for key in ("ii", "jj", "kk", "target", "weight"):
            assert np.array_equal(getattr(d.inactive, key + "_")[:ni].cpu().numpy(), getattr(o, key + "_inac")), (rnd, key)


def test_append_zeroes_rows_and_flags_overflow_and_bad_indices():
    pg = load_pkg("patchgraph")
    ix = torch.arange(10, device="cuda").repeat_interleave(4).contiguous()
    d = pg.PatchGraphEdges(ix, 4, 16, net_dim=8)
    d.active.net_.fill_(1.0); d.active.target_.fill_(2.0); d.active.weight_.fill_(3.0)
    d.append_grid(0, 4, 2, 5)                       # 12 edges
    assert d.active.count() == 12
    assert torch.equal(d.active.kk_[:12].cpu(), torch.arange(4).repeat_interleave(3))
    assert torch.equal(d.active.jj_[:12].cpu(), torch.arange(2, 5).repeat(4))
    assert float(d.active.net_[:12].abs().sum()) == 0 and float(d.active.target_[:12].abs().sum()) == 0
    assert float(d.active.net_[12:].min()) == 1.0
    d.check_info()
    d.append_grid(0, 4, 0, 2)                       # 8 more: 4 do not fit
    assert d.active.count() == 16
    with pytest.raises(RuntimeError, match="capacity"):
        d.check_info()
    d2 = pg.PatchGraphEdges(ix, 4, 16, net_dim=8)
    d2.append_factors(torch.tensor([1, 99], device="cuda"), torch.tensor([0, 0], device="cuda"))
    with pytest.raises(RuntimeError, match="outside ix"):
        d2.check_info()
    with pytest.raises(RuntimeError):
        d2.append_factors(torch.tensor([1, 2]), torch.tensor([0, 0]))          # CPU tensors refused
    with pytest.raises(RuntimeError):
        d2.append_factors(torch.tensor([1, 2], device="cuda"), torch.tensor([0], device="cuda"))


def test_reduce_edges_vs_reference():
    pg = load_pkg("patchgraph")
    m, g = edges_replay.golden_module(), edges_replay.golden()
    for i, c in enumerate(m.REDUCE_CASES):
        mag, ii, jj = m.reduce_case(c)
        dev = lambda x: torch.from_numpy(x).cuda()
        out, cnt = pg.reduce_edges(dev(mag), dev(ii), dev(jj), c["max_num_edges"], c["nms"], n_frames=c["frames"])
        n = int(cnt.item())
        assert np.array_equal(out[:n].cpu().numpy(), g[f"reduce{i}"]), i


@pytest.mark.parametrize("shape,mod,k,n", [((9, 5, 3), None, 2, 9), ((12, 7), 12, 9, 14), ((6, 1), None, 0, 6),
                                           ((8, 16, 4), 8, 5, 6), ((8, 4), None, 3, 3)])
def test_frames_shift_vs_oracle(shape, mod, k, n):
    """Byte path (odd slot sizes), 16-byte path, ring wrap-around (n > mod), single and empty moves."""
    from oracle import edges_oracle as eo
    pg = load_pkg("patchgraph")
    rng = np.random.default_rng(sum(shape))
    a = rng.integers(0, 255, shape).astype(np.uint8)
    d = torch.from_numpy(a.copy()).cuda()
    eo.frames_shift(a, k, n, mod)
    pg.frames_shift(d, k, n, mod)
    assert np.array_equal(d.cpu().numpy(), a)
    with pytest.raises(RuntimeError, match="ring slots"):
        pg.frames_shift(d, 0, shape[0] + 2, mod)
