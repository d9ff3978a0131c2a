"""Device-resident patch-graph edge list.

Mirror of the edge bookkeeping the reference spreads over `DPVO` and `PatchGraph`
(dpvo.py:215-239 append_factors / remove_factors, :262-310 keyframe, :362-375 __edges_forw/__edges_back,
patchgraph.py:41-53 the edge / inactive-edge fields, loop_closure/optim_utils.py:27-64 reduce_edges), with
the same method names and argument meaning.  The reference re-allocates every array with torch.cat /
boolean-mask indexing (each mask index synchronises with the host); here the arrays have a fixed capacity,
the live count is a device scalar and every edit is a kernel of libdpvo_b200.so, so a frame's bookkeeping
enqueues without a sync.  `count()` (one 8-byte read) is the only call that waits for the device.

No CPU fallback: CPU tensors are refused.
"""
import ctypes

import torch

from ._lib import lib, ptr, stream_of, check, require_cuda, c_i32, c_i64, c_p

lib.dpvo_edges_workspace_bytes.restype = ctypes.c_int64
lib.dpvo_edges_reduce_workspace_bytes.restype = ctypes.c_int64


class _EdgeStore(ctypes.Structure):
    """Mirror of dpvo_edge_store (include/dpvo_b200.h)."""
    _fields_ = [("ii", c_p), ("jj", c_p), ("kk", c_p), ("net", c_p), ("net_ld", c_i64), ("net_dim", c_i32),
                ("reserved", c_i32), ("target", c_p), ("weight", c_p), ("count", c_p), ("capacity", c_i64)]


class EdgeList:
    """ii / jj / kk (+ optional net, target, weight rows) in fixed-capacity device arrays."""

    def __init__(self, capacity, device, net_dim=0, with_factors=True):
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("patchgraph.EdgeList: needs a CUDA device (no CPU path)")
        if net_dim % 2:
            raise RuntimeError("patchgraph.EdgeList: net_dim must be even")
        self.capacity, self.device, self.net_dim = int(capacity), dev, int(net_dim)
        z = lambda *s, dt=torch.int64: torch.zeros(s, dtype=dt, device=dev)
        self.ii_, self.jj_, self.kk_ = z(capacity), z(capacity), z(capacity)
        self.net_ = z(capacity, net_dim, dt=torch.float16) if net_dim else None
        self.target_ = z(capacity, 2, dt=torch.float32) if with_factors else None
        self.weight_ = z(capacity, 2, dt=torch.float32) if with_factors else None
        self.count_ = z(1)
        self.bound = 0                      # host-known upper bound on the live count
        self.c = _EdgeStore(ptr(self.ii_), ptr(self.jj_), ptr(self.kk_), ptr(self.net_), c_i64(net_dim),
                            c_i32(net_dim), c_i32(0), ptr(self.target_), ptr(self.weight_), ptr(self.count_),
                            c_i64(capacity))

    def count(self):
        """Live edges (synchronises; also tightens the host-side bound)."""
        n = int(self.count_.item())
        self.bound = n
        return n

    def _live(self, t, n=None):
        n = self.count() if n is None else n
        return None if t is None else t[:n]

    @property
    def ii(self): return self._live(self.ii_)
    @property
    def jj(self): return self._live(self.jj_)
    @property
    def kk(self): return self._live(self.kk_)
    @property
    def net(self): return None if self.net_ is None else self._live(self.net_)[None]
    @property
    def target(self): return self._live(self.target_)[None]
    @property
    def weight(self): return self._live(self.weight_)[None]


class PatchGraphEdges:
    """Active + inactive edge lists of one `DPVO` instance and the reference's edits on them.

    M = patches per frame, ix = the flat patch->frame table (`PatchGraph.ix`, patchgraph.py:117-119).
    """

    def __init__(self, ix, M, capacity, inactive_capacity=None, net_dim=384, device=None):
        require_cuda(ix)
        if ix.dtype != torch.int64 or not ix.is_contiguous():
            raise RuntimeError("patchgraph: ix must be a contiguous int64 tensor")
        dev = ix.device if device is None else torch.device(device)
        self.ix, self.M = ix, int(M)
        self.active = EdgeList(capacity, dev, net_dim=net_dim)
        self.inactive = EdgeList(capacity if inactive_capacity is None else inactive_capacity, dev, net_dim=0)
        self.info = torch.zeros(1, dtype=torch.int32, device=dev)
        self._mask = torch.zeros(self.active.capacity, dtype=torch.uint8, device=dev)
        nbytes = int(lib.dpvo_edges_workspace_bytes(c_i64(self.active.capacity), c_i32(net_dim)))
        self._ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)

    # ---- reference-named edits -----------------------------------------------------------------
    def append_factors(self, ii, jj):
        """dpvo.py:215-221: `ii` are PATCH indices (they become kk), jj target frames."""
        require_cuda(ii, jj)
        if ii.numel() != jj.numel():
            raise RuntimeError("patchgraph.append_factors: ii and jj differ in length")
        ii, jj = ii.to(torch.int64).contiguous(), jj.to(torch.int64).contiguous()
        a = self.active
        check(lib.dpvo_edges_append(ctypes.byref(a.c), ptr(self.ix), c_i64(self.ix.numel()), ptr(ii), ptr(jj),
                                    c_i64(ii.numel()), ptr(self.info), stream_of(self.ix)), "patchgraph.append_factors")
        a.bound = min(a.capacity, a.bound + ii.numel())

    def append_grid(self, kk0, kk1, jj0, jj1):
        """append_factors(*flatmeshgrid(arange(kk0, kk1), arange(jj0, jj1), indexing='ij'))."""
        a = self.active
        check(lib.dpvo_edges_append_grid(ctypes.byref(a.c), ptr(self.ix), c_i64(self.ix.numel()), c_i64(kk0), c_i64(kk1),
                                         c_i64(jj0), c_i64(jj1), ptr(self.info), stream_of(self.ix)),
              "patchgraph.append_grid")
        a.bound = min(a.capacity, a.bound + max(kk1 - kk0, 0) * max(jj1 - jj0, 0))

    def edges_forw(self, n, patch_lifetime):
        """dpvo.py:362-368: patches of the last `patch_lifetime` frames -> the newest frame n-1."""
        self.append_grid(self.M * max(n - patch_lifetime, 0), self.M * max(n - 1, 0), n - 1, n)

    def edges_back(self, n, patch_lifetime):
        """dpvo.py:370-375: patches of the newest frame -> the last `patch_lifetime` frames."""
        self.append_grid(self.M * max(n - 1, 0), self.M * max(n, 0), max(n - patch_lifetime, 0), n)

    def remove_factors(self, m, store):
        """dpvo.py:223-239.  m: bool / uint8 mask over the live edges (device tensor)."""
        require_cuda(m)
        a = self.active
        if m.dtype == torch.bool:
            m = m.view(torch.uint8)
        if m.dtype != torch.uint8 or not m.is_contiguous() or m.numel() < a.bound:
            raise RuntimeError("patchgraph.remove_factors: mask must be a contiguous bool/uint8 tensor over the live edges")
        self._remove(m, store)

    def _remove(self, mask, store):
        a = self.active
        inac = ctypes.byref(self.inactive.c) if store else None
        check(lib.dpvo_edges_remove(ctypes.byref(a.c), ptr(mask), inac, c_i64(min(a.bound, mask.numel())), ptr(self.info),
                                    ptr(self._ws), c_i64(self._ws.numel()), stream_of(self.ix)), "patchgraph.remove_factors")
        if store:
            self.inactive.bound = min(self.inactive.capacity, self.inactive.bound + a.bound)

    def remove_frame(self, k):
        """keyframe(), dpvo.py:280-285: drop the edges touching frame k, renumber the frames / patches after it."""
        a, st = self.active, stream_of(self.ix)
        check(lib.dpvo_edges_mask_keyframe(ctypes.byref(a.c), c_i64(k), ptr(self._mask), c_i64(a.bound), st),
              "patchgraph.remove_frame")
        self._remove(self._mask, False)
        check(lib.dpvo_edges_renumber(ctypes.byref(a.c), c_i64(k), c_i64(self.M), c_i64(a.bound), st),
              "patchgraph.remove_frame")

    def remove_old(self, n, removal_window, loop_closure=False, optimization_window=0):
        """keyframe(), dpvo.py:304-310: edges whose source frame left the window move to the inactive list."""
        a = self.active
        check(lib.dpvo_edges_mask_window(ctypes.byref(a.c), ptr(self.ix), c_i64(self.ix.numel()), c_i64(n),
                                         c_i64(removal_window), c_i32(1 if loop_closure else 0),
                                         c_i64(optimization_window), ptr(self._mask), c_i64(a.bound), stream_of(self.ix)),
              "patchgraph.remove_old")
        self._remove(self._mask, True)

    def check_info(self):
        """Raises if an earlier (asynchronous) edit overflowed a capacity or saw a bad patch index.  Synchronises."""
        v = int(self.info.item())
        if v == 1:
            raise RuntimeError("patchgraph: edge capacity exceeded")
        if v == 2:
            raise RuntimeError("patchgraph: patch index outside ix")


def frames_shift(buf, k, n, mod=None):
    """keyframe(), dpvo.py:287-297: `for i in range(k, n-1): buf[i % mod] = buf[(i+1) % mod]` on the leading
    dimension of a contiguous device tensor (poses_, patches_, intrinsics_, colors_, imap_, gmap_, fmap rings)."""
    require_cuda(buf)
    if not buf.is_contiguous():
        raise RuntimeError("patchgraph.frames_shift: buffer must be contiguous")
    mod = buf.shape[0] if mod is None else int(mod)
    if mod > buf.shape[0]:
        raise RuntimeError("patchgraph.frames_shift: mod exceeds the buffer's slots")
    slot = buf[0].numel() * buf.element_size()
    check(lib.dpvo_frames_shift(ptr(buf), c_i64(slot), c_i32(mod), c_i32(k), c_i32(n), stream_of(buf)),
          "patchgraph.frames_shift")


def reduce_edges(flow_mag, ii, jj, max_num_edges, nms, n_frames=None):
    """loop_closure/optim_utils.py:27-64 on the device.  Returns (edges [max_num_edges, 2] int64, count [1] int64)
    as device tensors: rows [:count] are the reference's result."""
    require_cuda(flow_mag, ii, jj)
    n = flow_mag.numel()
    if ii.numel() != n or jj.numel() != n:
        raise RuntimeError("patchgraph.reduce_edges: flow_mag, ii, jj differ in length")
    flow_mag = flow_mag.to(torch.float32).contiguous()
    ii, jj = ii.to(torch.int64).contiguous(), jj.to(torch.int64).contiguous()
    dev = flow_mag.device
    out = torch.zeros((max(max_num_edges, 1), 2), dtype=torch.int64, device=dev)
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)
    if n == 0:
        return out[:max_num_edges], cnt
    if n_frames is None:
        n_frames = int(max(ii.max().item(), jj.max().item())) + 1
    nbytes = int(lib.dpvo_edges_reduce_workspace_bytes(c_i64(n), c_i64(n_frames), c_i64(n_frames)))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    check(lib.dpvo_edges_reduce(ptr(flow_mag), ptr(ii), ptr(jj), c_i64(n), c_i64(n_frames), c_i64(n_frames),
                                c_i32(max_num_edges), c_i32(nms), ptr(out), ptr(cnt), ptr(ws), c_i64(nbytes),
                                stream_of(flow_mag)), "patchgraph.reduce_edges")
    return out[:max_num_edges], cnt
