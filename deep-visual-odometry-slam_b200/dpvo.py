"""`DPVO` — the reference's tracking class (dpvo/dpvo.py:20-473 + dpvo/patchgraph.py) on the B200 operators.

Same constructor, `__call__(tstamp, image, intrinsics)` and `terminate()` as the reference class, the same state
machine (8-frame initialisation with 12 updates, motion probe, damped-linear motion model, keyframe removal,
removal window, proximity loop closure, global bundle adjustment), the same buffer names (`pg.poses_`,
`pg.patches_`, `pg.intrinsics_`, `pg.tstamps_`, `imap_`, `gmap_`, `fmap1_`, `fmap2_`) and the same random-number
consumption (patch centres, depth initialisation), so a run is comparable with the reference class frame by frame:
tests/test_refdpvo_gpu.py asserts bit-identical ii / jj / kk against the reference class on the reference extensions.

What differs is the buffer layout and who does the work (the north star's "DPVO update loop's buffer layout"):
  * feature rings are channels-last (the lookup reads them directly; no NCHW shadow), written from the encoders'
    channels-last output with one copy;
  * the edge list lives in fixed-capacity device arrays (patchgraph.PatchGraphEdges): append / remove / renumber /
    window culling are kernels, not torch.cat + boolean-mask indexing; the host reads the live count once per frame;
  * reprojection is one fused kernel (projective_ops.transform), the 2-level lookup one launch (altcorr.corr_pyramid),
    the Update operator runs on tcgen05 (update.Update), BA is cuda_ba.forward, the encoders are encoder.Patchifier;
  * the frame-slot shifts of keyframe() are `frames_shift` kernels.
The host keeps what the reference keeps on the host: n, m, the time stamps, the keyframe decision (one scalar read)
and the loop-closure edge selection count.  CUDA only.
"""
import numpy as np
import torch

from . import altcorr, fastba, lietorch_backends as lie, patchgraph, projective_ops as pops
from .encoder import Patchifier
from .update import Update

SE3 = 3


class _Graph:
    """`PatchGraph` fields (patchgraph.py:14-54) + the device edge store; ii/jj/kk & co. are views of the live part."""

    def __init__(self, cfg, P, DIM, pmem, device, capacity, inactive_capacity):
        self.cfg, self.P, self.DIM, self.pmem = cfg, P, DIM, pmem
        self.n = 0
        self.m = 0
        self.M, self.N = cfg.PATCHES_PER_FRAME, cfg.BUFFER_SIZE
        dev = device
        self.tstamps_ = np.zeros(self.N, dtype=np.int64)
        self.poses_ = torch.zeros(self.N, 7, dtype=torch.float, device=dev)
        self.patches_ = torch.zeros(self.N, self.M, 3, P, P, dtype=torch.float, device=dev)
        self.intrinsics_ = torch.zeros(self.N, 4, dtype=torch.float, device=dev)
        self.points_ = torch.zeros(self.N * self.M, 3, dtype=torch.float, device=dev)
        self.colors_ = torch.zeros(self.N, self.M, 3, dtype=torch.uint8, device=dev)
        # index_[f] = f for every frame slot (the reference fills row n+1 when frame n arrives, dpvo.py:407)
        self.index_ = torch.arange(self.N, device=dev).view(-1, 1).repeat(1, self.M).contiguous()
        self.index_map_ = torch.zeros(self.N, dtype=torch.long, device=dev)
        self.poses_[:, 6] = 1.0
        self.delta = {}
        self.edges = patchgraph.PatchGraphEdges(self.index_.view(-1), self.M, capacity, inactive_capacity=inactive_capacity,
                                                net_dim=DIM)
        self.E = 0            # live active edges (host copy, refreshed once per frame)
        self.E_inac = 0
        self.min_ii = self.BIG  # smallest source frame among the live edges (host copy, same refresh)

    # ---- views of the live edges (host-known counts: no synchronisation)
    @property
    def ii(self): return self.edges.active.ii_[:self.E]
    @property
    def jj(self): return self.edges.active.jj_[:self.E]
    @property
    def kk(self): return self.edges.active.kk_[:self.E]
    @property
    def net(self): return self.edges.active.net_[:self.E][None]
    @property
    def target(self): return self.edges.active.target_[:self.E][None]
    @property
    def weight(self): return self.edges.active.weight_[:self.E][None]
    @property
    def ii_inac(self): return self.edges.inactive.ii_[:self.E_inac]
    @property
    def jj_inac(self): return self.edges.inactive.jj_[:self.E_inac]
    @property
    def kk_inac(self): return self.edges.inactive.kk_[:self.E_inac]
    @property
    def target_inac(self): return self.edges.inactive.target_[:self.E_inac][None]
    @property
    def weight_inac(self): return self.edges.inactive.weight_[:self.E_inac][None]

    @property
    def poses(self): return self.poses_.view(1, self.N, 7)
    @property
    def patches(self): return self.patches_.view(1, self.N * self.M, 3, 3, 3)
    @property
    def intrinsics(self): return self.intrinsics_.view(1, self.N, 4)
    @property
    def ix(self): return self.index_.view(-1)

    BIG = 1 << 40

    def sync_counts(self):
        """The one host read per frame: live counts of both lists and the smallest source frame of a live edge (what
        `update()` needs for its "are there long-range edges" test, dpvo.py:348 — read here instead of with a second
        synchronisation inside the next update)."""
        a = self.edges.active
        b = min(max(a.bound, 1), a.capacity)
        live = torch.arange(b, device=a.ii_.device) < a.count_
        mn = torch.where(live, a.ii_[:b], torch.full_like(a.ii_[:b], self.BIG)).min().view(1)
        both = torch.cat([a.count_, self.edges.inactive.count_, mn]).cpu()
        self.E, self.E_inac, self.min_ii = int(both[0]), int(both[1]), int(both[2])
        self.edges.active.bound, self.edges.inactive.bound = self.E, self.E_inac
        return self.E


class DPVO:
    def __init__(self, cfg, network, ht=480, wd=640, viz=False, edge_capacity=393216, inactive_capacity=8_000_000):
        if viz:
            raise RuntimeError("DPVO (dpvo_b200): the Pangolin / OpenGL viewer is unavailable on this box; use viewer_ops")
        self.cfg = cfg
        self.load_weights(network)
        self.is_initialized = False
        self.enable_timing = False
        self.M = cfg.PATCHES_PER_FRAME
        self.N = cfg.BUFFER_SIZE
        self.ht, self.wd = ht, wd
        DIM, RES = self.DIM, self.RES
        self.tlist = []
        self.counter = 0
        self.ran_global_ba = np.zeros(100000, dtype=bool)
        h, w = ht // RES, wd // RES
        dev = self.device
        self.dtype = torch.half                                   # the B200 operators are fp16 (cfg.MIXED_PRECISION path)
        self.kwargs = {"device": dev, "dtype": torch.half}
        self.pmem = self.mem = 36
        if cfg.LOOP_CLOSURE:
            self.last_global_ba = -1000
            self.pmem = cfg.MAX_EDGE_AGE
        if getattr(cfg, "CLASSIC_LOOP_CLOSURE", False):
            raise RuntimeError("DPVO (dpvo_b200): CLASSIC_LOOP_CLOSURE (DBoW2 retrieval) is outside this build")
        self.imap_ = torch.zeros(self.pmem, self.M, DIM, **self.kwargs)
        self.gmap_ = torch.zeros(self.pmem, self.M, 128, self.P, self.P, **self.kwargs)
        self.pg = _Graph(cfg, self.P, DIM, self.pmem, dev, edge_capacity, inactive_capacity)
        # channels-last rings behind the reference's [1, mem, 128, h, w] shape
        self._fmap1 = torch.zeros(1, self.mem, h, w, 128, **self.kwargs)
        self._fmap2 = torch.zeros(1, self.mem, h // 4, w // 4, 128, **self.kwargs)
        self.fmap1_ = self._fmap1.permute(0, 1, 4, 2, 3)
        self.fmap2_ = self._fmap2.permute(0, 1, 4, 2, 3)
        self.pyramid = (self.fmap1_, self.fmap2_)
        self.viewer = None
        self._lmbda = torch.as_tensor([1e-4], device=dev)
        self._plan_out = torch.zeros(8, dtype=torch.int32, device=dev)      # keyframe(): decision + edge counts, one read
        self.verify_plan = False                                            # tests: compare the plan with the device counts
        # relative poses of removed frames (`pg.delta`, dpvo.py:274-277): rows of one device tensor, so that normalize()
        # rescales all of them with one kernel instead of one per removed frame
        self._delta_poses = torch.zeros(4096, 7, device=dev)
        self._delta_n = 0
        self._corr_buf = self._ctx_buf = None

    def _add_delta(self, t, t0, dP):
        if self._delta_n == self._delta_poses.shape[0]:
            self._delta_poses = torch.cat([self._delta_poses, torch.zeros_like(self._delta_poses)], 0)
        self._delta_poses[self._delta_n] = dP.reshape(7)
        self.pg.delta[t] = (t0, self._delta_n)
        self._delta_n += 1

    def load_weights(self, network):
        """`network`: the reference's VONet (or any object with `.patchify` / `.update` modules carrying the reference
        state-dict keys) or a checkpoint path (dpvo.py:90-112).  Its two sub-modules are re-bound to the B200 operators."""
        if isinstance(network, str):
            raise RuntimeError("DPVO (dpvo_b200): pass a constructed network (checkpoint loading needs the reference's VONet)")
        self.network = network
        self.DIM, self.RES, self.P = 384, 4, 3
        if hasattr(network, "cuda"):
            network.cuda()
        if hasattr(network, "eval"):
            network.eval()
        if not isinstance(network.update, Update):
            network.update = Update.from_reference(network.update)
        # The operator's operands live in persistent buffers here, so its opt-in CUDA-graph cache (update.Update.use_graphs)
        # COULD be used; it is not: keyframe removal makes the edge count differ from frame to frame (measured on the
        # 1,000-frame sequence: a new signature almost every frame, 146 -> 93 frames/s from the captures alone).
        if not isinstance(network.patchify, Patchifier):
            network.patchify = Patchifier.from_reference(network.patchify)
        self.device = network.update.device

    # ------------------------------------------------------------------ reference-named accessors
    @property
    def poses(self): return self.pg.poses
    @property
    def patches(self): return self.pg.patches
    @property
    def intrinsics(self): return self.pg.intrinsics
    @property
    def ix(self): return self.pg.ix
    @property
    def imap(self): return self.imap_.view(1, self.pmem * self.M, self.DIM)
    @property
    def gmap(self): return self.gmap_.view(1, self.pmem * self.M, 128, 3, 3)
    @property
    def n(self): return self.pg.n
    @n.setter
    def n(self, v): self.pg.n = v
    @property
    def m(self): return self.pg.m
    @m.setter
    def m(self, v): self.pg.m = v

    # ------------------------------------------------------------------ SE3 helpers on raw [k,7] tensors
    @staticmethod
    def _mul(a, b): return lie.mul(SE3, a.contiguous(), b.contiguous())
    @staticmethod
    def _inv(a): return lie.inv(SE3, a.contiguous())

    # ------------------------------------------------------------------ update pieces
    def corr(self, coords, indicies=None):
        ii, jj = indicies if indicies is not None else (self.pg.kk, self.pg.jj)
        ii1 = ii % (self.M * self.pmem)
        jj1 = jj % self.mem
        E = ii.numel()
        out = None
        if indicies is None:
            # the live edge list's lookup lands in a persistent buffer: the Update operator then sees the same operand
            # pointers frame after frame and replays its CUDA graph (update.Update._graph_lookup)
            if self._corr_buf is None or self._corr_buf.shape[1] < E:
                self._corr_buf = torch.empty((1, max(int(E * 1.25), 65536), 896), **self.kwargs)
            out = self._corr_buf
        return altcorr.corr_pyramid(self.gmap, self.pyramid, coords, ii1, jj1, 3, row_stride=896, out=out)

    def _ctx_index(self):
        """kk % (M * pmem) of dpvo.py:333 into a persistent buffer (stable pointer, see corr())."""
        E = self.pg.E
        if self._ctx_buf is None or self._ctx_buf.numel() < E:
            self._ctx_buf = torch.empty(max(int(E * 1.25), 65536), dtype=torch.int64, device=self.device)
        out = self._ctx_buf[:E]
        torch.remainder(self.pg.kk, self.M * self.pmem, out=out)
        return out

    def reproject(self, indicies=None):
        ii, jj, kk = indicies if indicies is not None else (self.pg.ii, self.pg.jj, self.pg.kk)
        coords = pops.transform(self.poses, self.patches, self.intrinsics, ii, jj, kk)
        return coords.permute(0, 1, 4, 2, 3).contiguous()

    def motion_probe(self):
        kk = torch.arange(self.m - self.M, self.m, device=self.device)
        jj = self.n * torch.ones_like(kk)
        ii = self.ix[kk]
        net = torch.zeros(1, len(ii), self.DIM, **self.kwargs)
        coords = self.reproject(indicies=(ii, jj, kk))
        corr = self.corr(coords, indicies=(kk, jj))
        ctx = (self.imap[0], kk % (self.M * self.pmem))
        net, (delta, weight, _) = self.network.update(net, ctx, corr, None, ii, jj, kk, n_frames=self.n + 1, n_patches=self.m)
        return torch.quantile(delta.norm(dim=-1).float(), 0.5)

    def motionmag(self, i, j):
        """dpvo.py:257-264 for ONE direction (kept for callers that want it; keyframe() uses motionmag_pair)."""
        pg = self.pg
        k = torch.nonzero((pg.ii == i) & (pg.jj == j)).view(-1)
        flow, _ = pops.flow_mag(self.poses, self.patches, self.intrinsics, pg.ii[k], pg.jj[k], pg.kk[k], beta=0.5)
        return flow.mean()

    def motionmag_pair(self, i, j):
        """motionmag(i, j) + motionmag(j, i) without the two boolean-mask index synchronisations: ONE launch over the
        live edge list (`pops.motionmag_pair`: flow magnitude of the matching edges, the two means reduced in a fixed
        order); empty selections give NaN like the reference's mean over an empty tensor (=> no keyframe removal)."""
        pg = self.pg
        return pops.motionmag_pair(self.poses, self.patches, self.intrinsics, pg.ii, pg.jj, pg.kk, i, j, beta=0.5)

    def keyframe(self):
        """dpvo.py:266-310.  ONE host read per frame: `pops.keyframe_plan` returns the motion magnitude the reference
        reads at :270-272 together with the edge counts / smallest source frame that follow from either outcome (a dry
        run of the removal rules in the same pass over the edge list), so the edits below are issued without reading
        their result back.  `self.verify_plan = True` re-reads the device counts afterwards and compares (tests)."""
        cfg, pg = self.cfg, self.pg
        i = self.n - cfg.KEYFRAME_INDEX - 1
        j = self.n - cfg.KEYFRAME_INDEX + 1
        k = self.n - cfg.KEYFRAME_INDEX
        plan = pops.keyframe_plan(self.poses, self.patches, self.intrinsics, pg.ii, pg.jj, pg.kk, self.ix, i, j, k, self.n,
                                  self.M, cfg.REMOVAL_WINDOW, bool(cfg.LOOP_CLOSURE), cfg.OPTIMIZATION_WINDOW, beta=0.5,
                                  out=self._plan_out, live_count=pg.edges.active.count_, info=pg.edges.info)
        h = plan.cpu()                                                           # the decision is a host read (dpvo.py:270)
        m = float(h[:1].view(torch.float32)[0])
        moved_a, min_a, dropped_b, moved_b, min_b, dev_count, info = (int(v) for v in h[1:8].tolist())
        if info != 0 or dev_count != pg.E:
            # the host mirrors the edge counts instead of reading them back: fail loudly if the mirror is not true
            raise RuntimeError(f"patch graph out of sync: device holds {dev_count} live edges, host expects {pg.E}; "
                               f"edge-kernel flag {info} (1 = a capacity was exceeded, 2 = patch index outside ix)")
        if m / 2 < cfg.KEYFRAME_THRESH:
            t0, t1 = pg.tstamps_[k - 1], pg.tstamps_[k]
            self._add_delta(t1, t0, self._mul(pg.poses_[k:k + 1], self._inv(pg.poses_[k - 1:k])))
            pg.edges.remove_frame(k)
            n = self.n
            pg.tstamps_[k:n - 1] = pg.tstamps_[k + 1:n]
            shift = patchgraph.frames_shift
            shift(pg.colors_, k, n); shift(pg.poses_, k, n); shift(pg.patches_, k, n); shift(pg.intrinsics_, k, n)
            shift(self.imap_, k, n, mod=self.pmem); shift(self.gmap_, k, n, mod=self.pmem)
            shift(self._fmap1[0], k, n, mod=self.mem); shift(self._fmap2[0], k, n, mod=self.mem)
            self.n -= 1
            self.m -= self.M
            pg.edges.active.bound = pg.E - dropped_b
            E, moved, mn = pg.E - dropped_b - moved_b, moved_b, min_b
        else:
            E, moved, mn = pg.E - moved_a, moved_a, min_a
        pg.edges.remove_old(self.n, cfg.REMOVAL_WINDOW, loop_closure=bool(cfg.LOOP_CLOSURE),
                            optimization_window=cfg.OPTIMIZATION_WINDOW)
        planned = (E, pg.E_inac + moved, pg.BIG if mn == pops.PLAN_NONE else mn)
        if self.verify_plan:
            pg.sync_counts()
            if (pg.E, pg.E_inac, pg.min_ii) != planned:
                raise RuntimeError(f"keyframe plan {planned} != device counts {(pg.E, pg.E_inac, pg.min_ii)}")
        pg.E, pg.E_inac, pg.min_ii = planned
        pg.edges.active.bound, pg.edges.inactive.bound = pg.E, pg.E_inac

    def edges_loop(self):
        """patchgraph.py:56-82."""
        cfg, pg, M = self.cfg, self.pg, self.M
        dev = self.device
        l = self.n - cfg.REMOVAL_WINDOW
        if l <= 0:
            return torch.empty(2, 0, dtype=torch.long, device=dev)
        fj = torch.arange(self.n - cfg.GLOBAL_OPT_FREQ, self.n - cfg.KEYFRAME_INDEX, device=dev)
        fk = torch.arange(max(l - cfg.MAX_EDGE_AGE, 0) * M, l * M, device=dev)
        jj, kk = (x.reshape(-1) for x in torch.meshgrid(fj, fk, indexing="ij"))
        ii = self.ix[kk]
        centres = self.patches[..., 1, 1].reshape(1, -1, 3, 1, 1)
        flow_mg, val = pops.flow_mag(self.poses, centres, self.intrinsics, ii, jj, kk, beta=0.5)
        nf = jj.numel() // M
        valf = val.view(nf, M).float()
        flow_sum = (flow_mg.view(nf, M) * valf).sum(1).float()
        num_val = valf.sum(1).clamp(min=1)
        flow_mag = torch.where(num_val > (M * 0.75), flow_sum / num_val, torch.full_like(flow_sum, float("inf")))
        mask = flow_mag < cfg.BACKEND_THRESH
        fi, fjj = ii[::M][mask], jj[::M][mask]
        if fi.numel() == 0:
            return torch.empty(2, 0, dtype=torch.long, device=dev)
        es, cnt = patchgraph.reduce_edges(flow_mag[mask], fi, fjj, max_num_edges=1000, nms=1, n_frames=self.n + 1)
        # the selection count is a host read (optim_utils.py returns host arrays); the smallest selected source frame
        # rides along, so that the host's copy of min(ii) stays exact and __run_global_BA needs no read of its own
        rows = torch.arange(es.shape[0], device=dev) < cnt
        lo = torch.where(rows, es[:, 0], torch.full_like(es[:, 0], pg.BIG)).min().view(1)
        c, lo = (int(v) for v in torch.cat([cnt.view(1), lo]).tolist())
        es = es[:c]
        if es.shape[0] == 0:
            return torch.empty(2, 0, dtype=torch.long, device=dev)
        self._lc_min_src = lo
        ei = es[:, 0:1].repeat(1, M)
        ej = es[:, 1:2].repeat(1, M)
        kk = ei * M + torch.arange(M, device=dev)
        return kk.flatten(), ej.flatten()

    def append_factors(self, ii, jj):
        """Loop-closure edges (patchgraph.py:56-82): jj - source frame >= 30 and jj <= n - KEYFRAME_INDEX - 1, so every one
        of them is a long-range edge in the sense of dpvo.py:348."""
        self.pg.edges.append_factors(ii, jj)
        self.pg.E += ii.numel()
        if ii.numel():
            self.pg.min_ii = min(self.pg.min_ii, self._lc_min_src)          # exact: read with the selection count

    def normalize(self):
        """patchgraph.py:84-96."""
        pg, n = self.pg, self.n
        s = pg.patches_[:n, :, 2].mean()
        pg.patches_[:n, :, 2] /= s
        pg.poses_[:n, :3] *= s
        self._delta_poses[:self._delta_n, :3] *= s
        inv0 = self._inv(pg.poses_[0:1]).expand(n, 7)
        pg.poses_[:n] = self._mul(pg.poses_[:n], inv0)
        self._points()

    def _points(self):
        m = self.m
        pts = pops.point_cloud(self.poses, self.patches[:, :m], self.intrinsics, self.ix[:m])
        self.pg.points_[:m] = (pts[0, :, 1, 1, :3] / pts[0, :, 1, 1, 3:])

    def _run_global_BA(self):
        pg = self.pg
        full_target = torch.cat((pg.target_inac, pg.target), dim=1)
        full_weight = torch.cat((pg.weight_inac, pg.weight), dim=1)
        full_ii = torch.cat((pg.ii_inac, pg.ii))
        full_jj = torch.cat((pg.jj_inac, pg.jj))
        full_kk = torch.cat((pg.kk_inac, pg.kk))
        self.normalize()
        t0 = pg.min_ii                   # = pg.ii.min() (dpvo.py:318): kept exact on the host by keyframe() and the appends
        if self.verify_plan and t0 != int(pg.ii.min().item()):
            raise RuntimeError(f"host copy of min(ii) {t0} != device {int(pg.ii.min().item())}")
        fastba.BA(self.poses, self.patches, self.intrinsics, full_target, full_weight, self._lmbda, full_ii, full_jj, full_kk,
                  t0, self.n, M=self.M, iterations=2, eff_impl=True)
        self.ran_global_ba[self.n] = True

    def update(self):
        pg, cfg = self.pg, self.cfg
        E = pg.E
        coords = self.reproject()
        corr = self.corr(coords)
        ctx = (self.imap[0], self._ctx_index())
        nf = (self.n + 64) // 64 * 64              # coarse bounds: the operator's graph signature changes every 64 frames
        net, (delta, weight, _) = self.network.update(pg.net, ctx, corr, None, pg.ii, pg.jj, pg.kk,
                                                      n_frames=nf, n_patches=nf * self.M)
        weight = weight.float()
        target = coords[..., self.P // 2, self.P // 2] + delta.float()
        act = pg.edges.active
        act.net_[:E].copy_(net[0])
        act.target_[:E].copy_(target[0])
        act.weight_[:E].copy_(weight[0])
        try:
            # (pg.ii < n - REMOVAL_WINDOW - 1).any() of dpvo.py:348 from the host copy of min(ii): refreshed by the
            # frame's one host read (keyframe -> sync_counts); edges appended since are either the new frame's
            # (source frame >= n - PATCH_LIFETIME) or loop-closure edges (accounted for in append_factors)
            far = bool(cfg.LOOP_CLOSURE) and pg.min_ii < self.n - cfg.REMOVAL_WINDOW - 1
            if far and not self.ran_global_ba[self.n]:
                self._run_global_BA()
            else:
                t0 = self.n - cfg.OPTIMIZATION_WINDOW if self.is_initialized else 1
                t0 = max(t0, 1)
                fastba.BA(self.poses, self.patches, self.intrinsics, pg.target, pg.weight, self._lmbda, pg.ii, pg.jj, pg.kk,
                          t0, self.n, M=self.M, iterations=2, eff_impl=False)
        except RuntimeError as ex:                                     # the reference swallows BA failures (dpvo.py:355)
            print("Warning BA failed...", str(ex)[:120])
        self._points()

    # ------------------------------------------------------------------ one frame
    def __call__(self, tstamp, image, intrinsics):
        cfg, pg, M = self.cfg, self.pg, self.M
        if (self.n + 1) >= self.N:
            raise Exception(f'The buffer size is too small. You can increase it using "--opts BUFFER_SIZE={self.N * 2}"')
        image = 2 * (image[None, None] / 255.0) - 0.5
        fmap, gmap, imap, patches, _, clr = self.network.patchify(image, patches_per_image=M,
                                                                  centroid_sel_strat=cfg.CENTROID_SEL_STRAT, return_color=True)
        pf = self.network.patchify
        n = self.n
        self.tlist.append(tstamp)
        pg.tstamps_[n] = self.counter
        pg.intrinsics_[n] = intrinsics / self.RES
        clr = (clr[0].flip(-1) + 0.5) * (255.0 / 2)
        pg.colors_[n] = clr.to(torch.uint8)
        pg.index_map_[n + 1] = self.m + M
        if n > 1:
            if cfg.MOTION_MODEL == "DAMPED_LINEAR":
                P1, P2 = pg.poses_[n - 1:n], pg.poses_[n - 2:n - 1]
                *_, a, b, c = [1] * 3 + self.tlist
                fac = (c - b) / (b - a)
                xi = cfg.MOTION_DAMPING * fac * lie.logm(SE3, self._mul(P1, self._inv(P2)))
                pg.poses_[n] = self._mul(lie.expm(SE3, xi.contiguous()), P1)[0]
            else:
                pg.poses_[n] = pg.poses_[n - 1]
        patches[:, :, 2] = torch.rand_like(patches[:, :, 2, 0, 0, None, None])
        if self.is_initialized:
            patches[:, :, 2] = torch.median(pg.patches_[n - 3:n, :, 2])
        pg.patches_[n] = patches
        self.imap_[n % self.pmem] = imap.squeeze()
        self.gmap_[n % self.pmem] = gmap.squeeze()
        self._fmap1[0, n % self.mem].copy_(pf.fmap_nhwc[0])
        self._fmap2[0, n % self.mem].copy_(pf.fmap2_nhwc[0])
        self.counter += 1
        if n > 0 and not self.is_initialized:
            if float(self.motion_probe()) < 2.0:
                self._add_delta(self.counter - 1, self.counter - 2, torch.tensor([0, 0, 0, 0, 0, 0, 1.0], device=self.device))
                return
        self.n += 1
        self.m += M
        if cfg.LOOP_CLOSURE:
            if self.n - self.last_global_ba >= cfg.GLOBAL_OPT_FREQ:
                lii, ljj = self.edges_loop()
                if lii.numel() > 0:
                    self.last_global_ba = self.n
                    self.append_factors(lii, ljj)
        r = cfg.PATCH_LIFETIME
        pg.edges.edges_forw(self.n, r)
        pg.edges.edges_back(self.n, r)
        added = M * (max(self.n - 1, 0) - max(self.n - r, 0)) + M * (self.n - max(self.n - r, 0))
        pg.E += added
        if added:
            pg.min_ii = min(pg.min_ii, max(self.n - r, 0))                    # sources of the new edges: frames >= n - r
        pg.edges.active.bound = pg.E
        if self.n == 8 and not self.is_initialized:
            self.is_initialized = True
            for _ in range(12):
                self.update()
        elif self.is_initialized:
            self.update()
            self.keyframe()

    # ------------------------------------------------------------------ end of sequence
    def get_pose(self, t):
        if t in self.traj:
            return self.traj[t]
        t0, row = self.pg.delta[t]
        self.traj[t] = self._mul(self._delta_host[row:row + 1], self.get_pose(t0))
        return self.traj[t]

    def terminate(self):
        if self.cfg.LOOP_CLOSURE:
            lii, ljj = self.edges_loop()
            if lii.numel() > 0:
                self.append_factors(lii, ljj)
        for _ in range(12):
            self.ran_global_ba[self.n] = False
            self.update()
        # interpolate the poses of the removed frames (dpvo.py:186-191); tiny chains: host form of the SE3 product
        import sys
        sys.setrecursionlimit(max(sys.getrecursionlimit(), self.counter + 100))
        poses_h = self.pg.poses_[:self.n].cpu()
        self._delta_host = self._delta_poses[:self._delta_n].cpu()
        self.traj = {int(self.pg.tstamps_[i]): poses_h[i:i + 1] for i in range(self.n)}
        poses = torch.cat([self.get_pose(t) for t in range(self.counter)], 0)
        poses = self._inv(poses).numpy()
        tstamps = np.array(self.tlist, dtype=np.float64)
        return poses, tstamps
