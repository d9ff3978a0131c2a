"""One frame of DPVO's update hot path, driven through this package's operator surface.

Mirrors what `DPVO.__call__` -> `DPVO.update` does per frame between the network calls
(/root/reference/methods/dpvo/dpvo/dpvo.py:328-360,435-438, net.py:76,135-145):

    ring write      new frame's fmaps into the feature rings           dpvo.py:437-438
    patchify        gmap (r=1, C=128) and imap (r=0, C=384) gathers    net.py:135-145
    reproject       coords of every active edge                        dpvo.py:209-213
    corr            2-level correlation lookup -> [1,E,882] fp16       dpvo.py:200-207
    neighbors       temporal neighbours of every edge                  net.py:76
    BA              2 Gauss-Newton iterations, 10-pose window          dpvo.py:351-354

`UpdateHotPath` is the north-star path: the Update network's outputs `target` and `weight` are inputs.
`UpdateIteration` adds the Update operator itself (update.py) between corr and BA, i.e.
the whole body of `DPVO.update` (dpvo.py:328-354); the encoders (N3) stay outside.
Used by bench.py and the tests; all tensors live on one CUDA device, all launches go to
torch's current stream, nothing synchronises with the host.
"""
import numpy as np
import torch

from . import altcorr, cuda_corr, fastba, synth, _lib


class UpdateHotPath:
    """Steady-state DPVO window (default config: 96 patches/frame, 45,312 active edges)."""

    def __init__(self, device, n_frames=40, seed=0, opt_window=synth.OPT_WINDOW, M=synth.M_DEFAULT):
        self.device = torch.device(device)
        self.n = n_frames
        self.M = M
        d = synth.ba_inputs(n_frames, seed=seed, M=M)
        c = synth.corr_inputs(0, seed=seed + 1, edges=(d["ii"], d["jj"], d["kk"]), M=M, oob_frac=0.0)
        dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(self.device)
        self.E = len(d["kk"])
        self.poses0 = dev(d["poses"])[None]
        self.patches0 = dev(d["patches"])[None]
        self.poses = self.poses0.clone()
        self.patches = self.patches0.clone()
        self.intrinsics = dev(d["intrinsics"])[None]
        self.target = dev(d["target"])[None]
        self.weight = dev(d["weight"])[None]
        self.ii, self.jj, self.kk = dev(d["ii"]), dev(d["jj"]), dev(d["kk"])
        self.lmbda = torch.tensor([1e-4], device=self.device)
        self.t0, self.t1 = max(n_frames - opt_window, 1), n_frames
        # feature rings, channels-last storage behind the reference's logical [1,mem,C,H,W] shape
        nhwc = lambda t: t.permute(0, 1, 3, 4, 2).contiguous().permute(0, 1, 4, 2, 3)
        self.fmap1 = nhwc(dev(c["fmap1"]))
        self.fmap2 = nhwc(dev(c["fmap2"]))
        self.gmap = dev(c["gmap"])                        # [1, mem*M, 128, 3, 3]
        self.mem = self.fmap1.shape[1]
        self.n_patches = self.gmap.shape[1]
        # per-frame inputs (what the encoders hand over): NCHW fmaps of the newest frame, an imap,
        # and the patch centres
        rng = np.random.default_rng(seed + 2)
        self.new_f1 = dev(rng.standard_normal((1, synth.C, synth.H1, synth.W1), dtype=np.float32).astype(np.float16))
        self.new_f2 = dev(rng.standard_normal((1, synth.C, synth.H2, synth.W2), dtype=np.float32).astype(np.float16))
        self.new_imap = dev(rng.standard_normal((1, 384, synth.H1, synth.W1), dtype=np.float32).astype(np.float16))
        xy = np.stack([rng.uniform(1, synth.W1 - 2, (1, M)), rng.uniform(1, synth.H1 - 2, (1, M))], -1)
        self.centres = dev(xy.astype(np.float32))
        self.host_coords = c["coords"]                    # for the CPU baseline / byte accounting
        self.host = dict(d=d, c=c)
        self.stage_events = None
        self.overlap_neighbors = True
        self._ba_ws = None
        self._side = None
        self._side2 = None
        self._side3 = None
        self._nb = None

    # ------------------------------------------------------------------ stages
    def ring_write(self):
        slot = (self.n - 1) % self.mem
        for src, ring in ((self.new_f1, self.fmap1), (self.new_f2, self.fmap2)):
            C, H, W = src.shape[1:]
            dst = ring[0, slot]                            # logical [C,H,W], memory [H,W,C]
            _lib.check(_lib.lib.dpvo_nchw_to_nhwc(_lib.c_i32(_lib.dtype_code(src)), _lib.ptr(src),
                                                  _lib.c_p(dst.data_ptr()), _lib.c_i64(1), _lib.c_i32(C),
                                                  _lib.c_i64(H * W), _lib.stream_of(src)), "ring_write")

    def patchify(self):
        g, = cuda_corr.patchify_forward(self.new_f1, self.centres, 1)
        i, = cuda_corr.patchify_forward(self.new_imap, self.centres, 0)
        return g, i

    def reproject(self):
        return fastba.reproject(self.poses, self.patches, self.intrinsics, self.ii, self.jj, self.kk)

    def corr(self, coords):
        us = self.kk % self.n_patches
        vs = self.jj % self.mem
        return altcorr.corr_pyramid(self.gmap, (self.fmap1, self.fmap2), coords, us, vs, 3)

    def neighbors(self):
        return fastba.neighbors(self.kk, self.jj)

    def ba(self, prepared=None):
        fastba.BA(self.poses, self.patches, self.intrinsics, self.target, self.weight, self.lmbda,
                  self.ii, self.jj, self.kk, self.t0, self.t1, self.M, 2, False, prepared=prepared)

    def ba_prepare(self):
        """BA's edge-list-only part (patch set, compaction, per-patch edge lists, cleared accumulators)."""
        self._ba_ws = fastba.BA_prepare(self.poses, self.patches, self.ii, self.jj, self.kk, self.t0, self.t1, self.M,
                                        workspace=self._ba_ws)
        return self._ba_ws

    def _front(self, main, fork):
        """Ring write, patchify and reproject only depend on the frame's inputs / the current state, not on
        each other: three short branches after `fork` (ring write stays on the main stream), joined on
        the main stream before the lookup.  Returns the reprojected coords."""
        if self._side2 is None:
            self._side2 = torch.cuda.Stream(self.device)
            self._side3 = torch.cuda.Stream(self.device)
        j2, j3 = torch.cuda.Event(), torch.cuda.Event()
        self._side2.wait_event(fork)
        self._side3.wait_event(fork)
        with torch.cuda.stream(self._side2):
            self._patches_new = self.patchify()
            j2.record(self._side2)
        with torch.cuda.stream(self._side3):
            coords = self.reproject()
            coords.record_stream(main)
            j3.record(self._side3)
        self.ring_write()
        main.wait_event(j2)
        main.wait_event(j3)
        return coords

    # ------------------------------------------------------------------ one frame
    def step(self, events=None):
        """All hot-path operators of one frame.  `events`: optional dict name -> list of
        (start, end) CUDA event pairs appended per stage (recorded on the current stream)."""
        self.poses.copy_(self.poses0)
        self.patches.copy_(self.patches0)

        def timed(name, fn):
            if events is None:
                return fn()
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn()
            e1.record()
            events.setdefault(name, []).append((e0, e1))
            return r

        if events is None and self.overlap_neighbors:
            # `neighbors` only depends on the edge list (net.py:76): fork it onto a side stream so its six
            # small launches run under the correlation kernel; join before BA.  Inside a CUDA-graph
            # capture the fork / join become graph edges.
            main = torch.cuda.current_stream(self.device)
            if self._side is None:
                self._side = torch.cuda.Stream(self.device)
            fork, join = torch.cuda.Event(), torch.cuda.Event()
            fork.record(main)
            self._side.wait_event(fork)
            with torch.cuda.stream(self._side):
                self._nb = self.neighbors()
                ws = self.ba_prepare()                   # also edge-list-only: rides under the lookup too
                join.record(self._side)
            coords = self._front(main, fork)
            corr = self.corr(coords)
            main.wait_event(join)
            self.ba(prepared=ws)
            return corr
        timed("ring_write", self.ring_write)
        timed("patchify", self.patchify)
        coords = timed("reproject", self.reproject)
        corr = timed("corr", lambda: self.corr(coords))
        timed("neighbors", self.neighbors)
        timed("ba", self.ba)
        return corr

    # ------------------------------------------------------------------ CUDA-graph replay
    def capture(self):
        """Capture one frame (all ~24 launches of step()) into a CUDA graph; `replay()` then issues
        the frame with a single launch.  Every buffer the frame touches is persistent (state,
        rings, per-frame inputs, the BA / neighbours workspaces sized during the warm-up call), the
        outputs of the captured frame live in the graph's private pool."""
        self.step()                                   # sizes the workspaces outside the capture
        torch.cuda.synchronize(self.device)
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._graph_out = self.step()
        return self._graph

    def replay(self):
        self._graph.replay()
        return self._graph_out

    # ------------------------------------------------------------------ host-buffer variant
    _PER_FRAME = ("new_f1", "new_f2", "new_imap", "centres", "target", "weight")

    def make_host_inputs(self):
        """One pinned host buffer holding the per-frame inputs (encoder outputs + network outputs)
        back to back, two device staging buffers of the same layout (double buffering) and the
        copy stream of the streaming pipeline.  One frame = ONE host->device copy."""
        offs, total = {}, 0
        for k in self._PER_FRAME:
            t = getattr(self, k)
            offs[k] = total
            total += (t.numel() * t.element_size() + 255) & ~255
        view = lambda buf, k: buf[offs[k]:offs[k] + getattr(self, k).numel() * getattr(self, k).element_size()] \
            .view(getattr(self, k).dtype).view(getattr(self, k).shape)
        host = torch.empty(total, dtype=torch.uint8).pin_memory()
        for k in self._PER_FRAME:
            view(host, k).copy_(getattr(self, k).cpu())
        self._stage = [torch.empty(total, dtype=torch.uint8, device=self.device) for _ in range(2)]
        self._sets = [{k: view(buf, k) for k in self._PER_FRAME} for buf in self._stage]
        self._h2d_bytes = sum(getattr(self, k).numel() * getattr(self, k).element_size() for k in self._PER_FRAME)
        self._copy_stream = torch.cuda.Stream(self.device)
        self._copied = [torch.cuda.Event() for _ in range(2)]
        self._consumed = [torch.cuda.Event() for _ in range(2)]
        self._frame = 0
        return host

    def step_host(self, hin, hout):
        """Same frame with host<->device copies inside, as a streaming front end would run it: frame
        k's inputs go from pinned host memory into staging buffer k%2 on a copy stream (overlapping
        frame k-1's kernels on the compute stream), the compute stream waits for that copy, runs the
        frame and reads the updated window poses and patch inverse depths back.  Every frame's bytes
        cross PCIe inside the caller's timed region.  Returns (h2d_bytes, d2h_bytes)."""
        b = self._frame & 1
        self._frame += 1
        compute = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(self._consumed[b])       # buffer b's previous frame is done with it
            self._stage[b].copy_(hin, non_blocking=True)
            self._copied[b].record(self._copy_stream)
        compute.wait_event(self._copied[b])
        if getattr(self, "_host_graphs", None):
            self._host_graphs[b].replay()
        else:
            for k in self._PER_FRAME:
                setattr(self, k, self._sets[b][k])
            self.step()
        self._consumed[b].record(compute)
        pw = self.poses[0, self.t0:self.t1]
        dz = self.patches[0, :, 2, 0, 0]
        hout["poses"].copy_(pw, non_blocking=True)
        hout["depths"].copy_(dz, non_blocking=True)
        d2h = pw.numel() * 4 + dz.numel() * 4
        return self._h2d_bytes, d2h

    def capture_host(self):
        """One CUDA graph per staging buffer for step_host()."""
        graphs = []
        for b in range(2):
            for k in self._PER_FRAME:
                setattr(self, k, self._sets[b][k])
            self.step()
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.step()
            graphs.append(g)
        self._host_graphs = graphs

    def make_host_outputs(self):
        return dict(poses=torch.empty((self.t1 - self.t0, 7), dtype=torch.float32).pin_memory(),
                    depths=torch.empty(self.patches.shape[1], dtype=torch.float32).pin_memory())

    def corr_algorithmic_bytes(self):
        """exact algorithmic bytes of one fused 2-level lookup over the edge list
        at the coords produced by reproject() on the initial state."""
        coords = self.reproject().cpu().numpy()
        return synth.corr_algorithmic_bytes(coords, [(synth.H1, synth.W1), (synth.H2, synth.W2)], [1.0, 0.25])


class UpdateIteration(UpdateHotPath):
    """The whole body of `DPVO.update` (dpvo.py:328-354) for one frame:

        reproject -> corr -> Update operator -> target = centre + delta, weight -> BA

    with the hidden state `net` carried from call to call like `self.pg.net`.  Random-initialised
    operator parameters (synth.update_state_dict) — there is no checkpoint on the box."""

    def __init__(self, device, n_frames=40, seed=0, **kw):
        super().__init__(device, n_frames=n_frames, seed=seed, **kw)
        from . import update as _update
        sd = {k: torch.from_numpy(v) for k, v in synth.update_state_dict(seed + 3).items()}
        self.update_op = _update.Update.from_state_dict(sd, device=self.device)
        rng = np.random.default_rng(seed + 4)
        half = lambda *shape: torch.from_numpy(rng.standard_normal(shape, dtype=np.float32)).to(self.device).half()
        self.net = 0.5 * half(1, self.E, 384)
        self.imap = half(self.n_patches, 384)              # context ring, dpvo.py:63 ([pmem*M, DIM] flattened)
        self.ctx_index = (self.kk % self.n_patches).contiguous()
        self.n_patch_ids = int(self.kk.max().item()) + 1

    def step_full(self):
        self.poses.copy_(self.poses0)
        self.patches.copy_(self.patches0)
        main = torch.cuda.current_stream(self.device)
        fork, join = torch.cuda.Event(), torch.cuda.Event()
        fork.record(main)
        if self._side is None:
            self._side = torch.cuda.Stream(self.device)
        self._side.wait_event(fork)
        with torch.cuda.stream(self._side):
            ws = self.ba_prepare()
            join.record(self._side)
        coords = self._front(main, fork)
        us = self.kk % self.n_patches
        vs = self.jj % self.mem
        corr = altcorr.corr_pyramid(self.gmap, (self.fmap1, self.fmap2), coords, us, vs, 3, row_stride=896)
        if torch.cuda.is_current_stream_capturing():
            # the hidden state read by a captured frame is baked into the graph as a pointer, and the attribute is
            # rebound to the graph's output just below: keep the input alive, or the caching allocator may hand its
            # block out again — or unmap it at the next capture's empty_cache() — under the graph's feet
            self._captured_inputs = getattr(self, "_captured_inputs", []) + [self.net]
        self.net, (delta, weight, _) = self.update_op(self.net, (self.imap, self.ctx_index), corr, None,
                                                      self.ii, self.jj, self.kk, n_frames=self.n,
                                                      n_patches=self.n_patch_ids)
        P = coords.shape[-1]
        target = coords[..., P // 2, P // 2] + delta       # dpvo.py:341
        main.wait_event(join)
        fastba.BA(self.poses, self.patches, self.intrinsics, target.contiguous(), weight, self.lmbda,
                  self.ii, self.jj, self.kk, self.t0, self.t1, self.M, 2, False, prepared=ws)
        return delta, weight


class ImageFrontHotPath(UpdateHotPath):
    """The hot-path frame fed from the IMAGE (SURVEY.md §8 row N3 in front of the north-star path): `Patchifier`'s
    two encoders run on the device, their channels-last outputs are copied straight into the feature-ring slot
    (dpvo.py:437-438: no transpose, level 2 = the fused 4x4 average pool) and the patch gathers read the strided views.
    Same state, edge list and BA as UpdateHotPath; only the per-frame inputs come from the encoders instead of being
    given (random-initialised encoder weights: synthetic, like everything else in the bench)."""

    def __init__(self, device, n_frames=40, seed=0):
        super().__init__(device, n_frames=n_frames, seed=seed)
        from . import encoder
        mk = lambda dim, norm, s: encoder.BasicEncoder4(
            {k: torch.from_numpy(v) for k, v in synth.encoder_state_dict(dim, s).items()}, norm, device=device,
            output_scale=0.25)
        self._enc = encoder
        self.fnet, self.inet = mk(128, "instance", 100 + seed), mk(384, "none", 200 + seed)
        self.image = torch.from_numpy(synth.encoder_images(1, 4 * synth.H1, 4 * synth.W1, seed)).to(device)

    def encode(self):
        """The two encoders are independent and each layer fills a fraction of the GPU (150-600 CTAs of 4 warps):
        inet runs on a side stream beside fnet (two branches of the frame graph)."""
        main = torch.cuda.current_stream(self.device)
        if getattr(self, "_enc_side", None) is None:
            self._enc_side = torch.cuda.Stream(self.device)
        fork, join = torch.cuda.Event(), torch.cuda.Event()
        fork.record(main)
        self._enc_side.wait_event(fork)
        with torch.cuda.stream(self._enc_side):
            self._i = self.inet(self.image)
            self._i.nhwc.record_stream(main)
            join.record(self._enc_side)
        self._f = self.fnet(self.image)
        self._f2 = self._enc.avg_pool4(self._f.nhwc)
        main.wait_event(join)

    def ring_write(self):
        slot = (self.n - 1) % self.mem
        self.fmap1[0, slot].permute(1, 2, 0).copy_(self._f.nhwc[0])       # ring memory is [H, W, C]
        self.fmap2[0, slot].permute(1, 2, 0).copy_(self._f2[0])

    def patchify(self):
        g, = cuda_corr.patchify_forward(self._f.nchw5[0], self.centres, 1)
        i, = cuda_corr.patchify_forward(self._i.nchw5[0], self.centres, 0)
        return g, i

    def step(self, events=None):
        self.encode()
        return super().step(events)


class FullFrame(ImageFrontHotPath, UpdateIteration):
    """One steady-state frame of `DPVO.__call__` (dpvo.py:377-473) from the uint8 IMAGE to the updated window, every
    stage on the device and the whole frame capturable as ONE CUDA graph:

        image normalisation 2*(u8/255)-0.5                                    dpvo.py:389
        Patchifier: fnet + inet, imap / gmap / grid / colour gathers          net.py:106-153
        depth init = median of the last three frames, motion-model pose       dpvo.py:405-432
        ring / imap / gmap slot writes                                        dpvo.py:435-438
        append_factors(edges_forw), append_factors(edges_back)                dpvo.py:449-450
        update(): reproject, corr, Update operator, BA, point cloud           dpvo.py:328-360
        keyframe(): motion magnitude, remove frame n-4 + renumber + slot
                    shifts, move edges that left the window to `inactive`     dpvo.py:262-310

    Steady state: the reference drops keyframe candidate n-4 and lowers n again, so every frame runs at the same n and
    the same edge count; that is the frame replayed here (the decision `motionmag < KEYFRAME_THRESH` is a host read
    in the reference; here the magnitude is computed and the removal always taken, the more expensive branch).
    The Update operator, correlation and BA consume the fixed 45,312-edge window of UpdateIteration (their kernels
    take E from the host); the edge edits run on a device edge list of the same size and rules (patchgraph.py) whose
    count lives on the device.  Random-initialised network weights, synthetic image: a throughput composition, the
    parity of each stage is pinned by its own tests."""

    PATCH_LIFETIME, REMOVAL_WINDOW, KEYFRAME_INDEX = 13, 22, 4

    def __init__(self, device, n_frames=40, seed=0, inactive_capacity=6_000_000):
        super().__init__(device, n_frames=n_frames, seed=seed)
        from . import patchgraph, projective_ops, lietorch_backends
        self._pg, self._pops, self._lie = patchgraph, projective_ops, lietorch_backends
        dev, M, n = self.device, self.M, self.n
        self.image_u8 = ((self.image + 0.5) * 127.5).round().clamp(0, 255).to(torch.uint8)
        # integer patch centres like the reference's randint (net.py:113-115): the bilinear blend of altcorr.patchify is
        # then the identity on the leading (2r+1)^2 window of the raw (2r+2)^2 gather, taken here as a slice
        self.centres = self.centres.floor().contiguous()
        h, w = synth.H1, synth.W1
        ys, xs = torch.meshgrid(torch.arange(h, device=dev, dtype=torch.float32),
                                torch.arange(w, device=dev, dtype=torch.float32), indexing="ij")
        self._grid = torch.stack([xs, ys, torch.ones_like(xs)], 0)[None].contiguous()       # utils.py:39-50, disps = 1
        self._patch_slot = torch.zeros(1, M, 3, 3, 3, device=dev)
        self._pose_slot = torch.zeros(1, 7, device=dev)
        self._colors = torch.zeros(M, 3, dtype=torch.uint8, device=dev)
        self._points = torch.zeros(self.patches.shape[1], 3, device=dev)
        n_buf = self.poses.shape[1]
        self._ix_state = torch.arange(n_buf, device=dev).repeat_interleave(M).contiguous()
        # device edge list in its steady state
        self._ix = torch.arange(4096, device=dev).repeat_interleave(M).contiguous()
        e = patchgraph.PatchGraphEdges(self._ix, M, 60000, inactive_capacity=inactive_capacity, net_dim=384)
        k = 0
        for _ in range(30):
            k += 1
            e.edges_forw(k, self.PATCH_LIFETIME); e.edges_back(k, self.PATCH_LIFETIME); e.remove_old(k, self.REMOVAL_WINDOW)
        self.edges, self._n_edges = e, k
        for _ in range(3):
            self._edges_append()
            self._edges_keyframe()
        self.steady_edges = e.active.count()
        e.inactive.count()
        e.check_info()
        # motionmag(i, j) + motionmag(j, i) edge subsets of the fixed window (dpvo.py:257-264, 266-272)
        hi, hj = self.host["d"]["ii"], self.host["d"]["jj"]
        i, j = n - self.KEYFRAME_INDEX - 1, n - self.KEYFRAME_INDEX + 1
        sel = np.nonzero(((hi == i) & (hj == j)) | ((hi == j) & (hj == i)))[0]
        sel = torch.from_numpy(sel).to(dev)
        self._mm = (self.ii[sel].contiguous(), self.jj[sel].contiguous(), self.kk[sel].contiguous())
        self._edge_side = None

    # ---- stages -----------------------------------------------------------------------------------
    def normalise(self):
        self.image = (2.0 * (self.image_u8.float() / 255.0) - 0.5).contiguous()              # dpvo.py:389

    def patchify(self):
        g, i = super().patchify()
        slot = (self.n - 1) % self.mem
        self.gmap[0, slot * self.M:(slot + 1) * self.M].copy_(g[0, :, :, :3, :3])                          # dpvo.py:436
        self.imap[slot * self.M:(slot + 1) * self.M].copy_(i[0, :, :, 0, 0])                              # dpvo.py:435
        p, = cuda_corr.patchify_forward(self._grid, self.centres, 1)                                       # net.py:146-150
        c, = cuda_corr.patchify_forward(self.image[0], 4 * (self.centres + 0.5), 0)                        # net.py:132-134
        self._colors.copy_(((c[0, :, :, 0, 0].flip(-1) + 0.5) * (255.0 / 2)).to(torch.uint8))        # dpvo.py:425-427
        n = self.n
        depth = torch.median(self.patches0.view(-1, self.M, 3, 3, 3)[n - 3:n, :, 2])                       # dpvo.py:430-432
        self._patch_slot.copy_(p[:, :, :, :3, :3])
        self._patch_slot[:, :, 2] = depth
        return g, i

    def motion_model(self):
        """dpvo.py:412-421 (MOTION_MODEL 'DAMPED_LINEAR'): pose of the new frame from the last two."""
        P1, P2 = self.poses0[0, self.n - 1:self.n].contiguous(), self.poses0[0, self.n - 2:self.n - 1].contiguous()
        L = self._lie
        xi = 0.5 * L.logm(3, L.mul(3, P1, L.inv(3, P2)))
        self._pose_slot.copy_(L.mul(3, L.expm(3, xi), P1))

    def _edges_append(self):
        n = self._n_edges + 1
        self.edges.edges_forw(n, self.PATCH_LIFETIME)
        self.edges.edges_back(n, self.PATCH_LIFETIME)

    def _edges_keyframe(self):
        n = self._n_edges + 1
        self.edges.remove_frame(n - self.KEYFRAME_INDEX)
        self.edges.remove_old(n - 1, self.REMOVAL_WINDOW)

    def keyframe(self):
        ii, jj, kk = self._mm
        flow, _ = self._pops.flow_mag(self.poses, self.patches, self.intrinsics, ii, jj, kk, beta=0.5)
        self.motion = flow.mean()                                     # device scalar; the reference reads it (.item())
        # the slot shifts touch the frame buffers, the edge edits the edge list: two branches
        main = torch.cuda.current_stream(self.device)
        if getattr(self, "_shift_side", None) is None:
            self._shift_side = torch.cuda.Stream(self.device)
        fork, join = torch.cuda.Event(), torch.cuda.Event()
        fork.record(main)
        self._shift_side.wait_event(fork)
        with torch.cuda.stream(self._shift_side):
            k, n, M = self.n - self.KEYFRAME_INDEX, self.n, self.M
            shift = self._pg.frames_shift
            shift(self.fmap1.permute(0, 1, 3, 4, 2)[0], k, n, mod=self.mem)
            shift(self.fmap2.permute(0, 1, 3, 4, 2)[0], k, n, mod=self.mem)
            shift(self.gmap[0].view(self.mem, -1), k, n, mod=self.mem)
            shift(self.imap.view(self.mem, -1), k, n, mod=self.mem)
            shift(self.poses[0], k, n)
            shift(self.patches[0].view(-1, M * 27), k, n)
            shift(self.intrinsics[0], k, n)
            join.record(self._shift_side)
        self._edges_keyframe()
        main.wait_event(join)

    def point_cloud(self):
        """dpvo.py:356-358."""
        m = self.n * self.M
        pts = self._pops.point_cloud(self.poses, self.patches[:, :m], self.intrinsics, self._ix_state[:m])
        self._points[:m] = (pts[0, :, 1, 1, :3] / pts[0, :, 1, 1, 3:])

    # ---- one frame --------------------------------------------------------------------------------
    def step_frame(self):
        main = torch.cuda.current_stream(self.device)
        if self._edge_side is None:
            self._edge_side = torch.cuda.Stream(self.device)
        fork, join = torch.cuda.Event(), torch.cuda.Event()
        fork.record(main)
        self._edge_side.wait_event(fork)
        with torch.cuda.stream(self._edge_side):              # edge appends and the motion model only need the state
            self._edges_append()
            self.motion_model()
            join.record(self._edge_side)
        self.normalise()
        self.encode()
        out = self.step_full()
        self.point_cloud()
        main.wait_event(join)
        self.keyframe()
        return out

    def _rest(self):
        """step_frame() without normalise + encode (they only depend on the image: see make_pipeline)."""
        main = torch.cuda.current_stream(self.device)
        if self._edge_side is None:
            self._edge_side = torch.cuda.Stream(self.device)
        fork, join = torch.cuda.Event(), torch.cuda.Event()
        fork.record(main)
        self._edge_side.wait_event(fork)
        with torch.cuda.stream(self._edge_side):
            self._edges_append()
            self.motion_model()
            join.record(self._edge_side)
        out = self.step_full()
        self.point_cloud()
        main.wait_event(join)
        self.keyframe()
        return out

    def capture_frame(self):
        for _ in range(2):
            self.step_frame()
        torch.cuda.synchronize(self.device)
        self.edges.active.count()
        self.edges.check_info()
        self._frame_graph = torch.cuda.CUDAGraph()
        self._captured_inputs = getattr(self, "_captured_inputs", []) + [self.image_u8]   # the graph reads this buffer
        with torch.cuda.graph(self._frame_graph):
            self._frame_out = self.step_frame()
        return self._frame_graph

    def replay_frame(self):
        self._frame_graph.replay()
        return self._frame_out

    # ---- host-buffer variant: the uint8 image in, the window's poses out -----------------------------
    def make_host_image(self):
        self._u8 = [torch.empty_like(self.image_u8) for _ in range(2)]
        self._cp = torch.cuda.Stream(self.device)
        self._cp_done = [torch.cuda.Event() for _ in range(2)]
        self._cp_free = [torch.cuda.Event() for _ in range(2)]
        self._fi = 0
        graphs = []
        for b in range(2):
            self.image_u8 = self._u8[b]
            self.capture_frame()
            graphs.append(self._frame_graph)
        self._frame_graphs = graphs
        return self.image_u8.cpu().pin_memory(), torch.empty((self.t1 - self.t0, 7), dtype=torch.float32).pin_memory()

    def step_frame_host(self, himg, hposes):
        b = self._fi & 1
        self._fi += 1
        compute = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self._cp):
            self._cp.wait_event(self._cp_free[b])
            self._u8[b].copy_(himg, non_blocking=True)
            self._cp_done[b].record(self._cp)
        compute.wait_event(self._cp_done[b])
        self._frame_graphs[b].replay()
        self._cp_free[b].record(compute)
        hposes.copy_(self.poses[0, self.t0:self.t1], non_blocking=True)
        return himg.numel(), hposes.numel() * 4

    # ---- software pipeline: frame t+1's encoders under frame t's update ------------------------------------
    def make_pipeline(self):
        """The encoders are the only stateless part of a frame (they read the image, nothing else): as their own CUDA
        graph on their own stream, frame t+1's normalise + fnet + inet run UNDER frame t's update iteration; the rest of
        the frame (gathers, slot writes, edge edits, update, keyframe) stays strictly sequential from frame to frame, as
        DPVO's state requires.  Two sets of image / encoder-output buffers; events order copy -> encoders -> rest per
        buffer and keep a buffer from being overwritten while the frame two back still reads it.  Per-frame latency is
        unchanged, throughput goes from the sum towards the maximum of the two parts."""
        dev = self.device
        src = self.image_u8
        self._pu8 = [src.clone() for _ in range(2)]
        self._pcp, self._enc_stream = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        ev = lambda: [torch.cuda.Event() for _ in range(2)]
        self._pcopied, self._enc_done, self._rest_done = ev(), ev(), ev()
        self._enc_graphs, self._rest_graphs, self._enc_out = [], [], []
        for b in range(2):
            self.image_u8 = self._pu8[b]
            for _ in range(2):
                self.step_frame()
            torch.cuda.synchronize(dev)
            ge = torch.cuda.CUDAGraph()
            with torch.cuda.graph(ge):
                self.normalise()
                self.encode()
            ge.replay()
            self._enc_out.append((self.image, self._f, self._i, self._f2))
            torch.cuda.synchronize(dev)
            self.edges.active.count()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                self._pipe_out = self._rest()
            self._enc_graphs.append(ge)
            self._rest_graphs.append(gr)
        self._pfi = 0
        return src.cpu().pin_memory(), torch.empty((self.t1 - self.t0, 7), dtype=torch.float32).pin_memory()

    def step_pipelined(self, himg, hposes):
        b = self._pfi & 1
        self._pfi += 1
        compute = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self._pcp):
            self._pcp.wait_event(self._enc_done[b])            # frame t-2's encoders have read image buffer b
            self._pu8[b].copy_(himg, non_blocking=True)
            self._pcopied[b].record(self._pcp)
        with torch.cuda.stream(self._enc_stream):
            self._enc_stream.wait_event(self._pcopied[b])
            self._enc_stream.wait_event(self._rest_done[b])   # frame t-2's gathers / ring write have read outputs b
            self._enc_graphs[b].replay()
            self._enc_done[b].record(self._enc_stream)
        compute.wait_event(self._enc_done[b])
        self._rest_graphs[b].replay()
        self._rest_done[b].record(compute)
        hposes.copy_(self.poses[0, self.t0:self.t1], non_blocking=True)
        return himg.numel(), hposes.numel() * 4
