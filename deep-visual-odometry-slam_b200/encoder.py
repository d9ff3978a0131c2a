"""Per-frame feature encoders on the device.

`BasicEncoder4` mirrors methods/dpvo/dpvo/extractor.py:200-264 as `Patchifier` uses it (net.py:95-96): the
reference `state_dict` keys, the reference call (`images [b, n, 3, H, W]` float32 in, feature map out), inference
only.  Every layer is a kernel of libdpvo_b200.so (csrc/encoder.cu); activations are channels-last fp16 and the
result is returned as a `[b, n, C, H/4, W/4]` VIEW of the channels-last buffer (`.nhwc` holds the buffer itself, which
is what the feature ring stores).  No CPU path.
"""
import torch

from ._lib import lib, ptr, stream_of, check, require_cuda, c_i32, c_i64, c_f32, c_p

_EPS = 1e-5      # nn.InstanceNorm2d default
_REPLICAS = 16   # DPVO_ENCODER_STATS_REPLICAS


def _conv(x, w, b, n, H, W, cin, cout, k, s, scale, stats, in_stats=None, relu=False, skip=None):
    pad = k // 2
    OH, OW = (H + 2 * pad - k) // s + 1, (W + 2 * pad - k) // s + 1
    out = torch.empty((n, OH, OW, cout), dtype=torch.float16, device=x.device)
    z = c_p(0)
    check(lib.dpvo_encoder_conv(ptr(x), ptr(w), ptr(b), c_i32(n), c_i32(H), c_i32(W), c_i32(cin), c_i32(cout), c_i32(k),
                                c_i32(s), c_f32(scale), ptr(in_stats) if in_stats is not None else z, c_f32(_EPS),
                                c_i32(1 if relu else 0), ptr(skip) if skip is not None else z, ptr(out),
                                ptr(stats) if stats is not None else z, stream_of(x)), "encoder.conv")
    return out


def _norm_act(raw, stats, skip=None, skip_stats=None):
    n, H, W, C = raw.shape
    out = torch.empty_like(raw)
    z = c_p(0)
    check(lib.dpvo_encoder_norm_act(ptr(raw), ptr(stats) if stats is not None else z, ptr(skip) if skip is not None else z,
                                    ptr(skip_stats) if skip_stats is not None else z, c_i32(n), c_i64(H * W), c_i32(C),
                                    c_f32(_EPS), ptr(out), stream_of(raw)), "encoder.norm_act")
    return out


def avg_pool4(x_nhwc):
    """F.avg_pool2d(x, 4, 4) on a channels-last fp16 map [n, H, W, C] (dpvo.py:438)."""
    require_cuda(x_nhwc)
    if x_nhwc.dtype != torch.float16 or x_nhwc.dim() != 4 or not x_nhwc.is_contiguous() or x_nhwc.shape[3] % 8:
        raise RuntimeError("encoder.avg_pool4: expected a contiguous fp16 [n, H, W, C] tensor, C a multiple of 8")
    n, H, W, C = x_nhwc.shape
    out = torch.empty((n, H // 4, W // 4, C), dtype=torch.float16, device=x_nhwc.device)
    check(lib.dpvo_avgpool4_nhwc_f16(ptr(x_nhwc), c_i32(n), c_i32(H), c_i32(W), c_i32(C), ptr(out), stream_of(x_nhwc)),
          "encoder.avg_pool4")
    return out


class FeatureMap:
    """Result of an encoder call: `.nhwc` [n, h, w, C] fp16 buffer, `.nchw5` the reference-shaped [1, n, C, h, w] view."""
    def __init__(self, nhwc):
        self.nhwc = nhwc
        self.nchw5 = nhwc.permute(0, 3, 1, 2)[None]


class BasicEncoder4:
    _CONVS = ["conv1", "layer1.0.conv1", "layer1.0.conv2", "layer1.1.conv1", "layer1.1.conv2", "layer2.0.conv1",
              "layer2.0.conv2", "layer2.0.downsample.0", "layer2.1.conv1", "layer2.1.conv2", "conv2"]

    def __init__(self, state_dict, norm_fn, device="cuda", output_scale=1.0):
        """state_dict: the reference module's (`fnet.` / `inet.` prefix stripped).  norm_fn: 'instance' or 'none'.
        output_scale: folded into the last layer (Patchifier divides both maps by 4, net.py:108-109)."""
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("encoder.BasicEncoder4: needs a CUDA device (no CPU path)")
        if norm_fn not in ("instance", "none"):
            raise RuntimeError("encoder.BasicEncoder4: norm_fn must be 'instance' or 'none' (the two DPVO uses)")
        self.norm, self.device, self.scale = norm_fn == "instance", dev, float(output_scale)
        self.w, self.b = {}, {}
        for name in self._CONVS:
            for suffix in (".weight", ".bias"):
                if name + suffix not in state_dict:
                    raise RuntimeError(f"encoder.BasicEncoder4: state dict lacks {name + suffix}")
            w = torch.as_tensor(state_dict[name + ".weight"]).to(dev, torch.float32)
            self.b[name] = torch.as_tensor(state_dict[name + ".bias"]).to(dev, torch.float32).contiguous()
            if name == "conv1":
                if tuple(w.shape) != (32, 3, 7, 7):
                    raise RuntimeError("encoder.BasicEncoder4: conv1.weight must be [32, 3, 7, 7]")
                packed = torch.zeros((7, 32, 32), dtype=torch.float16, device=dev)      # [ky][cout][kx * 3 + c]
                packed[:, :, :21] = w.permute(2, 0, 3, 1).reshape(7, 32, 21).to(torch.float16)
                self.w[name] = packed
            else:                                    # [cout, cin, k, k] -> fp16 [cout, k*k, cin]
                co, ci, k, _ = w.shape
                self.w[name] = w.permute(0, 2, 3, 1).reshape(co, k * k, ci).to(torch.float16).contiguous()
        self.out_dim = self.w["conv2"].shape[0]
        if self.out_dim % 64:
            raise RuntimeError("encoder.BasicEncoder4: output_dim must be a multiple of 64")

    @classmethod
    def from_reference(cls, module, output_scale=1.0):
        return cls({k: v.detach() for k, v in module.state_dict().items()}, module.norm_fn, output_scale=output_scale)

    def _stats(self, n, c):
        """One plane-statistics block out of the per-call arena (zeroed once per call)."""
        if not self.norm:
            return None
        size = n * _REPLICAS * c * 2
        out = self._arena[self._used:self._used + size]
        self._used += size
        return out

    def _block(self, x, name, cin, cout, stride):
        """ResidualBlock.forward (extractor.py:47-55).  With the instance norm: conv1 -> [norm + ReLU folded into conv2's
        patch staging] -> conv2 -> one norm / ReLU / add / ReLU pass.  Without a norm everything is a conv epilogue."""
        n, H, W, _ = x.shape
        w, b = self.w, self.b
        if not self.norm:
            y = _conv(x, w[name + ".conv1"], b[name + ".conv1"], n, H, W, cin, cout, 3, stride, 1.0, None, relu=True)
            _, OH, OW, _ = y.shape
            if stride != 1:
                x = _conv(x, w[name + ".downsample.0"], b[name + ".downsample.0"], n, H, W, cin, cout, 1, stride, 1.0, None)
            return _conv(y, w[name + ".conv2"], b[name + ".conv2"], n, OH, OW, cout, cout, 3, 1, 1.0, None, relu=True, skip=x)
        s1 = self._stats(n, cout)
        r1 = _conv(x, w[name + ".conv1"], b[name + ".conv1"], n, H, W, cin, cout, 3, stride, 1.0, s1)
        _, OH, OW, _ = r1.shape
        s2 = self._stats(n, cout)
        r2 = _conv(r1, w[name + ".conv2"], b[name + ".conv2"], n, OH, OW, cout, cout, 3, 1, 1.0, s2, in_stats=s1)
        if stride == 1:
            return _norm_act(r2, s2, skip=x)
        s3 = self._stats(n, cout)
        r3 = _conv(x, w[name + ".downsample.0"], b[name + ".downsample.0"], n, H, W, cin, cout, 1, stride, 1.0, s3)
        return _norm_act(r2, s2, skip=r3, skip_stats=s3)

    def __call__(self, images):
        require_cuda(images)
        if images.dim() != 5 or images.shape[2] != 3 or images.dtype != torch.float32:
            raise RuntimeError("encoder.BasicEncoder4: images must be float32 [b, n, 3, H, W]")
        b, n, _, H, W = images.shape
        x = images.reshape(b * n, 3, H, W).contiguous()
        OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
        if self.norm:                       # 32 x 5 + 64 x 5 channels of statistics over the ten normalised layers
            self._arena = torch.zeros(b * n * _REPLICAS * 2 * (32 * 5 + 64 * 5), dtype=torch.float64, device=x.device)
            self._used = 0
        s0 = self._stats(b * n, 32)
        r0 = torch.empty((b * n, OH, OW, 32), dtype=torch.float16, device=x.device)
        check(lib.dpvo_encoder_conv7x7s2(ptr(x), ptr(self.w["conv1"]), ptr(self.b["conv1"]), c_i32(b * n), c_i32(H), c_i32(W),
                                         c_i32(0 if self.norm else 1), ptr(r0), ptr(s0) if s0 is not None else c_p(0),
                                         stream_of(x)), "encoder.conv1")
        y = _norm_act(r0, s0) if self.norm else r0
        y = self._block(y, "layer1.0", 32, 32, 1)
        y = self._block(y, "layer1.1", 32, 32, 1)
        y = self._block(y, "layer2.0", 32, 64, 2)
        y = self._block(y, "layer2.1", 64, 64, 1)
        _, h, w, _ = y.shape
        out = _conv(y, self.w["conv2"], self.b["conv2"], b * n, h, w, 64, self.out_dim, 1, 1, self.scale, None)
        return FeatureMap(out)


class Patchifier(torch.nn.Module):
    """`Patchifier.forward` (net.py:106-153) for inference: both encoders, the three patch gathers, the patch grid.

    fnet / inet: BasicEncoder4 instances built with output_scale=0.25 (the reference divides both maps by 4).
    Returns the reference's tuple (fmap, gmap, imap, patches, index[, clr]); `fmap` is the [1, n, 128, h, w] view of the
    channels-last buffer, which is also kept as `.fmap_nhwc` together with its 4x4 average pool `.fmap2_nhwc` — the two
    pyramid levels in the layout the feature ring stores (dpvo.py:437-438), so the ring write is a plain copy.
    Patch centres: `coords` ([n, M, 2] float32, integer valued) or, like the reference's 'RANDOM' strategy, uniform in
    [1, w-1) x [1, h-1).  'GRADIENT_BIAS' (net.py:117-129) is not built."""

    def __init__(self, fnet, inet, patch_size=3):
        super().__init__()      # an nn.Module (no parameters) so it can replace `network.patchify` (net.py:175)
        self.fnet, self.inet, self.patch_size = fnet, inet, patch_size

    @classmethod
    def from_reference(cls, module):
        return cls(BasicEncoder4.from_reference(module.fnet, 0.25), BasicEncoder4.from_reference(module.inet, 0.25),
                   module.patch_size)

    use_graph = True        # replay the two encoders as one CUDA graph per image shape (≈30 launches -> 1)

    def _encode(self, images):
        """fnet and inet are independent and each layer fills a fraction of the GPU: inet runs beside fnet on a side
        stream (two branches when captured)."""
        main = torch.cuda.current_stream(images.device)          # inet beside fnet on a side stream (independent nets)
        if getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream(images.device)
        fork, join = torch.cuda.Event(), torch.cuda.Event()
        fork.record(main)
        self._side.wait_event(fork)
        with torch.cuda.stream(self._side):
            i = self.inet(images)
            i.nhwc.record_stream(main)
            join.record(self._side)
        f = self.fnet(images)
        f2 = avg_pool4(f.nhwc)
        main.wait_event(join)
        return f, i, f2

    def _encode_graphed(self, images):
        """One graph per (shape, device): the image is copied into the graph's input buffer, the encoders' outputs live
        in the graph's buffers and are valid until the next call (DPVO copies them into its rings right away)."""
        if not hasattr(self, "_graphs"):
            self._graphs = {}
        key = (tuple(images.shape), images.device.index)
        ent = self._graphs.get(key)
        if ent is None:
            static = images.clone()
            self._encode(static)                                  # warm-up outside the capture (lazy initialisations)
            torch.cuda.synchronize(images.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                out = self._encode(static)
            ent = self._graphs[key] = (g, static, out)
        g, static, out = ent
        static.copy_(images)
        g.replay()
        return out

    def forward(self, images, patches_per_image=80, disps=None, centroid_sel_strat="RANDOM", return_color=False,
                coords=None):
        from . import altcorr
        if centroid_sel_strat != "RANDOM":
            raise NotImplementedError(f"Patch centroid selection not built: {centroid_sel_strat}")
        if self.use_graph and not torch.cuda.is_current_stream_capturing():
            f, i, f2 = self._encode_graphed(images)
        else:
            f, i, f2 = self._encode(images)
        self.fmap_nhwc = f.nhwc
        self.fmap2_nhwc = f2
        fmap, imap = f.nchw5, i.nchw5
        b, n, c, h, w = fmap.shape
        P = self.patch_size
        dev = images.device
        integer = coords is None
        if coords is None:
            x = torch.randint(1, w - 1, size=[n, patches_per_image], device=dev)
            y = torch.randint(1, h - 1, size=[n, patches_per_image], device=dev)
            coords = torch.stack([x, y], dim=-1).float()
        if integer:
            # integer centres (the reference's randint, net.py:113-115): the bilinear blend of altcorr.patchify has zero
            # fractional weights, i.e. it IS the leading (2r+1)^2 window of the raw (2r+2)^2 gather — one launch per map
            from . import cuda_corr

            def gather(net, xy, r):
                d = 2 * r + 1
                return cuda_corr.patchify_forward(net, xy, r)[0][..., :d, :d]
        else:
            gather = altcorr.patchify
        imap_p = gather(imap[0], coords, 0).reshape(b, -1, imap.shape[2], 1, 1)
        gmap = gather(fmap[0], coords, P // 2).reshape(b, -1, c, P, P)
        gkey = (b, n, h, w, dev.index)
        grid = getattr(self, "_grid_cache", {}).get(gkey) if disps is None else None
        if grid is None:
            d = torch.ones(b, n, h, w, device=dev) if disps is None else disps
            ys, xs = torch.meshgrid(torch.arange(h, device=dev, dtype=torch.float32),
                                    torch.arange(w, device=dev, dtype=torch.float32), indexing="ij")
            grid = torch.stack([xs.expand(b, n, h, w), ys.expand(b, n, h, w), d], dim=2)     # utils.py:39-50
            if disps is None:                                    # inference: the unit-disparity grid never changes
                if not hasattr(self, "_grid_cache"):
                    self._grid_cache = {}
                self._grid_cache[gkey] = grid
        patches = gather(grid[0], coords, P // 2).reshape(b, -1, 3, P, P)
        index = torch.arange(n, device=dev).view(n, 1).repeat(1, coords.shape[1]).reshape(-1)
        if return_color:
            clr = gather(images[0], 4 * (coords + 0.5), 0).reshape(b, -1, 3)
            return fmap, gmap, imap_p, patches, index, clr
        return fmap, gmap, imap_p, patches, index
