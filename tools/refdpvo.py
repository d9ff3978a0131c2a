#!/usr/bin/env python
"""Run the reference's OWN Python — `dpvo.dpvo.DPVO` (dpvo/dpvo.py:20-473), `dpvo.net.VONet` — unmodified,
on top of one of two back ends for the three bare-name extension modules it imports
(`cuda_corr`, `cuda_ba`, `lietorch_backends`):

    --backend ref      the reference's own CUDA extensions, compiled untouched into oracle/_ref/*.so
    --backend dropin   this repo's sm_100a implementation (deep-visual-odometry-slam_b200/dropin/)

HARNESS / TEST INFRASTRUCTURE, not product code.  The reference package itself is the untouched copy
that oracle/build_ref.py installs under oracle/_ref/site/ (git-ignored; it travels to the GPU box like
the prebuilt .so files).  Two stubs only, as SURVEY.md §8(c) specifies: an empty `pypose` module
(imported by loop_closure/optim_utils.py:4, never called on the default path) and a plain namespace
carrying dpvo/config.py:3-36 + config/default.yaml instead of a yacs node.

Synthetic 480x640 sequence (no dataset in the image): a textured plane seen from a camera that flies
a closed loop with fast and slow stretches, so that the DPVO state machine goes through everything
it has: the 8-frame initialisation with 12 updates (dpvo.py:461-466), keyframe removal
(dpvo.py:266-310), proximity loop-closure edges (patchgraph.py:56-82) and global bundle adjustment
(dpvo.py:312-326).  Random-initialised weights cannot track, so the Update operator's output is
TEACHER-FORCED: the harness wraps `network.update` (the wrapped module still runs — correlation,
Update operator, neighbours — and its outputs still enter the state), and replaces the predicted flow
by  (ground-truth reprojection - current reprojection) + ALPHA * tanh(predicted flow).  Both back ends
get the identical wrapper, images, seeds and weights.

Modes
    parity   per-frame log of n, m, edge counts, sha1 of ii/jj/kk, poses (npz)
    fps      wall-clock frames/s of  for t: slam(t, image.cuda(), intrinsics)  + terminate()
"""
import argparse
import hashlib
import importlib
import json
import math
import os
import sys
import time
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "deep-visual-odometry-slam_b200"
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
SITE = os.path.join(REF_DIR, "site")
DROPIN = os.path.join(ROOT, PKG, "dropin")


# --------------------------------------------------------------------------- configuration
def make_cfg(**over):
    """dpvo/config.py:3-36 defaults overlaid with config/default.yaml (what evaluate_tartan.py does with
    cfg.merge_from_file); a plain namespace because DPVO only reads attributes."""
    import yaml
    cfg = dict(BUFFER_SIZE=4096, CENTROID_SEL_STRAT="RANDOM", PATCHES_PER_FRAME=80, REMOVAL_WINDOW=20,
               OPTIMIZATION_WINDOW=12, PATCH_LIFETIME=12, KEYFRAME_INDEX=4, KEYFRAME_THRESH=12.5,
               MOTION_MODEL="DAMPED_LINEAR", MOTION_DAMPING=0.5, MIXED_PRECISION=True, LOOP_CLOSURE=False,
               BACKEND_THRESH=64.0, MAX_EDGE_AGE=1000, GLOBAL_OPT_FREQ=15, CLASSIC_LOOP_CLOSURE=False,
               LOOP_CLOSE_WINDOW_SIZE=3, LOOP_RETR_THRESH=0.04)
    with open(os.path.join(SITE, "config", "default.yaml")) as f:
        cfg.update(yaml.safe_load(f))
    cfg.update(over)
    return types.SimpleNamespace(**cfg)


def have_reference_install():
    return os.path.isdir(os.path.join(SITE, "dpvo"))


def have_reference_ext():
    return all(os.path.exists(os.path.join(REF_DIR, n + ".so")) for n in ("cuda_corr", "cuda_ba", "lietorch_backends"))


# --------------------------------------------------------------------------- back-end selection
class _PyposeStub(types.ModuleType):
    """`import pypose as pp` at loop_closure/optim_utils.py:4 only needs the NAMES used in annotations
    (`pp.SE3`, `pp.Sim3`) to exist at import; nothing on the default DPVO path calls them."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)

        def _absent(*a, **k):
            raise RuntimeError(f"pypose.{name} is not available in this image (harness stub)")
        return _absent


def select_backend(backend, cpu_dryrun=False):
    """Must run before the first `import dpvo`.  One back end per process (the reference binds the
    extension modules at import time: altcorr/correlation.py:2, fastba/ba.py:2-5, lietorch/group_ops.py:1)."""
    if "dpvo" in sys.modules:
        raise RuntimeError("refdpvo: a back end was already selected in this process")
    if not have_reference_install():
        raise RuntimeError(f"{SITE} is missing: run `python oracle/build_ref.py site` where /root/reference exists")
    if "pypose" not in sys.modules:
        try:
            import pypose  # noqa: F401
        except Exception:
            sys.modules["pypose"] = _PyposeStub("pypose")
    if cpu_dryrun:
        _install_cpu_dryrun_modules()
        sys.path.insert(0, REF_DIR)          # lietorch_backends.so of the reference has a CPU backend
    elif backend == "ref":
        sys.path.insert(0, REF_DIR)
    elif backend == "dropin":
        sys.path.insert(0, DROPIN)
    else:
        raise ValueError(backend)
    sys.path.insert(0, SITE)
    import torch  # noqa: F401  (libtorch before the pybind modules)
    mods = {n: importlib.import_module(n) for n in ("cuda_corr", "cuda_ba", "lietorch_backends")}
    return {n: getattr(m, "__file__", "<memory>") for n, m in mods.items()}


def _install_cpu_dryrun_modules():
    """CPU stand-ins for cuda_corr / cuda_ba built on the oracle restatements — ONLY to exercise this
    harness's own logic (teacher, sequence, logging) in a container without a GPU."""
    import numpy as np
    import torch
    sys.path.insert(0, ROOT)
    from oracle import altcorr_torch as ot, altcorr_oracle as ao, fastba_oracle as fo

    cc = types.ModuleType("cuda_corr")

    def forward(fmap1, fmap2, coords, ii, jj, radius):
        prep = ot.prepare(fmap1, [fmap2])
        out = ot.corr_pyramid(prep, [1.0], coords, ii, jj, radius, dtype=fmap1.dtype)
        D = 2 * radius + 1
        P = coords.shape[-1]
        return [out.view(1, -1, D, D, P, P)]

    def patchify_forward(net, coords, radius):
        return [torch.as_tensor(ao.patchify_forward(net.numpy(), coords.numpy(), radius)).to(net.dtype)]

    cc.forward, cc.patchify_forward = forward, patchify_forward
    cb = types.ModuleType("cuda_ba")

    def ba_forward(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, PPF, t0, t1, iterations, eff_impl):
        try:
            return _ba_forward(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, PPF, t0, t1, iterations, eff_impl)
        except Exception:
            import traceback
            traceback.print_exc()
            raise

    def _ba_forward(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, PPF, t0, t1, iterations, eff_impl):
        p, z = fo.ba(poses[0].numpy(), patches[0].numpy(), intrinsics[0].numpy(), target[0].numpy(), weight[0].numpy(),
                     lmbda.numpy(), ii.numpy(), jj.numpy(), kk.numpy(), PPF, t0, t1, iterations, eff_impl)
        poses[0].copy_(torch.as_tensor(p).float())
        patches[0].copy_(torch.as_tensor(z).float())
        return []

    def neighbors(ii, jj):
        a, b = fo.neighbors(ii.numpy(), jj.numpy())
        return [torch.as_tensor(a), torch.as_tensor(b)]

    cb.forward, cb.neighbors = ba_forward, neighbors
    cb.reproject = None
    sys.modules["cuda_corr"], sys.modules["cuda_ba"] = cc, cb


class _CpuRemap:
    """torch-function mode that sends every device='cuda' request to the CPU (dry run only)."""

    def __new__(cls):
        import torch
        from torch.overrides import TorchFunctionMode

        def is_cuda(d):
            return (isinstance(d, str) and d.startswith("cuda")) or (isinstance(d, torch.device) and d.type == "cuda")

        class Mode(TorchFunctionMode):
            def __torch_function__(self, func, types_, args=(), kwargs=None):
                kwargs = dict(kwargs or {})
                if is_cuda(kwargs.get("device")):
                    kwargs["device"] = "cpu"
                if func is torch.Tensor.cuda:
                    return args[0]
                if func is torch.Tensor.to:
                    args = tuple("cpu" if is_cuda(a) else a for a in args)
                return func(*args, **kwargs)

        return Mode()


# --------------------------------------------------------------------------- synthetic sequence
class SynthSequence:
    """T frames of a textured plane, camera on a closed loop in the x-y plane with fast / slow stretches.

    gt[t]      world->camera pose (tx ty tz qx qy qz qw), float64
    depth(x, y, t)  inverse depth of the pixel (x, y) (1/4-resolution pixel units) of frame t — a smooth
               function; only each patch's own 3-D point is ever needed (teacher), so no cross-frame
               consistency of the depth maps is required.
    image(t)   uint8 [3, H, W] crop of one big texture, shifted with the camera (host tensor)
    """

    def __init__(self, frames, seed=0, ht=480, wd=640, radius=3.0, fast=0.30, slow=0.06, fast_len=44, slow_len=10):
        import numpy as np
        self.T, self.ht, self.wd = frames, ht, wd
        self.intrinsics = np.array([wd / 2.0, wd / 2.0, wd / 2.0, ht / 2.0], np.float32)   # 320 320 320 240
        rng = np.random.RandomState(seed)
        # arc length along the loop
        s = np.zeros(frames)
        for t in range(1, frames):
            ph = (t % (fast_len + slow_len))
            s[t] = s[t - 1] + (fast if ph < fast_len else slow)
        phi = s / radius
        c = np.stack([radius * np.cos(phi) - radius, radius * np.sin(phi), 0.15 * np.sin(0.05 * np.arange(frames))], -1)
        yaw = 0.03 * np.sin(0.07 * np.arange(frames))
        roll = 0.02 * np.sin(0.045 * np.arange(frames) + 1.0)
        gt = np.zeros((frames, 7))
        for t in range(frames):
            Rz = np.array([[math.cos(roll[t]), -math.sin(roll[t]), 0], [math.sin(roll[t]), math.cos(roll[t]), 0], [0, 0, 1]])
            Ry = np.array([[math.cos(yaw[t]), 0, math.sin(yaw[t])], [0, 1, 0], [-math.sin(yaw[t]), 0, math.cos(yaw[t])]])
            R = Rz @ Ry
            gt[t, :3] = -R @ c[t]
            gt[t, 3:] = _rot_to_quat(R)
        self.centres, self.gt = c, gt
        self.z0 = 2.5
        # texture: smooth blobs + fine grain, large enough for every crop
        self.px_per_unit = (wd / 2.0) / self.z0
        span = int(2 * radius * self.px_per_unit) + 8
        th, tw = ht + span + 64, wd + span + 64
        low = rng.rand(3, th // 16 + 2, tw // 16 + 2).astype(np.float32)
        tex = np.kron(low, np.ones((1, 16, 16), np.float32))[:, :th, :tw]
        tex = 0.6 * tex + 0.4 * rng.rand(3, th, tw).astype(np.float32)
        self.tex = (tex * 255).astype(np.uint8)
        self._span = span

    def depth(self, x, y, t):
        import torch
        return 0.4 + 0.1 * torch.sin(0.05 * x + 0.3 * t) * torch.cos(0.07 * y)

    def image(self, t):
        import numpy as np
        c = self.centres[t]
        ox = int(round(c[0] * self.px_per_unit)) + self._span + 16 - 1
        oy = int(round(c[1] * self.px_per_unit)) + self._span // 2 + 16
        ox = min(max(ox, 0), self.tex.shape[2] - self.wd)
        oy = min(max(oy, 0), self.tex.shape[1] - self.ht)
        return np.ascontiguousarray(self.tex[:, oy:oy + self.ht, ox:ox + self.wd])

    def host_frames(self, pin=True):
        import torch
        buf = torch.empty((self.T, 3, self.ht, self.wd), dtype=torch.uint8)
        for t in range(self.T):
            buf[t] = torch.from_numpy(self.image(t))
        if pin and torch.cuda.is_available():
            buf = buf.pin_memory()
        return buf


def _rot_to_quat(R):
    import numpy as np
    from scipy.spatial.transform import Rotation
    return Rotation.from_matrix(R).as_quat()       # x y z w


# --------------------------------------------------------------------------- teacher
def _load_teacher_lib():
    """tools/libteacher.so (tools/teacher_kernel.cu, built by __graft_entry__.build()): the teacher in one launch.
    REFDPVO_TORCH_TEACHER=1 keeps the torch formulation (tests compare the two)."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libteacher.so")
    if os.environ.get("REFDPVO_TORCH_TEACHER") == "1" or not os.path.exists(path):
        return None
    import ctypes
    lib = ctypes.CDLL(path)
    lib.teacher_targets.restype = ctypes.c_int
    return lib


_TEACHER_LIB = _load_teacher_lib()


def _teacher_kind(args):
    if args.no_teacher:
        return "none"
    return "one-launch kernel (tools/teacher_kernel.cu)" if (_TEACHER_LIB is not None and not args.cpu_dryrun) else "torch ops"


def _qrot(q, v):
    """rotate v [.,3] by unit quaternion q [.,4] (x y z w)"""
    import torch
    u = q[..., :3]
    w = q[..., 3:]
    t = 2.0 * torch.cross(u, v, dim=-1)
    return v + w * t + torch.cross(u, t, dim=-1)


def make_teacher(inner, slam, seq, alpha):
    import torch

    class Teacher(torch.nn.Module):
        """Wraps the Update operator (net.py:70-88 call signature).  The wrapped module runs unchanged."""

        def __init__(self):
            super().__init__()
            self.inner = inner
            self.gt = None
            self._conj = None
            self._pin, self._pin_i = None, 0

        def _tstamps(self, n_hi, dev):
            """time stamps of the window on the device WITHOUT a device synchronisation: a pageable host-to-device copy
            (torch.as_tensor(numpy, device=cuda)) waits for everything queued before it, i.e. it stalled the tracker in
            the middle of every update; here the stamps go through a small ring of pinned buffers."""
            ts = slam.pg.tstamps_[:n_hi + 1]
            if dev.type != "cuda":
                return torch.as_tensor(ts.copy(), device=dev)
            if self._pin is None:
                cap = len(slam.pg.tstamps_)
                self._pin = [torch.empty(cap, dtype=torch.from_numpy(ts[:1]).dtype).pin_memory() for _ in range(8)]
            buf = self._pin[self._pin_i % len(self._pin)]
            self._pin_i += 1
            buf[:len(ts)].copy_(torch.from_numpy(ts))
            return buf[:len(ts)].to(dev, non_blocking=True)

        def gt_centres(self, ii, jj, kk):
            dev = ii.device
            if self.gt is None or self.gt.device != dev or self._conj is None:
                self.gt = torch.as_tensor(seq.gt, dtype=torch.float64, device=dev)
                self._conj = torch.tensor([-1.0, -1.0, -1.0, 1.0], dtype=torch.float64, device=dev)
            n_hi = int(slam.n) + 1
            ts = self._tstamps(n_hi, dev)
            G = self.gt[ts]                                            # [n+2, 7]
            P = slam.pg.patches_.view(-1, 3, 3, 3)[kk]                # [E,3,3,3]
            x = P[:, 0, 1, 1].double()
            y = P[:, 1, 1, 1].double()
            K = slam.pg.intrinsics_[ii].double()
            Kj = slam.pg.intrinsics_[jj].double()
            d = seq.depth(x, y, ts[ii].double())
            Xi = torch.stack([(x - K[:, 2]) / K[:, 0], (y - K[:, 3]) / K[:, 1], torch.ones_like(x)], -1) / d[:, None]
            Gi, Gj = G[ii], G[jj]
            qi_inv = Gi[:, 3:] * self._conj
            Xw = _qrot(qi_inv, Xi - Gi[:, :3])
            Xj = _qrot(Gj[:, 3:], Xw) + Gj[:, :3]
            z = Xj[:, 2]
            ok = z > 0.2
            zc = torch.where(ok, z, torch.ones_like(z))
            u = Kj[:, 0] * Xj[:, 0] / zc + Kj[:, 2]
            v = Kj[:, 1] * Xj[:, 1] / zc + Kj[:, 3]
            return torch.stack([u, v], -1).float()[None], ok[None, :, None]

        def fused(self, delta, weight, coords, ii, jj, kk):
            """the same targets from ONE launch (tools/teacher_kernel.cu, float64 like gt_centres) — ~70 torch kernels
            per update otherwise, which is harness time on the host-bound arm"""
            dev = ii.device
            if self.gt is None or self.gt.device != dev:
                self.gt = torch.as_tensor(seq.gt, dtype=torch.float64, device=dev).contiguous()
            ts = self._tstamps(int(slam.n) + 1, dev)
            pt, K = slam.pg.patches_, slam.pg.intrinsics_
            E = ii.numel()
            if not (pt.is_contiguous() and K.is_contiguous() and coords.is_contiguous() and pt.dtype == torch.float32
                    and K.dtype == torch.float32 and coords.dtype == torch.float32 and tuple(coords.shape[2:]) == (2, 3, 3)
                    and tuple(pt.shape[-3:]) == (3, 3, 3) and ii.is_contiguous() and jj.is_contiguous() and kk.is_contiguous()):
                raise RuntimeError("teacher kernel: unexpected buffer layout")
            d32 = delta.float().contiguous().view(-1, 2)
            w32 = weight.float().contiguous().view(-1, 2)
            d = torch.empty((1, E, 2), dtype=torch.float32, device=dev)
            w = torch.empty((1, E, 2), dtype=torch.float32, device=dev)
            import ctypes as C
            p = lambda t: C.c_void_p(t.data_ptr())
            rc = _TEACHER_LIB.teacher_targets(
                p(self.gt), C.c_int64(self.gt.shape[0]), p(ts), C.c_int64(ts.numel()), p(pt), C.c_int64(pt.numel() // 27),
                p(K), p(ii), p(jj), p(kk), C.c_int64(E), p(coords), C.c_int64(18), C.c_int64(9), C.c_int64(4),
                p(d32), p(w32), C.c_float(alpha), p(d), p(w), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
            if rc != 0:
                raise RuntimeError(f"teacher kernel launch failed: CUDA error {rc}")
            return d, w

        def forward(self, net, inp, corr, flow, ii, jj, kk, **kw):
            net, (delta, weight, _) = self.inner(net, inp, corr, flow, ii, jj, kk, **kw)
            coords = slam._last_coords                                  # [1,E,2,P,P]
            if _TEACHER_LIB is not None and ii.is_cuda:
                d, w = self.fused(delta, weight, coords, ii, jj, kk)
                return net, (d, w, None)
            c = coords[:, :, :, 1, 1]
            gt, ok = self.gt_centres(ii, jj, kk)
            d = (gt - c) + alpha * torch.tanh(delta.float())
            d = torch.where(ok, d, torch.zeros_like(d))
            w = torch.where(ok, weight.float(), torch.zeros_like(weight, dtype=torch.float))
            return net, (d.to(delta.dtype) if delta.dtype == torch.float32 else d, w, None)

    return Teacher()


# --------------------------------------------------------------------------- building a DPVO instance
def build_slam(seed=1234, cfg=None, ht=480, wd=640, seq=None, alpha=0.25, teacher=True, update_impl="reference",
               patchify_impl="reference", pipeline="reference", cpu_dryrun=False):
    """Unmodified reference DPVO (dpvo.py:22-112) with a seeded random-init VONet (no checkpoint in the image)."""
    import torch
    dpvo_mod = importlib.import_module("dpvo.dpvo")
    net_mod = importlib.import_module("dpvo.net")
    torch.manual_seed(seed)
    network = net_mod.VONet()
    cfg = cfg or make_cfg()
    if pipeline == "b200":
        # this repo's DPVO class (deep-visual-odometry-slam_b200/dpvo.py): same API and state machine on the B200
        # operators and the device-resident buffer layout; the reference VONet only supplies the seeded weights
        sys.path.insert(0, ROOT)
        slam = importlib.import_module(PKG + ".dpvo").DPVO(cfg, network, ht=ht, wd=wd, viz=False)
        update_impl = patchify_impl = "bound"
    else:
        slam = dpvo_mod.DPVO(cfg, network, ht=ht, wd=wd, viz=False)
    if update_impl == "b200":
        # the one-line binding of the Update operator (next row N1)
        import dpvo_b200_update
        slam.network.update = dpvo_b200_update.Update.from_reference(slam.network.update)
    if patchify_impl == "b200":
        import dpvo_b200_patchify
        slam.network.patchify = dpvo_b200_patchify.Patchifier.from_reference(slam.network.patchify)
    if teacher:
        if seq is None:
            raise ValueError("teacher needs the sequence (ground truth)")
        orig = slam.reproject

        def reproject(indicies=None):
            c = orig(indicies=indicies)
            slam._last_coords = c
            return c

        slam.reproject = reproject
        slam.network.update = make_teacher(slam.network.update, slam, seq, alpha)
    torch.manual_seed(seed + 1)
    return slam


def _sha(t):
    return hashlib.sha1(t.detach().cpu().numpy().tobytes()).hexdigest()[:16]


def run_parity(slam, seq, frames, keep_every=10, dev="cuda"):
    """Feed `frames` frames; after each, log the bookkeeping state.  Returns a dict of numpy arrays."""
    import numpy as np
    import torch
    intr = torch.as_tensor(seq.intrinsics).to(dev)
    log = dict(n=[], m=[], E=[], Einac=[], h_ii=[], h_jj=[], h_kk=[], removed=[], gba=[], loop_added=[])
    poses, kept = [], {}
    gba_count = 0
    for t in range(frames):
        n0 = slam.n
        last_gba0 = getattr(slam, "last_global_ba", None)
        ran0 = int(slam.ran_global_ba.sum())
        img = torch.from_numpy(seq.image(t)).to(dev)
        slam(t, img, intr)
        pg = slam.pg
        log["n"].append(slam.n); log["m"].append(slam.m)
        log["E"].append(pg.ii.numel()); log["Einac"].append(pg.ii_inac.numel())
        log["h_ii"].append(_sha(pg.ii)); log["h_jj"].append(_sha(pg.jj)); log["h_kk"].append(_sha(pg.kk))
        log["removed"].append(int(slam.is_initialized and slam.n == n0 and t > 0 and n0 >= 8))
        ran1 = int(slam.ran_global_ba.sum())
        log["gba"].append(int(ran1 > ran0)); gba_count += int(ran1 > ran0)
        log["loop_added"].append(int(getattr(slam, "last_global_ba", None) != last_gba0))
        # monocular scale is a gauge freedom of the window (only pose 0 is held fixed, dpvo.py:352): log the poses in
        # the normalisation DPVO itself applies before every global BA (patchgraph.py:84-90: mean inverse depth = 1)
        P = torch.zeros(64, 7)
        k = min(slam.n, 64)
        s = pg.patches_[:slam.n, :, 2].mean() if slam.n > 0 else torch.ones(())
        Pn = pg.poses_[slam.n - k:slam.n].clone()
        Pn[:, :3] *= s
        P[:k] = Pn.cpu()
        poses.append(P.numpy())
        log.setdefault("scale", []).append(float(s))
        if t % keep_every == 0 or log["removed"][-1] or log["gba"][-1] or log["loop_added"][-1]:
            kept[f"ii_{t}"] = pg.ii.cpu().numpy().astype(np.int32)
            kept[f"jj_{t}"] = pg.jj.cpu().numpy().astype(np.int32)
            kept[f"kk_{t}"] = pg.kk.cpu().numpy().astype(np.int32)
    out = {k: np.array(v) for k, v in log.items()}
    out["poses_tail"] = np.stack(poses)
    out["final_n"] = np.array(slam.n)
    out["final_poses"] = slam.pg.poses_[:slam.n].cpu().numpy()
    out["final_depths"] = slam.pg.patches_[:slam.n, :, 2, 1, 1].cpu().numpy()
    out["tstamps"] = slam.pg.tstamps_[:slam.n].copy()
    out.update(kept)
    return out


def run_fps(slam, seq, frames, host_frames, terminate=True, dev="cuda"):
    """evaluate_tartan.py-style loop: per frame one H2D image copy + slam(t, image, intrinsics); terminate()
    (12 more updates, dpvo.py:173-198) inside the timed region.  Wall clock, device synchronised both sides."""
    import torch
    intr = torch.as_tensor(seq.intrinsics).to(dev)
    if dev != "cpu":
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for t in range(frames):
        img = host_frames[t].to(dev, non_blocking=True)
        slam(t, img, intr)
    if dev != "cpu":
        torch.cuda.synchronize()
    t1 = time.perf_counter()
    out = dict(frames=frames, loop_s=t1 - t0, fps_loop=frames / (t1 - t0), final_n=int(slam.n),
               edges=int(slam.pg.ii.numel()), inactive_edges=int(slam.pg.ii_inac.numel()),
               global_ba_calls=int(slam.ran_global_ba.sum()))
    if terminate:
        poses, _ = slam.terminate()
        if dev != "cpu":
            torch.cuda.synchronize()
        t2 = time.perf_counter()
        import numpy as np
        out.update(terminate_s=t2 - t1, total_s=t2 - t0, fps=frames / (t2 - t0),
                   finite=bool(np.isfinite(poses).all()))
    else:
        out.update(fps=out["fps_loop"])
    return out


def instrument(slam):
    """--mode profile: wall time per stage with a device synchronisation on both sides of every stage (so the
    total is slower than --mode fps; the SHARES are what this is for)."""
    import torch
    acc = {}

    def wrap(fn, key):
        def timed(*a, **k):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            try:
                return fn(*a, **k)
            finally:
                torch.cuda.synchronize()
                c = acc.setdefault(key, [0, 0.0])
                c[0] += 1
                c[1] += time.perf_counter() - t0
        return timed

    ours = type(slam).__module__.startswith(PKG)
    dm = sys.modules[type(slam).__module__] if ours else sys.modules["dpvo.dpvo"]
    slam.update = wrap(slam.update, "update (whole)")
    slam.keyframe = wrap(slam.keyframe, "keyframe")
    slam.corr = wrap(slam.corr, "  corr x2 + stack")
    slam.reproject = wrap(slam.reproject, "  reproject (pops.transform)")
    slam.motion_probe = wrap(slam.motion_probe, "motion_probe")
    if ours:
        slam.edges_loop = wrap(slam.edges_loop, "edges_loop")
        slam.normalize = wrap(slam.normalize, "  normalize")
    else:
        slam.pg.edges_loop = wrap(slam.pg.edges_loop, "edges_loop")
    slam.network.patchify.forward = wrap(slam.network.patchify.forward, "patchify (encoders + gathers)")
    slam.network.update.forward = wrap(slam.network.update.forward, "  Update operator (+teacher)")
    orig_ba = dm.fastba.BA

    def ba(*a, **k):
        key = "  BA global (eff_impl)" if k.get("eff_impl") else "  BA window"
        return wrap(orig_ba, key)(*a, **k)

    dm.fastba.BA = ba
    dm.pops.point_cloud = wrap(dm.pops.point_cloud, "  point_cloud")
    return acc


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="dropin", choices=["ref", "dropin"])
    ap.add_argument("--mode", default="parity", choices=["parity", "fps", "profile"])
    ap.add_argument("--frames", type=int, default=120)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--alpha", type=float, default=0.25)
    ap.add_argument("--update", default="reference", choices=["reference", "b200"])
    ap.add_argument("--patchify", default="reference", choices=["reference", "b200"])
    ap.add_argument("--pipeline", default="reference", choices=["reference", "b200"],
                    help="b200: this repo's DPVO class instead of the reference's (needs --backend dropin)")
    ap.add_argument("--no-teacher", action="store_true")
    ap.add_argument("--no-loop-closure", action="store_true")
    ap.add_argument("--patches", type=int, default=None)
    ap.add_argument("--out", default=None)
    ap.add_argument("--cpu-dryrun", action="store_true", help="no GPU: oracle stand-ins, every cuda device -> cpu")
    args = ap.parse_args()

    import torch
    files = select_backend(args.backend, cpu_dryrun=args.cpu_dryrun)
    over = {}
    if args.no_loop_closure:
        over["LOOP_CLOSURE"] = False
    if args.patches:
        over["PATCHES_PER_FRAME"] = args.patches
    if args.cpu_dryrun:
        over["MIXED_PRECISION"] = False
    seq = SynthSequence(args.frames, seed=args.seed)
    dev = "cpu" if args.cpu_dryrun else "cuda"
    ctx = _CpuRemap() if args.cpu_dryrun else None
    if ctx is not None:
        ctx.__enter__()
    torch.set_grad_enabled(False)             # evaluate_tartan.py:46 runs DPVO under @torch.no_grad()
    try:
        slam = build_slam(seed=args.seed, cfg=make_cfg(**over), seq=seq, alpha=args.alpha, teacher=not args.no_teacher,
                          update_impl=args.update, patchify_impl=args.patchify, pipeline=args.pipeline,
                          cpu_dryrun=args.cpu_dryrun)
        if args.mode == "parity":
            if hasattr(slam, "verify_plan"):
                slam.verify_plan = True        # this repo's class: every keyframe() compares its one-read plan with the device counts
            res = run_parity(slam, seq, args.frames, dev=dev)
            import numpy as np
            summary = dict(backend=args.backend, pipeline=args.pipeline, frames=args.frames, final_n=int(res["final_n"]),
                           removed=int(res["removed"].sum()), gba=int(res["gba"].sum()),
                           loop_added=int(res["loop_added"].sum()), E_last=int(res["E"][-1]),
                           Einac_last=int(res["Einac"][-1]), modules=files, teacher=_teacher_kind(args))
            if args.out:
                np.savez_compressed(args.out, **res)
            print(json.dumps(summary), flush=True)
        else:
            hf = seq.host_frames(pin=not args.cpu_dryrun)
            acc = instrument(slam) if args.mode == "profile" else None
            res = run_fps(slam, seq, args.frames, hf, dev=dev)
            if acc is not None:
                res["stages"] = {k: {"calls": c, "s": round(t, 3), "ms_per_call": round(1e3 * t / max(c, 1), 3)}
                                 for k, (c, t) in sorted(acc.items(), key=lambda kv: -kv[1][1])}
            res.update(backend=args.backend, update=args.update, patchify=args.patchify, pipeline=args.pipeline, modules=files,
                       teacher=_teacher_kind(args))
            if args.out:
                with open(args.out, "w") as f:
                    json.dump(res, f)
            print(json.dumps(res), flush=True)
    finally:
        if ctx is not None:
            ctx.__exit__(None, None, None)
    return 0


if __name__ == "__main__":
    sys.exit(main())
