#!/usr/bin/env python
"""BASELINE config 4: the reference's training step — `dpvo.net.VONet.forward` (dpvo/net.py:184-268, 15-frame clip,
80 patches/frame, STEPS=18, AMP) + the loss and optimiser sequence of train.py:292-346 — run UNMODIFIED on top of
either back end (`--backend ref`: the reference's CUDA extensions in oracle/_ref; `--backend dropin`: this repo's
sm_100a `cuda_corr` / `cuda_ba` / `lietorch_backends`).  HARNESS / TEST INFRASTRUCTURE (see tools/refdpvo.py for how the
reference package reaches the GPU box).

Data parallel (new — the reference is single-GPU): under torchrun every rank trains its own synthetic
clip and `parallel.FlatGradAllReducer` averages one flat fp32 gradient buffer over NCCL between `backward()` and
`scaler.unscale_()` (train.py:341-342).

Modes
    grads   one step; dumps loss + flat gradient (npz) for the 1-GPU gradient-parity test
    bench   W warm-up + K timed steps, CUDA events, max over ranks; one JSON line (clips/s)
"""
import argparse
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import refdpvo  # noqa: E402

PKG = refdpvo.PKG


def synth_clip(seed, n_frames=15, ht=480, wd=640, device="cuda"):
    """clip [1,15,3,480,640] float (0..255), camera-to-world poses along a smooth trajectory,
    disparities ~ U(0.2, 1) at full resolution, TartanAir intrinsics."""
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    low = torch.rand(1, n_frames, 3, ht // 8, wd // 8, generator=g)
    images = torch.nn.functional.interpolate(low.view(-1, 3, ht // 8, wd // 8), size=(ht, wd), mode="bilinear",
                                             align_corners=False).view(1, n_frames, 3, ht, wd)
    images = (255.0 * (0.7 * images + 0.3 * torch.rand(1, n_frames, 3, ht, wd, generator=g))).float()
    t = torch.arange(n_frames, dtype=torch.float32)
    poses = torch.zeros(1, n_frames, 7)
    poses[0, :, 0] = 0.05 * t
    poses[0, :, 1] = 0.01 * torch.sin(0.5 * t)
    poses[0, :, 2] = 0.02 * t
    ang = 0.02 * t
    poses[0, :, 4] = torch.sin(ang / 2)           # rotation about y
    poses[0, :, 6] = torch.cos(ang / 2)
    disps = 0.2 + 0.8 * torch.rand(1, n_frames, ht, wd, generator=g)
    intr = torch.tensor([wd / 2.0, wd / 2.0, wd / 2.0, ht / 2.0]).view(1, 1, 4).repeat(1, n_frames, 1)
    return [x.to(device) for x in (images, poses, disps, intr)]


class Trainer:
    """train.py:230-346 without the data loader / logger: model, AdamW, OneCycleLR, GradScaler, and `step`."""

    def __init__(self, seed=1234, amp=True, lr=8e-5, steps=240000, weight_decay=1e-6, clip=10.0, flow_weight=0.1,
                 pose_weight=10.0, STEPS=18, M=1024, reducer=False, structure_only=False):
        import torch
        net_mod = importlib.import_module("dpvo.net")
        self.SE3 = importlib.import_module("dpvo.lietorch").SE3
        self.kabsch_umeyama = importlib.import_module("utils.utils").kabsch_umeyama
        torch.manual_seed(seed)
        import numpy as np
        np.random.seed(seed)
        self.net = net_mod.VONet()
        self.net.train()
        self.net.cuda()
        self.optimizer = torch.optim.AdamW(self.net.parameters(), lr=lr, weight_decay=weight_decay)
        self.scheduler = torch.optim.lr_scheduler.OneCycleLR(self.optimizer, lr, steps, pct_start=0.01, cycle_momentum=False,
                                                             anneal_strategy="linear")
        self.scaler = torch.amp.GradScaler("cuda", enabled=amp)
        self.amp, self.clip, self.fw, self.pw, self.STEPS, self.M, self.so = amp, clip, flow_weight, pose_weight, STEPS, M, structure_only
        self.reducer = None
        if reducer:
            par = importlib.import_module(PKG + ".parallel")
            self.reducer = par.FlatGradAllReducer(self.net.parameters())

    def loss(self, traj):
        """train.py:305-339, verbatim semantics."""
        import torch
        net, so = self.net, self.so
        loss = 0.0
        for i, (v, x, y, P1, P2, kl) in enumerate(traj):
            e = (x - y).norm(dim=-1)
            e = e.reshape(-1, net.P ** 2)[(v > 0.5).reshape(-1)].min(dim=-1).values
            N = P1.shape[1]
            ii, jj = torch.meshgrid(torch.arange(N), torch.arange(N), indexing="ij")
            ii = ii.reshape(-1).cuda()
            jj = jj.reshape(-1).cuda()
            k = ii != jj
            ii = ii[k]
            jj = jj[k]
            P1 = P1.inv()
            P2 = P2.inv()
            t1 = P1.matrix()[..., :3, 3]
            t2 = P2.matrix()[..., :3, 3]
            s = self.kabsch_umeyama(t2[0], t1[0]).detach().clamp(max=10.0)
            P1 = P1.scale(s.view(1, 1))
            dP = P1[:, ii].inv() * P1[:, jj]
            dG = P2[:, ii].inv() * P2[:, jj]
            e1 = (dP * dG.inv()).log()
            tr = e1[..., 0:3].norm(dim=-1)
            ro = e1[..., 3:6].norm(dim=-1)
            loss += self.fw * e.mean()
            if not so and i >= 2:
                loss += self.pw * (tr.mean() + ro.mean())
        loss += kl
        return loss

    def forward_backward(self, clip):
        import torch
        images, poses, disps, intrinsics = clip
        if self.reducer is not None:
            self.reducer.zero_grad()
        else:
            self.optimizer.zero_grad()
        poses = self.SE3(poses).inv()
        with torch.amp.autocast("cuda", enabled=self.amp):
            traj = self.net(images, poses, disps, intrinsics, M=self.M, STEPS=self.STEPS, structure_only=self.so)
            loss = self.loss(traj)
        self.scaler.scale(loss).backward()
        return loss

    def step(self, clip):
        import torch
        loss = self.forward_backward(clip)
        if self.reducer is not None:
            # non-finite values are left in: they make every rank's unscale_ find them and skip together
            self.reducer.allreduce(sanitize=False)
        self.scaler.unscale_(self.optimizer)
        torch.nn.utils.clip_grad_norm_(self.net.parameters(), self.clip)
        self.scaler.step(self.optimizer)
        self.scaler.update()
        self.scheduler.step()
        return loss

    def flat_grad(self):
        import torch
        return torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1).float()
                          for p in self.net.parameters()])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="dropin", choices=["ref", "dropin"])
    ap.add_argument("--mode", default="bench", choices=["grads", "bench"])
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--no-amp", action="store_true")
    ap.add_argument("--update-steps", type=int, default=18)
    ap.add_argument("--structure-only", action="store_true")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()

    import numpy as np
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    files = refdpvo.select_backend(args.backend)
    par = None
    if world > 1:
        sys.path.insert(0, ROOT)
        par = importlib.import_module(PKG + ".parallel")
        par.init(backend="nccl", device=dev)
    tr = Trainer(seed=args.seed, amp=not args.no_amp, STEPS=args.update_steps, reducer=world > 1,
                 structure_only=args.structure_only)
    clip = synth_clip(args.seed + 100 + rank, device=dev)

    if args.mode == "grads":
        loss = tr.forward_backward(clip)
        torch.cuda.synchronize()
        g = tr.flat_grad()
        out = dict(loss=float(loss.item()), grad=g.cpu().numpy(), scale=float(tr.scaler.get_scale()))
        if args.out:
            np.savez_compressed(args.out, **out)
        print(json.dumps({"backend": args.backend, "loss": out["loss"], "grad_norm": float(g.norm().item()),
                          "finite": bool(torch.isfinite(g).all().item()), "numel": int(g.numel()), "modules": files}), flush=True)
        return 0

    losses = []
    for _ in range(args.warmup):
        losses.append(float(tr.step(clip).item()))
    torch.cuda.synchronize()
    if par is not None:
        par.barrier(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        losses.append(float(tr.step(clip).item()))
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    ms = e0.elapsed_time(e1)
    if par is not None:
        ms = par.max_over_ranks(ms, device=dev)
        par.barrier(dev)
    if rank == 0:
        line = {"metric": "DPVO training clips/sec (15 frames 480x640, 80 patches/frame, STEPS=%d, AMP=%s)" % (args.update_steps, not args.no_amp),
                "backend": args.backend, "value": world * args.steps / (ms * 1e-3), "unit": "clips/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "wall_s": wall,
                "scaling": "weak", "collective": None if world == 1 else "NCCL all-reduce (AVG) of one flat fp32 gradient buffer, %d bytes" % tr.reducer.nbytes(),
                "losses": [round(x, 4) if np.isfinite(x) else str(x) for x in losses], "grad_scale": float(tr.scaler.get_scale()),
                "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2), "modules": files}
        if args.out:
            with open(args.out, "w") as f:
                json.dump(line, f)
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
