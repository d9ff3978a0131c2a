"""Per-frame cost of the edge bookkeeping (N4) at the default window: M = 96 patches, PATCH_LIFETIME 13,
REMOVAL_WINDOW 22 (E ~ 45k edges, 384-wide fp16 net rows).  Device edge list (libdpvo_b200.so) against the
reference's formulation (torch.cat + boolean-mask indexing, dpvo.py:215-239/280-310) as torch ops on the
same GPU.  Each frame: edges_forw + edges_back, keyframe removal of frame n-4 with renumbering, window removal."""
import importlib
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pg = importlib.import_module("deep-visual-odometry-slam_b200.patchgraph")

M, R, RW, DIM, N = 96, 13, 22, 384, 4096
dev = torch.device("cuda:0")
ix = torch.arange(N, device=dev).repeat_interleave(M).contiguous()


def ours(frames):
    e = pg.PatchGraphEdges(ix, M, 60000, inactive_capacity=frames * 4000 + 200000, net_dim=DIM)
    n = 0
    for _ in range(30):                      # fill the window
        n += 1
        e.edges_forw(n, R); e.edges_back(n, R); e.remove_old(n, RW)
    torch.cuda.synchronize()
    t0, ev0, ev1 = time.perf_counter(), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(frames):
        n += 1
        e.edges_forw(n, R); e.edges_back(n, R)
        e.remove_frame(n - 4)
        n -= 1
        e.remove_old(n, RW)
    ev1.record()
    host = (time.perf_counter() - t0) / frames
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / frames
    e.check_info()
    return dict(edges=e.active.count(), gpu_ms=ev0.elapsed_time(ev1) / frames, host_issue_ms=host * 1e3, wall_ms=wall * 1e3)


def torch_formulation(frames):
    ii = torch.zeros(0, dtype=torch.long, device=dev); jj = ii.clone(); kk = ii.clone()
    net = torch.zeros(1, 0, DIM, dtype=torch.half, device=dev)
    st = dict(ii=ii, jj=jj, kk=kk, net=net, w=torch.zeros(1, 0, 2, device=dev), t=torch.zeros(1, 0, 2, device=dev),
              ii_i=ii.clone(), jj_i=ii.clone(), kk_i=ii.clone(), w_i=torch.zeros(1, 0, 2, device=dev), t_i=torch.zeros(1, 0, 2, device=dev))

    def append(a, b):
        st["jj"] = torch.cat([st["jj"], b]); st["kk"] = torch.cat([st["kk"], a]); st["ii"] = torch.cat([st["ii"], ix[a]])
        st["net"] = torch.cat([st["net"], torch.zeros(1, len(a), DIM, dtype=torch.half, device=dev)], dim=1)
        z = torch.zeros(1, len(a), 2, device=dev)
        st["w"] = torch.cat([st["w"], z], 1); st["t"] = torch.cat([st["t"], z], 1)

    def remove(m, store):
        if store:
            st["ii_i"] = torch.cat((st["ii_i"], st["ii"][m])); st["jj_i"] = torch.cat((st["jj_i"], st["jj"][m]))
            st["kk_i"] = torch.cat((st["kk_i"], st["kk"][m]))
            st["w_i"] = torch.cat((st["w_i"], st["w"][:, m]), 1); st["t_i"] = torch.cat((st["t_i"], st["t"][:, m]), 1)
        st["w"] = st["w"][:, ~m]; st["t"] = st["t"][:, ~m]
        st["ii"] = st["ii"][~m]; st["jj"] = st["jj"][~m]; st["kk"] = st["kk"][~m]; st["net"] = st["net"][:, ~m]

    def grid(a0, a1, b0, b1):
        a, b = torch.meshgrid(torch.arange(a0, a1, device=dev), torch.arange(b0, b1, device=dev), indexing="ij")
        return a.reshape(-1), b.reshape(-1)

    def frame(n, key):
        append(*grid(M * max(n - R, 0), M * max(n - 1, 0), n - 1, n))
        append(*grid(M * max(n - 1, 0), M * n, max(n - R, 0), n))
        if key:
            k = n - 4
            remove((st["ii"] == k) | (st["jj"] == k), False)
            st["kk"][st["ii"] > k] -= M; st["ii"][st["ii"] > k] -= 1; st["jj"][st["jj"] > k] -= 1
            n -= 1
        remove(ix[st["kk"]] < n - RW, True)
        return n

    n = 0
    for _ in range(30):
        n = frame(n + 1, False)
    torch.cuda.synchronize()
    t0, ev0, ev1 = time.perf_counter(), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(frames):
        n = frame(n + 1, True)
    ev1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / frames
    return dict(edges=int(st["ii"].numel()), gpu_ms=ev0.elapsed_time(ev1) / frames, wall_ms=wall * 1e3)


if __name__ == "__main__":
    frames = 50
    ours(5); torch_formulation(5)
    out = dict(frames=frames, M=M, patch_lifetime=R, removal_window=RW, ours=ours(frames), torch_formulation=torch_formulation(frames))
    print(json.dumps(out))
