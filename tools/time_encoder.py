"""Per-frame cost of the two encoders (N3) at 480x640: device path (CUDA graph) against the reference formulation
(torch nn modules, channels-first, autocast fp16, CUDA graph) on the same GPU."""
import importlib
import json
import os
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
enc = importlib.import_module("deep-visual-odometry-slam_b200.encoder")
synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
dev = torch.device("cuda:0")


def torch_encoder(sd, instance):
    p = {k: torch.from_numpy(v).to(dev) for k, v in sd.items()}
    norm = (lambda x: F.instance_norm(x, eps=1e-5)) if instance else (lambda x: x)
    def block(x, name, stride):
        y = F.relu(norm(F.conv2d(x, p[name + ".conv1.weight"], p[name + ".conv1.bias"], stride=stride, padding=1)))
        y = F.relu(norm(F.conv2d(y, p[name + ".conv2.weight"], p[name + ".conv2.bias"], padding=1)))
        if stride != 1:
            x = norm(F.conv2d(x, p[name + ".downsample.0.weight"], p[name + ".downsample.0.bias"], stride=stride))
        return F.relu(x + y)
    def f(img):
        with torch.autocast("cuda", dtype=torch.float16):
            x = F.relu(norm(F.conv2d(img, p["conv1.weight"], p["conv1.bias"], stride=2, padding=3)))
            for name, s in (("layer1.0", 1), ("layer1.1", 1), ("layer2.0", 2), ("layer2.1", 1)):
                x = block(x, name, s)
            return F.conv2d(x, p["conv2.weight"], p["conv2.bias"]) / 4.0
    return f


def graph_time(fn, iters=50):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(); s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fn()
    torch.cuda.synchronize()
    for _ in range(5):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


if __name__ == "__main__":
    img5 = torch.from_numpy(synth.encoder_images(1, 480, 640, 0)).to(dev)
    out = {}
    for name, dim, norm in (("fnet", 128, "instance"), ("inet", 384, "none")):
        sd = synth.encoder_state_dict(dim, 1)
        ours = enc.BasicEncoder4({k: torch.from_numpy(v) for k, v in sd.items()}, norm, output_scale=0.25)
        ref = torch_encoder(sd, norm == "instance")
        a = ours(img5).nchw5[0].float(); b = ref(img5[0]).float()
        rel = float((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt())
        out[name] = dict(ours_ms=graph_time(lambda: ours(img5)), torch_autocast_ms=graph_time(lambda: ref(img5[0])),
                         rms_diff_vs_torch_autocast=rel)
    print(json.dumps(out))
