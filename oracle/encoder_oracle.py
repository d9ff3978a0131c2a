"""TEST INFRASTRUCTURE ONLY (never imported by the product path).

float32 restatement of BasicEncoder4 / ResidualBlock (methods/dpvo/dpvo/extractor.py:6-55, 200-264) with
torch.nn.functional on CPU, from the reference state-dict keys.  Pinned: tests/test_oracle_cpu.py compares it with
tests/golden/ref_cpu_encoder.npz, produced by the reference module itself (tests/golden/make_golden_encoder.py).
"""
import torch
import torch.nn.functional as F


def _t(x):
    return torch.as_tensor(x, dtype=torch.float32)


def _norm(x, instance):
    return F.instance_norm(x, eps=1e-5) if instance else x          # nn.InstanceNorm2d(planes): affine=False


def _block(sd, name, x, stride, instance):
    """ResidualBlock.forward, extractor.py:47-55."""
    y = F.relu(_norm(F.conv2d(x, _t(sd[name + ".conv1.weight"]), _t(sd[name + ".conv1.bias"]), stride=stride, padding=1), instance))
    y = F.relu(_norm(F.conv2d(y, _t(sd[name + ".conv2.weight"]), _t(sd[name + ".conv2.bias"]), padding=1), instance))
    if stride != 1:
        x = _norm(F.conv2d(x, _t(sd[name + ".downsample.0.weight"]), _t(sd[name + ".downsample.0.bias"]), stride=stride), instance)
    return F.relu(x + y)


def basic_encoder4(sd, images, norm_fn):
    """BasicEncoder4.forward, extractor.py:246-264.  images [b, n, 3, H, W] -> [b, n, C, H/4, W/4] (numpy float32)."""
    instance = norm_fn == "instance"
    images = _t(images)
    b, n, c, h, w = images.shape
    x = images.reshape(b * n, c, h, w)
    x = F.relu(_norm(F.conv2d(x, _t(sd["conv1.weight"]), _t(sd["conv1.bias"]), stride=2, padding=3), instance))
    x = _block(sd, "layer1.0", x, 1, instance)
    x = _block(sd, "layer1.1", x, 1, instance)
    x = _block(sd, "layer2.0", x, 2, instance)
    x = _block(sd, "layer2.1", x, 1, instance)
    x = F.conv2d(x, _t(sd["conv2.weight"]), _t(sd["conv2.bias"]))
    return x.reshape(b, n, *x.shape[1:]).numpy()


def avg_pool4(x):
    """F.avg_pool2d(fmap[0], 4, 4), dpvo.py:438; x [n, C, H, W]."""
    return F.avg_pool2d(_t(x), 4, 4).numpy()
