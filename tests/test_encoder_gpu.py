"""GPU parity for the per-frame encoders (N3) through the C ABI (dpvo_encoder_*): BasicEncoder4 as fnet (instance
norm, 128) and inet (no norm, 384) against the reference module's own float32 outputs (tests/golden/ref_cpu_encoder.npz)
and, at other sizes, against oracle/encoder_oracle.py.  The device path stores fp16 activations between layers with fp32
accumulation (the reference runs these layers under autocast in fp16 too, dpvo.py:391); tolerance, written here: the
RMS error is below 1 % of the map's RMS and no element is off by more than 5 % of the map's RMS... measured values are
printed with -s."""
import os

import numpy as np
import pytest
import torch

from conftest import load_pkg
from test_oracle_cpu import _encoder_golden

pytestmark = pytest.mark.gpu


def _err(got, want):
    rms = float(np.sqrt(np.mean(want.astype(np.float64) ** 2)))
    d = got.astype(np.float64) - want
    return float(np.sqrt(np.mean(d ** 2))) / rms, float(np.abs(d).max()) / rms


def _run(name, out_dim, norm_fn, seed, images, scale=1.0):
    enc, synth = load_pkg("encoder"), load_pkg("synth")
    sd = synth.encoder_state_dict(out_dim, seed)
    e = enc.BasicEncoder4({k: torch.from_numpy(v) for k, v in sd.items()}, norm_fn, output_scale=scale)
    fm = e(torch.from_numpy(images).cuda())
    torch.cuda.synchronize()
    return sd, fm


@pytest.mark.parametrize("name", ["fnet", "inet"])
def test_encoder_vs_reference_golden(name):
    synth = load_pkg("synth")
    m, g = _encoder_golden()
    c = m.CASES[name]
    images = synth.encoder_images(m.N_IMG, m.H, m.W, m.IMG_SEED)
    _, fm = _run(name, c["out_dim"], c["norm_fn"], c["seed"], images)
    got = fm.nchw5[0].float().cpu().numpy()
    assert got.shape == g[name].shape and fm.nhwc.is_contiguous()
    rms, mx = _err(got, g[name])
    print(f"{name}: rms err {rms:.2e}, max err {mx:.2e} (relative to the map's rms)")
    assert rms < 1e-2 and mx < 5e-2


@pytest.mark.parametrize("n,H,W,norm_fn,out_dim", [(1, 480, 640, "instance", 128), (1, 480, 640, "none", 384),
                                                   (3, 50, 70, "instance", 128), (1, 33, 47, "none", 64)])
def test_encoder_vs_oracle_sizes(n, H, W, norm_fn, out_dim):
    """Full 480x640 frames and sizes that are not multiples of the tiles / strides; the '/ 4' folded into the last layer."""
    from oracle import encoder_oracle as eo
    synth = load_pkg("synth")
    images = synth.encoder_images(n, H, W, n + H)
    sd, fm = _run("x", out_dim, norm_fn, 7, images, scale=0.25)
    want = eo.basic_encoder4(sd, images, norm_fn)[0] / 4.0
    got = fm.nchw5[0].float().cpu().numpy()
    assert got.shape == want.shape
    rms, mx = _err(got, want)
    print(f"{n}x{H}x{W} {norm_fn}: rms err {rms:.2e}, max err {mx:.2e}")
    assert rms < 1e-2 and mx < 6e-2


def test_avg_pool4_and_errors():
    from oracle import encoder_oracle as eo
    enc = load_pkg("encoder")
    rng = np.random.default_rng(0)
    x = rng.standard_normal((2, 21, 30, 128)).astype(np.float16)
    got = enc.avg_pool4(torch.from_numpy(x).cuda()).float().cpu().numpy()
    want = eo.avg_pool4(x.astype(np.float32).transpose(0, 3, 1, 2)).transpose(0, 2, 3, 1)
    assert got.shape == (2, 5, 7, 128) and np.abs(got - want).max() < 2e-3
    synth = load_pkg("synth")
    sd = {k: torch.from_numpy(v) for k, v in synth.encoder_state_dict(128, 0).items()}
    with pytest.raises(RuntimeError):
        enc.BasicEncoder4(sd, "batch")
    bad = dict(sd); bad.pop("layer2.0.downsample.0.bias")
    with pytest.raises(RuntimeError):
        enc.BasicEncoder4(bad, "none")
    e = enc.BasicEncoder4(sd, "none")
    with pytest.raises(RuntimeError):
        e(torch.zeros(1, 1, 3, 16, 16))                       # CPU tensor
    with pytest.raises(RuntimeError):
        e(torch.zeros(1, 3, 16, 16, device="cuda"))
