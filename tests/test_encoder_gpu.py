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


def test_patchifier_matches_reference_gathers():
    """Patchifier (net.py:106-153): with integer centres the three gathers are pure copies of the encoder maps, the
    patch grid is (x + dx, y + dy, 1) and level 2 of the pyramid is the 4x4 average of level 1."""
    enc, synth = load_pkg("encoder"), load_pkg("synth")
    mk = lambda dim, norm, seed: enc.BasicEncoder4({k: torch.from_numpy(v) for k, v in synth.encoder_state_dict(dim, seed).items()},
                                                   norm, output_scale=0.25)
    pf = enc.Patchifier(mk(128, "instance", 1), mk(384, "none", 2))
    images = torch.from_numpy(synth.encoder_images(2, 64, 96, 3)).cuda()
    rng = np.random.default_rng(0)
    h, w, M = 16, 24, 10
    xy = np.stack([rng.integers(1, w - 1, (2, M)), rng.integers(1, h - 1, (2, M))], -1).astype(np.float32)
    fmap, gmap, imap, patches, index, clr = pf(images, patches_per_image=M, return_color=True, coords=torch.from_numpy(xy).cuda())
    assert fmap.shape == (1, 2, 128, h, w) and gmap.shape == (1, 2 * M, 128, 3, 3) and imap.shape == (1, 2 * M, 384, 1, 1)
    assert patches.shape == (1, 2 * M, 3, 3, 3) and index.tolist() == [0] * M + [1] * M and clr.shape == (1, 2 * M, 3)
    f = fmap[0].float().cpu().numpy()
    i_full = pf.inet(images).nchw5[0].float().cpu().numpy()
    g, im, p = gmap[0].float().cpu().numpy(), imap[0].float().cpu().numpy(), patches[0].cpu().numpy()
    for n in range(2):
        for m in range(M):
            x, y = int(xy[n, m, 0]), int(xy[n, m, 1])
            assert np.array_equal(g[n * M + m], f[n, :, y - 1:y + 2, x - 1:x + 2])
            assert np.array_equal(im[n * M + m, :, 0, 0], i_full[n, :, y, x])
            assert np.array_equal(p[n * M + m, 0], np.tile(np.arange(x - 1, x + 2, dtype=np.float32), (3, 1)))
            assert np.array_equal(p[n * M + m, 1], np.tile(np.arange(y - 1, y + 2, dtype=np.float32)[:, None], (1, 3)))
            assert np.all(p[n * M + m, 2] == 1)
    lvl2 = pf.fmap2_nhwc.float().cpu().numpy()
    want = pf.fmap_nhwc.float().cpu().numpy().reshape(2, h // 4, 4, w // 4, 4, 128).mean((2, 4))
    assert lvl2.shape == (2, h // 4, w // 4, 128) and np.abs(lvl2 - want).max() < 2e-3 * max(1.0, np.abs(want).max())
    with pytest.raises(NotImplementedError):
        pf(images, centroid_sel_strat="GRADIENT_BIAS")
    fmap_r, *_ = pf(images, patches_per_image=7)            # the reference's RANDOM strategy
    assert fmap_r.shape == fmap.shape


def test_image_fed_hot_path_writes_the_encoders_output_into_the_ring():
    """hotpath.ImageFrontHotPath: the ring slot of the newest frame holds fnet(image) / 4 and its 4x4 average,
    the gathered patches come from the same maps, and the frame still runs end to end (stream launches and graph)."""
    hotpath = load_pkg("hotpath")
    hp = hotpath.ImageFrontHotPath(torch.device("cuda", 0), n_frames=12, seed=1)
    hp.step()
    torch.cuda.synchronize()
    slot = (hp.n - 1) % hp.mem
    assert torch.equal(hp.fmap1[0, slot].permute(1, 2, 0), hp._f.nhwc[0])
    assert torch.equal(hp.fmap2[0, slot].permute(1, 2, 0), hp._f2[0])
    g, i = hp._patches_new
    xy = hp.centres[0].long()
    x, y = int(xy[5, 0]), int(xy[5, 1])
    assert torch.equal(g[0, 5], hp._f.nchw5[0, 0, :, y - 1:y + 3, x - 1:x + 3])
    assert torch.equal(i[0, 5, :, 0, 0], hp._i.nchw5[0, 0, :, y, x])
    poses = hp.poses.clone()
    hp.capture()
    hp.replay()
    torch.cuda.synchronize()
    assert torch.isfinite(hp.poses).all() and torch.allclose(hp.poses, poses, atol=1e-5)
