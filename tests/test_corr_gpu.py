"""GPU parity tests for the correlation / patchify operators (call path: Python
operator surface -> ctypes -> C ABI -> sm_100a kernels), checked against the CPU
oracle and, when oracle/_ref/cuda_corr.so is present, against the reference's own
CUDA extension on identical inputs.

Tolerance (north star): correlation within 1e-3 relative in fp16, judged norm-wise
(||a-b|| / ||b||) because element-wise relative error is meaningless at zero
crossings; plus a per-element bound of a few fp16 ulps of the output scale.
"""
import numpy as np
import pytest
import torch

from conftest import load_pkg
from oracle import altcorr_oracle as oc
from oracle.ref_ext import load_ref

pytestmark = pytest.mark.gpu

REL_TOL_F16 = 1e-3


def _rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def _dev(x):
    return torch.from_numpy(x).cuda()


@pytest.fixture(scope="module")
def small():
    synth = load_pkg("synth")
    return synth.corr_inputs(600, seed=3, mem=4, n_patches=64, oob_frac=0.1)


def _nhwc(t):
    """Same logical [B,N,C,H,W] tensor, channels-last memory."""
    return t.permute(0, 1, 3, 4, 2).contiguous().permute(0, 1, 4, 2, 3)


@pytest.mark.parametrize("layout", ["nchw", "nhwc"])
@pytest.mark.parametrize("level", [0, 1])
def test_forward_f16_vs_oracle(small, layout, level):
    pkg = load_pkg()
    d = small
    fm_np = d["fmap1"] if level == 0 else d["fmap2"]
    scale = np.float32(1.0 if level == 0 else 0.25)
    coords = (d["coords"] * scale).astype(np.float32)
    want = oc.corr_forward(d["gmap"], fm_np, coords, d["us"], d["vs"], 3)
    fm = _dev(fm_np)
    if layout == "nhwc":
        fm = _nhwc(fm)
    got, = pkg.cuda_corr.forward(_dev(d["gmap"]), fm, _dev(coords), _dev(d["us"]), _dev(d["vs"]), 3)
    assert got.shape == want.shape and got.dtype == torch.float16
    got = got.cpu().numpy()
    assert _rel(got, want) < REL_TOL_F16
    scale_out = np.abs(want.astype(np.float32)).max()
    assert np.abs(got.astype(np.float32) - want.astype(np.float32)).max() <= 4e-3 * scale_out
    # exactness: the overwhelming majority of outputs are bit-identical to the restatement
    assert (got == want).mean() > 0.90


def test_pyramid_matches_stacked_levels(small):
    pkg = load_pkg()
    d = small
    g = _dev(d["gmap"]); f1 = _nhwc(_dev(d["fmap1"])); f2 = _nhwc(_dev(d["fmap2"]))
    c = _dev(d["coords"]); us = _dev(d["us"]); vs = _dev(d["vs"])
    c1, = pkg.cuda_corr.forward(g, f1, c, us, vs, 3)
    c2, = pkg.cuda_corr.forward(g, f2, c / 4, us, vs, 3)
    want = torch.stack([c1, c2], -1).view(1, len(us), -1)
    got = pkg.altcorr.corr_pyramid(g, (f1, f2), c, us, vs, 3)
    assert got.shape == (1, len(us), 882)
    assert torch.equal(got, want)
    want_np = oc.corr_pyramid(d["gmap"], [d["fmap1"], d["fmap2"]], [1.0, 0.25], d["coords"], d["us"], d["vs"], 3)
    assert _rel(got.cpu().numpy(), want_np) < REL_TOL_F16


@pytest.mark.parametrize("dtype", [np.float32, np.float64, np.float16])
def test_forward_generic_path(dtype):
    """Odd sizes force the generic kernel: C=24, 2x2 patches, radius 2, batch 2."""
    pkg = load_pkg()
    rng = np.random.default_rng(11)
    B, N1, N2, C, PH, PW, H2, W2, M, R = 2, 7, 3, 24, 2, 2, 17, 23, 50, 2
    f1 = rng.standard_normal((B, N1, C, PH, PW)).astype(dtype)
    f2 = rng.standard_normal((B, N2, C, H2, W2)).astype(dtype)
    coords = np.stack([rng.uniform(-4, W2 + 3, (B, M, PH, PW)), rng.uniform(-4, H2 + 3, (B, M, PH, PW))], 2).astype(np.float32)
    us = rng.integers(0, N1, M).astype(np.int64)
    vs = rng.integers(0, N2, M).astype(np.int64)
    want = oc.corr_forward(f1, f2, coords, us, vs, R)
    got, = pkg.cuda_corr.forward(_dev(f1), _dev(f2), _dev(coords), _dev(us), _dev(vs), R)
    tol = {np.float32: 2e-6, np.float64: 2e-6, np.float16: REL_TOL_F16}[dtype]
    assert got.shape == want.shape
    assert _rel(got.cpu().numpy(), want) < tol


def test_forward_edge_cases(small):
    pkg = load_pkg()
    d = small
    g = _dev(d["gmap"]); f1 = _nhwc(_dev(d["fmap1"]))
    # empty edge list
    e = torch.zeros(0, dtype=torch.int64, device="cuda")
    out, = pkg.cuda_corr.forward(g, f1, torch.zeros(1, 0, 2, 3, 3, device="cuda"), e, e, 3)
    assert out.shape == (1, 0, 7, 7, 3, 3)
    # spread-out patch pixels (no compact bounding box) + everything out of bounds + fractional edges
    rng = np.random.default_rng(5)
    M = 64
    coords = np.zeros((1, M, 2, 3, 3), np.float32)
    coords[0, :, 0] = rng.uniform(-30, 190, (M, 3, 3))
    coords[0, :, 1] = rng.uniform(-30, 150, (M, 3, 3))
    coords[0, :8] = 1e6
    coords[0, 8:16] = -1e6
    coords[0, 16, 0] = 159.5  # right border: half the window leaves the map
    coords[0, 17, 1] = -0.5
    us = rng.integers(0, 64, M).astype(np.int64)
    vs = rng.integers(0, 4, M).astype(np.int64)
    want = oc.corr_forward(d["gmap"], d["fmap1"], coords, us, vs, 3)
    got, = pkg.cuda_corr.forward(g, f1, _dev(coords), _dev(us), _dev(vs), 3)
    got = got.cpu().numpy()
    assert np.all(got[0, :16] == 0)
    assert _rel(got, want) < REL_TOL_F16


def test_shadow_cache_tracks_inplace_writes(small):
    """NCHW ring (reference layout): the channels-last shadow must follow in-place slot writes."""
    pkg = load_pkg()
    d = small
    g = _dev(d["gmap"]); fm = _dev(d["fmap1"]).clone()
    c = _dev(d["coords"]); us = _dev(d["us"]); vs = _dev(d["vs"])
    a, = pkg.cuda_corr.forward(g, fm, c, us, vs, 3)
    n0 = pkg.cuda_corr.shadow_refreshes
    b, = pkg.cuda_corr.forward(g, fm, c, us, vs, 3)
    assert pkg.cuda_corr.shadow_refreshes == n0 and torch.equal(a, b)
    fm[0, 1] = fm[0, 2]  # ring slot shift, as dpvo.py:296
    c2, = pkg.cuda_corr.forward(g, fm, c, us, vs, 3)
    assert pkg.cuda_corr.shadow_refreshes == n0 + 1
    ref, = pkg.cuda_corr.forward(g, _nhwc(fm), c, us, vs, 3)
    assert torch.equal(c2, ref)


def test_forward_vs_reference_extension(small):
    """Same inputs through the reference's own CUDA kernel (built from untouched sources)."""
    ref = load_ref("cuda_corr")
    if ref is None:
        pytest.skip("oracle/_ref/cuda_corr.so not built")
    pkg = load_pkg()
    d = small
    for dtype in (torch.float16, torch.float32):
        g = _dev(d["gmap"]).to(dtype); fm = _dev(d["fmap1"]).to(dtype)
        c = _dev(d["coords"]); us = _dev(d["us"]); vs = _dev(d["vs"])
        want, = ref.forward(g, fm, c, us, vs, 3)
        torch.cuda.synchronize()
        got, = pkg.cuda_corr.forward(g, fm, c, us, vs, 3)
        assert got.shape == want.shape
        tol = REL_TOL_F16 if dtype == torch.float16 else 2e-6
        assert _rel(got.float().cpu().numpy(), want.float().cpu().numpy()) < tol
        # and the oracle agrees with the reference kernel (pins the restatement)
        if dtype == torch.float16:
            w2 = oc.corr_forward(d["gmap"], d["fmap1"], d["coords"], d["us"], d["vs"], 3)
            assert _rel(w2, want.cpu().numpy()) < REL_TOL_F16
            assert (w2 == want.cpu().numpy()).mean() > 0.90


@pytest.mark.parametrize("dtype", [np.float16, np.float32])
@pytest.mark.parametrize("radius,C", [(0, 384), (1, 128), (1, 3)])
def test_patchify_forward(dtype, radius, C):
    pkg = load_pkg()
    rng = np.random.default_rng(radius * 10 + C)
    B, H, W, M = 1, 30, 40, 96
    net = rng.standard_normal((B, C, H, W)).astype(dtype)
    coords = np.stack([rng.uniform(-2, W + 1, (B, M)), rng.uniform(-2, H + 1, (B, M))], -1).astype(np.float32)
    want = oc.patchify_forward(net, coords, radius)
    got, = pkg.cuda_corr.patchify_forward(_dev(net), _dev(coords), radius)
    assert np.array_equal(got.cpu().numpy(), want)
    # full operator with the bilinear blend (torch-level, as the reference)
    got2 = pkg.altcorr.patchify(_dev(net), _dev(coords), radius)
    want2 = oc.patchify(net, coords, radius)
    assert _rel(got2.cpu().numpy(), want2) < (REL_TOL_F16 if dtype == np.float16 else 1e-6)
    ref = load_ref("cuda_corr")
    if ref is not None:
        w3, = ref.patchify_forward(_dev(net), _dev(coords), radius)
        assert torch.equal(got, w3)


def test_patchify_backward():
    pkg = load_pkg()
    rng = np.random.default_rng(2)
    B, C, H, W, M, R = 1, 16, 12, 14, 40, 1
    net = rng.standard_normal((B, C, H, W)).astype(np.float32)
    coords = np.stack([rng.uniform(-2, W + 1, (B, M)), rng.uniform(-2, H + 1, (B, M))], -1).astype(np.float32)
    grad = rng.standard_normal((B, M, C, 2 * R + 2, 2 * R + 2)).astype(np.float32)
    want = oc.patchify_backward(net.shape, coords, grad, R)
    got, = pkg.cuda_corr.patchify_backward(_dev(net), _dev(coords), _dev(grad), R)
    assert _rel(got.cpu().numpy(), want) < 1e-6
    # autograd path
    t = _dev(net).requires_grad_(True)
    out = pkg.altcorr.patchify(t, _dev(coords), R, mode="nearest")
    out.backward(_dev(grad))
    assert _rel(t.grad.cpu().numpy(), want) < 1e-6


def test_full_size_properties():
    """BASELINE.json full size (45,312 edges, 36-frame rings): size-independent properties.
    (1) linearity in the patch features: corr(2*g) == 2*corr(g) in fp16 (power of two scaling);
    (2) an edge's result does not depend on its position in the edge list (permutation);
    (3) spot-check 128 random edges against the oracle."""
    pkg = load_pkg()
    synth = load_pkg("synth")
    edges = synth.dpvo_edges(40)
    d = synth.corr_inputs(0, seed=9, edges=edges)
    E = len(d["us"])
    assert E == 45312
    g = _dev(d["gmap"]); f1 = _nhwc(_dev(d["fmap1"])); f2 = _nhwc(_dev(d["fmap2"]))
    c = _dev(d["coords"]); us = _dev(d["us"]); vs = _dev(d["vs"])
    out = pkg.altcorr.corr_pyramid(g, (f1, f2), c, us, vs, 3)
    out2 = pkg.altcorr.corr_pyramid(g * 2, (f1, f2), c, us, vs, 3)
    fin = torch.isfinite(out2)
    # exact except where a blend product falls into fp16's subnormal range (spacing 2^-24): that
    # perturbation can flip a round-to-nearest tie of the running blend sum, i.e. at most one fp16
    # ulp of the result, and only on a vanishing fraction of outputs
    a = (out.float() * 2)[fin]; b = out2.float()[fin]
    assert ((a - b).abs() <= b.abs() * 2.0 ** -10 + 2.0 ** -23).all()
    assert (a != b).float().mean().item() < 1e-3
    perm = torch.randperm(E, device="cuda", generator=torch.Generator("cuda").manual_seed(0))
    outp = pkg.altcorr.corr_pyramid(g, (f1, f2), c[:, perm].contiguous(), us[perm], vs[perm], 3)
    assert torch.equal(outp, out[:, perm])
    sel = np.random.default_rng(0).choice(E, 128, replace=False)
    want = oc.corr_pyramid(d["gmap"], [d["fmap1"], d["fmap2"]], [1.0, 0.25], d["coords"][:, sel], d["us"][sel], d["vs"][sel], 3)
    assert _rel(out[:, torch.from_numpy(sel).cuda()].cpu().numpy(), want) < REL_TOL_F16


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
def test_corr_backward_vs_oracle_and_reference(dtype):
    """cuda_corr.backward (correlation_kernel.cu:140-191,237-287) and the autograd path of altcorr.corr.
    fp32: 1e-5 norm-wise against the float64-accumulating restatement and the reference extension.
    fp16: ours accumulates in fp32 and rounds once; the reference accumulates with fp16 atomics, so it
    is compared at the looser fp16-accumulation level."""
    pkg = load_pkg()
    rng = np.random.default_rng(21)
    B, N1, N2, C, P, H2, W2, M, R = 1, 12, 3, 128, 3, 30, 40, 96, 3
    f1 = rng.standard_normal((B, N1, C, P, P)).astype(dtype)
    f2 = rng.standard_normal((B, N2, C, H2, W2)).astype(dtype)
    cx = rng.uniform(-2, W2 + 1, (B, M, 1, 1)); cy = rng.uniform(-2, H2 + 1, (B, M, 1, 1))
    gy, gx = np.meshgrid(np.arange(P) - 1.0, np.arange(P) - 1.0, indexing="ij")
    coords = np.stack([cx + gx + rng.normal(0, 0.2, (B, M, P, P)), cy + gy + rng.normal(0, 0.2, (B, M, P, P))], 2).astype(np.float32)
    coords[0, :4] *= 7.0      # spread-out patches: no compact bounding box
    us = rng.integers(0, N1, M).astype(np.int64); vs = rng.integers(0, N2, M).astype(np.int64)
    grad = rng.standard_normal((B, M, 2 * R + 1, 2 * R + 1, P, P)).astype(dtype)
    w1, w2 = oc.corr_backward(f1, f2, coords, us, vs, grad, R)
    g1, g2 = pkg.cuda_corr.backward(_dev(f1), _dev(f2), _dev(coords), _dev(us), _dev(vs), _dev(grad), R)
    assert g1.shape == f1.shape and g2.shape == f2.shape and g1.dtype == _dev(f1).dtype
    tol = 1e-5 if dtype == np.float32 else 1e-3
    assert _rel(g1.float().cpu().numpy(), w1) < tol
    assert _rel(g2.float().cpu().numpy(), w2) < tol
    # non-contiguous (permuted) upstream gradient, as autograd hands it over
    gperm = _dev(np.ascontiguousarray(grad.transpose(0, 1, 3, 2, 4, 5))).permute(0, 1, 3, 2, 4, 5)
    h1, h2 = pkg.cuda_corr.backward(_dev(f1), _dev(f2), _dev(coords), _dev(us), _dev(vs), gperm, R)
    assert _rel(h1.float().cpu().numpy(), w1) < tol and _rel(h2.float().cpu().numpy(), w2) < tol
    ref = load_ref("cuda_corr")
    if ref is not None:
        r1, r2 = ref.backward(_dev(f1), _dev(f2), _dev(coords), _dev(us), _dev(vs), _dev(grad), R)
        rt = 1e-5 if dtype == np.float32 else 2e-2
        assert _rel(g1.float().cpu().numpy(), r1.float().cpu().numpy()) < rt
        assert _rel(g2.float().cpu().numpy(), r2.float().cpu().numpy()) < rt
    # autograd: d/dfmap of sum(corr * grad)
    if dtype == np.float32:
        t1 = _dev(f1).requires_grad_(True); t2 = _dev(f2).requires_grad_(True)
        out = pkg.altcorr.corr(t1, t2, _dev(coords), _dev(us), _dev(vs), R)
        (out * _dev(grad)).sum().backward()
        assert _rel(t1.grad.cpu().numpy(), w1) < tol and _rel(t2.grad.cpu().numpy(), w2) < tol


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(3, 128, 30, 40), (2, 128, 7, 9), (1, 130, 6, 10), (2, 96, 5, 13), (1, 2, 1, 2)])
def test_nchw_to_nhwc_bit_exact(dtype, shape):
    """Layout kernel (ring write / channels-last shadow): a pure permutation, so bit-exact against
    torch's permute for 16- and 32-bit elements, even and odd plane sizes, channel tails."""
    pkg = load_pkg()
    lib = pkg._lib
    N, C, H, W = shape
    g = torch.Generator(device="cpu").manual_seed(N * 1000 + C)
    src = torch.randn(shape, generator=g).to(dtype).cuda()
    dst = torch.full((N, H, W, C), float("nan"), dtype=dtype, device="cuda")
    lib.check(lib.lib.dpvo_nchw_to_nhwc(lib.c_i32(lib.dtype_code(src)), lib.ptr(src), lib.ptr(dst), lib.c_i64(N),
                                        lib.c_i32(C), lib.c_i64(H * W), lib.stream_of(src)), "nchw_to_nhwc")
    want = src.permute(0, 2, 3, 1).contiguous()
    iv = {2: torch.int16, 4: torch.int32}[src.element_size()]
    assert torch.equal(dst.view(iv), want.view(iv))


def test_ring_sync_refreshes_only_changed_slots():
    """The NCHW-ring shadow (cuda_corr._channels_last_view): after the first full sync, writing two slots the way
    the reference does (dpvo.py:435-438: `fmap1_[:, n % mem] = ...`) must refresh exactly those; a raw write the
    version counter does not see needs `invalidate`."""
    pkg = load_pkg()
    cc = pkg.cuda_corr
    torch.manual_seed(5)
    ring = torch.randn(1, 9, 128, 30, 40, device="cuda").half()
    want = lambda: ring.permute(0, 1, 3, 4, 2).contiguous()
    s0 = cc.shadow_stats()["syncs"]
    v = cc._channels_last_view(ring)
    assert v.shape == ring.shape and v.stride(2) == 1
    assert torch.equal(v.permute(0, 1, 3, 4, 2).contiguous(), want())
    assert cc._channels_last_view(ring) is v and cc.shadow_stats()["syncs"] == s0 + 1      # unchanged: cache hit
    # poison the shadow of an UNCHANGED slot: a selective sync must not touch it
    v[0, 3].fill_(7.0)
    ring[:, 2] = torch.randn(128, 30, 40, device="cuda").half()
    ring[0, 7] = ring[0, 8]
    v2 = cc._channels_last_view(ring)
    assert v2 is v and cc.shadow_stats()["syncs"] == s0 + 2
    got = v.permute(0, 1, 3, 4, 2).contiguous()
    w = want()
    for slot in (0, 1, 2, 4, 5, 6, 7, 8):
        assert torch.equal(got[0, slot], w[0, slot]), slot
    assert bool((got[0, 3] == 7.0).all())                       # untouched because its source did not change
    # a write torch's version counter does not see
    ring.data[0, 3, 0, 0, 0] += 1.0
    cc.invalidate(ring)
    v3 = cc._channels_last_view(ring)
    assert torch.equal(v3.permute(0, 1, 3, 4, 2).contiguous(), want())


def test_corr_backward_fast_path_matches_generic_kernel_at_training_shape():
    """Config 4's shape (15-frame clip, 80 patches/frame, fp16, level 1 = 120x160): the warp-per-edge kernel over the
    channels-last ring against this repo's generic one-CTA-per-edge kernel (both fp32-accumulating, independent code)."""
    import os
    pkg = load_pkg()
    synth = load_pkg("synth")
    E = 3600
    d = synth.corr_inputs(E, seed=4, mem=15, n_patches=15 * 80, oob_frac=0.03)
    g = _dev(d["gmap"]).half(); f = _dev(d["fmap1"]).half(); c = _dev(d["coords"]); us = _dev(d["us"]); vs = _dev(d["vs"])
    grad = torch.randn(1, E, 7, 7, 3, 3, device="cuda", generator=torch.Generator("cuda").manual_seed(1)).half()
    a1, a2 = pkg.cuda_corr.backward(g, f, c, us, vs, grad, 3)
    assert a2.shape == f.shape and a2.stride(2) == 1             # channels-last view
    os.environ["DPVO_CORR_BWD_GENERIC"] = "1"
    try:
        b1, b2 = pkg.cuda_corr.backward(g, f, c, us, vs, grad, 3)
    finally:
        del os.environ["DPVO_CORR_BWD_GENERIC"]
    assert _rel(a1.float().cpu().numpy(), b1.float().cpu().numpy()) < 2e-4
    assert _rel(a2.float().cpu().numpy(), b2.float().cpu().numpy()) < 2e-4
