"""CPU restatement (numpy) of the reference's altcorr operators.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module, and only as the checker.  The product path never does.

Parity status: the reference ships no golden vectors or known-answer tests for
correlation (correlation_test.py only checks shapes / no-throw).  This restatement
is pinned against outputs of the reference's own CUDA extension built from the
untouched sources (oracle/build_ref.py -> oracle/_ref/cuda_corr.so) executed on the
GPU box on identical seeded inputs (tests/test_corr_gpu.py::test_*_vs_reference_ext),
and against fixtures generated from it (tests/golden/).

Follows (paths under /root/reference/methods/dpvo/dpvo/altcorr/):
  corr_forward       correlation_kernel.cu:83-137  (window indexing, floor, bounds -> 0,
                                                    fp32 accumulate, cast to feature dtype)
                     correlation_kernel.cu:222-233 (bilinear blend in the feature dtype,
                                                    one rounding per op, final permute)
  corr_backward      correlation_kernel.cu:140-191,237-287
  patchify_forward   correlation_kernel.cu:17-48,289-308
  patchify_backward  correlation_kernel.cu:50-81,311-334
  patchify (blend)   correlation.py:55-74
"""
import numpy as np


def _floor_int(v):
    """static_cast<int>(floor(v)) with CUDA semantics (NaN -> 0, saturating), clamped to +-2^24."""
    f = np.floor(v.astype(np.float32))
    f = np.where(np.isnan(f), 0.0, f)
    return np.clip(f, -(1 << 24), (1 << 24)).astype(np.int64)


def _blend_weights(coords_scaled, dtype):
    """dx,dy -> the four blend weights, each op rounded in `dtype` (correlation_kernel.cu:224-231)."""
    x = coords_scaled[:, :, 0].astype(np.float32)
    y = coords_scaled[:, :, 1].astype(np.float32)
    with np.errstate(invalid="ignore"):
        dx = (x - np.floor(x)).astype(dtype)
        dy = (y - np.floor(y)).astype(dtype)
    one = dtype(1)
    mdx = (one - dx).astype(dtype)
    mdy = (one - dy).astype(dtype)
    w00 = (mdx * mdy).astype(dtype)
    w01 = (dx * mdy).astype(dtype)
    w10 = (mdx * dy).astype(dtype)
    w11 = (dx * dy).astype(dtype)
    return w00, w01, w10, w11  # each [B,M,P,P]


def corr_raw(fmap1, fmap2, coords, ii, jj, radius, chunk=256, _f64=None):
    """Raw window volume [B,M,D,D,P,P] (rows=y offset, cols=x offset) in fmap dtype.
    `_f64`: optional (fmap1, fmap2) already converted to float64 (callers that loop over edge chunks)."""
    B, N1, C, PH, PW = fmap1.shape
    _, N2, _, H2, W2 = fmap2.shape
    M = coords.shape[1]
    R = radius
    D = 2 * R + 2
    dtype = fmap1.dtype.type
    out = np.zeros((B, M, D, D, PH, PW), dtype=dtype)
    f1, f2 = _f64 if _f64 is not None else (fmap1.astype(np.float64), fmap2.astype(np.float64))
    off = np.arange(D) - R
    for b in range(B):
        for m0 in range(0, M, chunk):
            m1 = min(M, m0 + chunk)
            fx = _floor_int(coords[b, m0:m1, 0])  # [m,P,P]
            fy = _floor_int(coords[b, m0:m1, 1])
            i1 = fy[:, None, None] + off[None, :, None, None, None]  # [m,D,1,P,P]
            j1 = fx[:, None, None] + off[None, None, :, None, None]  # [m,1,D,P,P]
            i1 = np.broadcast_to(i1, (m1 - m0, D, D, PH, PW))
            j1 = np.broadcast_to(j1, (m1 - m0, D, D, PH, PW))
            ok = (i1 >= 0) & (i1 < H2) & (j1 >= 0) & (j1 < W2)
            i1c = np.clip(i1, 0, H2 - 1)
            j1c = np.clip(j1, 0, W2 - 1)
            v = jj[m0:m1][:, None, None, None, None]
            g = f2[b][v, :, i1c, j1c]  # [m,D,D,P,P,C]
            p = f1[b][ii[m0:m1]]  # [m,C,P,P]
            s = np.einsum("mabijc,mcij->mabij", g, p)
            s = np.where(ok, s, 0.0)
            out[b, m0:m1] = s.astype(np.float32).astype(dtype)
    return out


def corr_forward(fmap1, fmap2, coords, ii, jj, radius, _f64=None):
    """cuda_corr.forward: returns [B,M,2r+1(x),2r+1(y),P,P] in fmap dtype."""
    dtype = fmap1.dtype.type
    raw = corr_raw(fmap1, fmap2, coords, ii, jj, radius, _f64=_f64)
    D = 2 * radius + 2
    w00, w01, w10, w11 = _blend_weights(coords, dtype)
    e = lambda w: w[:, :, None, None]
    with np.errstate(invalid="ignore", over="ignore"):
        out = (e(w00) * raw[:, :, 0:D - 1, 0:D - 1]).astype(dtype)
        out = (out + (e(w01) * raw[:, :, 0:D - 1, 1:D]).astype(dtype)).astype(dtype)
        out = (out + (e(w10) * raw[:, :, 1:D, 0:D - 1]).astype(dtype)).astype(dtype)
        out = (out + (e(w11) * raw[:, :, 1:D, 1:D]).astype(dtype)).astype(dtype)
    return np.ascontiguousarray(out.transpose(0, 1, 3, 2, 4, 5))


def corr_pyramid(fmap1, pyramid, scales, coords, ii, jj, radius):
    """DPVO.corr (dpvo.py:200-207): stack levels last, flatten -> [B,M,49*P*P*L]."""
    outs = [corr_forward(fmap1, fm, (coords.astype(np.float32) * np.float32(s)).astype(np.float32), ii, jj, radius)
            for fm, s in zip(pyramid, scales)]
    st = np.stack(outs, -1)
    return st.reshape(st.shape[0], st.shape[1], -1)


def corr_pyramid_threads(fmap1, pyramid, scales, coords, ii, jj, radius, threads):
    """corr_pyramid over `threads` host threads (edge chunks are independent; numpy's gathers and
    einsum release the GIL).  Same result as corr_pyramid; used by bench.py's CPU baseline legs."""
    from concurrent.futures import ThreadPoolExecutor
    M = coords.shape[1]
    threads = max(1, min(int(threads), M)) if M else 1
    f1 = fmap1.astype(np.float64)
    f2s = [fm.astype(np.float64) for fm in pyramid]
    bounds = np.linspace(0, M, threads + 1).astype(np.int64)

    def work(t):
        a, b = int(bounds[t]), int(bounds[t + 1])
        outs = [corr_forward(fmap1, fm, (coords[:, a:b].astype(np.float32) * np.float32(s)).astype(np.float32),
                             ii[a:b], jj[a:b], radius, _f64=(f1, f2))
                for fm, f2, s in zip(pyramid, f2s, scales)]
        st = np.stack(outs, -1)
        return st.reshape(st.shape[0], st.shape[1], -1)

    with ThreadPoolExecutor(max_workers=threads) as ex:
        parts = list(ex.map(work, range(threads)))
    return np.concatenate(parts, 1)


def corr_backward(fmap1, fmap2, coords, ii, jj, grad, radius, acc_dtype=np.float64):
    """cuda_corr.backward.  grad: [B,M,2r+1(x),2r+1(y),P,P].  The blend adjoint follows the
    reference's rounding (feature dtype); the scatter accumulates in `acc_dtype`
    (the reference uses feature-dtype atomics, whose order is not defined)."""
    B, N1, C, PH, PW = fmap1.shape
    _, N2, _, H2, W2 = fmap2.shape
    M = coords.shape[1]
    R = radius
    D = 2 * R + 2
    dtype = fmap1.dtype.type
    g = np.ascontiguousarray(grad.transpose(0, 1, 3, 2, 4, 5)).astype(dtype)  # [B,M,y,x,P,P]
    w00, w01, w10, w11 = _blend_weights(coords, dtype)
    e = lambda w: w[:, :, None, None]
    cg = np.zeros((B, M, D, D, PH, PW), dtype=dtype)
    t = np.zeros_like(cg); t[:, :, 0:D - 1, 0:D - 1] = (e(w00) * g).astype(dtype); cg = (cg + t).astype(dtype)
    t = np.zeros_like(cg); t[:, :, 0:D - 1, 1:D] = (e(w01) * g).astype(dtype); cg = (cg + t).astype(dtype)
    t = np.zeros_like(cg); t[:, :, 1:D, 0:D - 1] = (e(w10) * g).astype(dtype); cg = (cg + t).astype(dtype)
    t = np.zeros_like(cg); t[:, :, 1:D, 1:D] = (e(w11) * g).astype(dtype); cg = (cg + t).astype(dtype)
    g1 = np.zeros(fmap1.shape, dtype=acc_dtype)
    g2 = np.zeros(fmap2.shape, dtype=acc_dtype)
    f1 = fmap1.astype(acc_dtype)
    f2 = fmap2.astype(acc_dtype)
    cga = cg.astype(acc_dtype)
    off = np.arange(D) - R
    for b in range(B):
        fx = _floor_int(coords[b, :, 0])
        fy = _floor_int(coords[b, :, 1])
        for m in range(M):
            u, v = int(ii[m]), int(jj[m])
            for i0 in range(PH):
                for j0 in range(PW):
                    i1 = fy[m, i0, j0] + off
                    j1 = fx[m, i0, j0] + off
                    vi = (i1 >= 0) & (i1 < H2)
                    vj = (j1 >= 0) & (j1 < W2)
                    if not vi.any() or not vj.any():
                        continue
                    yi = i1[vi]; xj = j1[vj]
                    G = cga[b, m][np.ix_(vi, vj)][:, :, i0, j0]  # [ny,nx]
                    win = f2[b, v][:, yi[:, None], xj[None, :]]  # [C,ny,nx]
                    g1[b, u, :, i0, j0] += (win * G[None]).sum((1, 2))
                    g2[b, v][:, yi[:, None], xj[None, :]] += f1[b, u, :, i0, j0][:, None, None] * G[None]
    return g1, g2


def patchify_forward(net, coords, radius):
    """cuda_corr.patchify_forward: [B,M,C,D,D], zeros outside the map."""
    B, C, H, W = net.shape
    M = coords.shape[1]
    R = radius
    D = 2 * R + 2
    out = np.zeros((B, M, C, D, D), dtype=net.dtype)
    off = np.arange(D) - R
    for b in range(B):
        fx = _floor_int(coords[b, :, 0])
        fy = _floor_int(coords[b, :, 1])
        i = fy[:, None, None] + off[None, :, None]
        j = fx[:, None, None] + off[None, None, :]
        i = np.broadcast_to(i, (M, D, D)); j = np.broadcast_to(j, (M, D, D))
        ok = (i >= 0) & (i < H) & (j >= 0) & (j < W)
        ic = np.clip(i, 0, H - 1); jc = np.clip(j, 0, W - 1)
        g = net[b][:, ic, jc]  # [C,M,D,D]
        g = np.where(ok[None], g, 0)
        out[b] = g.transpose(1, 0, 2, 3)
    return out


def patchify_backward(net_shape, coords, gradient, radius, acc_dtype=np.float64):
    B, C, H, W = net_shape
    M = coords.shape[1]
    R = radius
    D = 2 * R + 2
    grad = np.zeros((B, C, H, W), dtype=acc_dtype)
    for b in range(B):
        fx = _floor_int(coords[b, :, 0])
        fy = _floor_int(coords[b, :, 1])
        for m in range(M):
            for a in range(D):
                i = fy[m] + a - R
                if i < 0 or i >= H:
                    continue
                for c in range(D):
                    j = fx[m] + c - R
                    if 0 <= j < W:
                        grad[b, :, i, j] += gradient[b, m, :, a, c].astype(acc_dtype)
    return grad


def patchify(net, coords, radius, mode="bilinear"):
    """altcorr.patchify (correlation.py:55-74) including the optional bilinear blend in net dtype."""
    dtype = net.dtype.type
    patches = patchify_forward(net, coords.astype(np.float32), radius)
    if mode != "bilinear":
        return patches
    c32 = coords.astype(np.float32)
    offset = (c32 - np.floor(c32)).astype(dtype)
    dx = offset[:, :, None, None, None, 0]
    dy = offset[:, :, None, None, None, 1]
    d = 2 * radius + 1
    one = dtype(1)
    r = lambda a: a.astype(dtype)
    x00 = r(r(r(one - dy) * r(one - dx)) * patches[..., :d, :d])
    x01 = r(r(r(one - dy) * dx) * patches[..., :d, 1:])
    x10 = r(r(dy * r(one - dx)) * patches[..., 1:, :d])
    x11 = r(r(dy * dx) * patches[..., 1:, 1:])
    return r(r(r(x00 + x01) + x10) + x11)
