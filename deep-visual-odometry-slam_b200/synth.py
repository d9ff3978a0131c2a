"""Synthetic TartanAir-shaped inputs for the hot path (no datasets, no checkpoints).

Sizes follow the reference's default config: 480x640 images, RES=4 ->
fmap 120x160 and 30x40, 128 channels, 96 patches/frame of 3x3, radius 3, ring of 36
frames, PATCH_LIFETIME 13, REMOVAL_WINDOW 22, OPTIMIZATION_WINDOW 10.
Everything is generated on CPU with numpy from a seed so the GPU path and the CPU
oracle see identical bits.
"""
import numpy as np

M_DEFAULT = 96
MEM = 36
H1, W1 = 120, 160
H2, W2 = 30, 40
C = 128
P = 3
LIFETIME = 13
REMOVAL = 22
OPT_WINDOW = 10
INTRINSICS = (320.0 / 4, 320.0 / 4, 320.0 / 4, 240.0 / 4)  # fx fy cx cy at 1/4 res (dpvo.py:401)


def dpvo_edges(n, M=M_DEFAULT, lifetime=LIFETIME, removal=REMOVAL):
    """Active edge list (ii, jj, kk) after `n` frames, built with the reference's
    rules: __edges_forw/__edges_back (dpvo.py:362-375) appended per frame, then the
    removal-window cull of keyframe() (dpvo.py:305-310, no keyframe dropped).
    kk = frame*M + slot, ii = source frame of the patch, jj = target frame."""
    kk_l, jj_l = [], []
    kk = np.zeros(0, np.int64)
    jj = np.zeros(0, np.int64)
    for cur in range(1, n + 1):  # cur = self.n after the increment
        # forward: patches of frames [cur-r, cur-2] -> frame cur-1
        t0 = M * max(cur - lifetime, 0)
        t1 = M * max(cur - 1, 0)
        fk = np.arange(t0, t1, dtype=np.int64)
        fj = np.full_like(fk, cur - 1)
        # backward: patches of frame cur-1 -> frames [cur-r, cur-1]
        bk = np.repeat(np.arange(M * (cur - 1), M * cur, dtype=np.int64), cur - max(cur - lifetime, 0))
        bj = np.tile(np.arange(max(cur - lifetime, 0), cur, dtype=np.int64), M)
        kk = np.concatenate([kk, fk, bk])
        jj = np.concatenate([jj, fj, bj])
        keep = (kk // M) >= cur - removal
        kk, jj = kk[keep], jj[keep]
    del kk_l, jj_l
    ii = kk // M
    return ii, jj, kk


def corr_inputs(E, seed=0, mem=MEM, n_patches=None, hw1=(H1, W1), hw2=(H2, W2), channels=C,
                oob_frac=0.05, dtype=np.float16, edges=None, M=M_DEFAULT):
    """fmap rings (NCHW, as the reference allocates them), gmap, coords, patch/frame indices.
    If `edges`=(ii,jj,kk) is given, indices follow it (wrapped like dpvo.py:203-204)."""
    rng = np.random.default_rng(seed)
    n_patches = n_patches or mem * M
    fmap1 = rng.standard_normal((1, mem, channels, hw1[0], hw1[1]), dtype=np.float32).astype(dtype)
    fmap2 = rng.standard_normal((1, mem, channels, hw2[0], hw2[1]), dtype=np.float32).astype(dtype)
    gmap = rng.standard_normal((1, n_patches, channels, P, P), dtype=np.float32).astype(dtype)
    if edges is None:
        us = rng.integers(0, n_patches, E).astype(np.int64)
        vs = rng.integers(0, mem, E).astype(np.int64)
    else:
        ii, jj, kk = edges
        E = len(kk)
        us = (kk % n_patches).astype(np.int64)
        vs = (jj % mem).astype(np.int64)
    cx = rng.uniform(0, hw1[1] - 1, E).astype(np.float32)
    cy = rng.uniform(0, hw1[0] - 1, E).astype(np.float32)
    n_oob = int(E * oob_frac)
    if n_oob:
        sel = rng.choice(E, n_oob, replace=False)
        cx[sel] += rng.choice([-1.0, 1.0], n_oob).astype(np.float32) * hw1[1]
        cy[sel[: n_oob // 2]] -= hw1[0]
    gy, gx = np.meshgrid(np.arange(P, dtype=np.float32) - 1, np.arange(P, dtype=np.float32) - 1, indexing="ij")
    coords = np.zeros((1, E, 2, P, P), np.float32)
    coords[0, :, 0] = cx[:, None, None] + gx[None] + rng.normal(0, 0.25, (E, P, P)).astype(np.float32)
    coords[0, :, 1] = cy[:, None, None] + gy[None] + rng.normal(0, 0.25, (E, P, P)).astype(np.float32)
    return dict(fmap1=fmap1, fmap2=fmap2, gmap=gmap, coords=coords, us=us, vs=vs)


# ------------------------------------------------------------------ SE3 helpers (numpy, float64)
def _quat_mul(a, b):
    ax, ay, az, aw = np.moveaxis(a, -1, 0)
    bx, by, bz, bw = np.moveaxis(b, -1, 0)
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw,
                     aw * bw - ax * bx - ay * by - az * bz], -1)


def _quat_rot(q, v):
    qv = q[..., :3]
    uv = 2.0 * np.cross(qv, v)
    return v + q[..., 3:4] * uv + np.cross(qv, uv)


def trajectory(n, seed=0, step=0.05, rot=0.02):
    """n world-to-camera poses [n,7] (tx ty tz qx qy qz qw): smooth forward motion."""
    rng = np.random.default_rng(seed + 1000)
    poses = np.zeros((n, 7))
    poses[0, 6] = 1.0
    for i in range(1, n):
        w = rng.normal(0, rot, 3)
        th = np.linalg.norm(w)
        dq = np.concatenate([np.sin(th / 2) * w / max(th, 1e-12), [np.cos(th / 2)]])
        dt = np.array([0.0, 0.0, -step]) + rng.normal(0, step * 0.2, 3)
        q = _quat_mul(dq, poses[i - 1, 3:])
        t = _quat_rot(dq, poses[i - 1, :3]) + dt
        poses[i, :3] = t
        poses[i, 3:] = q / np.linalg.norm(q)
    return poses


def reproject_centres(poses, patches, intr, ii, jj, kk):
    """float64 reprojection of patch centres (the geometry of fastba/ba_cuda.cu:276-300)."""
    fx, fy, cx, cy = intr
    px = patches[kk, 0, 1, 1]
    py = patches[kk, 1, 1, 1]
    d = patches[kk, 2, 1, 1]
    X = np.stack([(px - cx) / fx, (py - cy) / fy, np.ones_like(px), d], -1)
    qi, ti = poses[ii, 3:], poses[ii, :3]
    qj, tj = poses[jj, 3:], poses[jj, :3]
    qi_inv = qi * np.array([-1, -1, -1, 1.0])
    qij = _quat_mul(qj, qi_inv)
    tij = tj - _quat_rot(qij, ti)
    Y = _quat_rot(qij, X[:, :3]) + tij * X[:, 3:4]
    return np.stack([fx * Y[:, 0] / Y[:, 2] + cx, fy * Y[:, 1] / Y[:, 2] + cy], -1)


def ba_inputs(n_frames, seed=0, M=M_DEFAULT, n_buf=None, pose_noise=0.01, px_noise=0.5, hw=(H1, W1),
              edges=None):
    """A DPVO-shaped BA problem: `n_frames` frames, DPVO-rule edges, smooth trajectory,
    inverse depths U(0.2,1), targets = true reprojection + N(0,px_noise), weights U(0,1).
    Poses handed to BA are the truth perturbed by `pose_noise`."""
    rng = np.random.default_rng(seed)
    n_buf = n_buf or n_frames + 2
    ii, jj, kk = edges if edges is not None else dpvo_edges(n_frames, M)
    true_poses = trajectory(n_buf, seed)
    patches = np.zeros((n_buf * M, 3, P, P))
    cxs = rng.uniform(8, hw[1] - 8, n_buf * M)
    cys = rng.uniform(8, hw[0] - 8, n_buf * M)
    gy, gx = np.meshgrid(np.arange(P) - 1.0, np.arange(P) - 1.0, indexing="ij")
    patches[:, 0] = cxs[:, None, None] + gx
    patches[:, 1] = cys[:, None, None] + gy
    patches[:, 2] = rng.uniform(0.2, 1.0, n_buf * M)[:, None, None]
    intr = np.array(INTRINSICS)
    target = reproject_centres(true_poses, patches, intr, ii, jj, kk)
    target = target + rng.normal(0, px_noise, target.shape)
    weight = rng.uniform(0, 1, target.shape)
    poses = true_poses.copy()
    poses[1:, :3] += rng.normal(0, pose_noise, (n_buf - 1, 3))
    dq = rng.normal(0, pose_noise, (n_buf - 1, 3))
    q = poses[1:, 3:].copy()
    q[:, :3] += 0.5 * dq
    poses[1:, 3:] = q / np.linalg.norm(q, axis=1, keepdims=True)
    noisy_patches = patches.copy()
    noisy_patches[:, 2] *= rng.uniform(0.8, 1.25, n_buf * M)[:, None, None]
    intrinsics = np.tile(intr[None], (n_buf, 1))
    return dict(poses=poses.astype(np.float32), patches=noisy_patches.astype(np.float32),
                intrinsics=intrinsics.astype(np.float32), target=target.astype(np.float32),
                weight=weight.astype(np.float32), ii=ii.astype(np.int64), jj=jj.astype(np.int64),
                kk=kk.astype(np.int64), n=n_frames, M=M)


# ------------------------------------------------------------------ roofline accounting
def corr_algorithmic_bytes(coords, shapes, scales, R=3, C=128, P=3, esz=2):
    """gmap patch + union of in-bounds window pixels per level + output + coords + indices."""
    E = coords.shape[1]
    D = 2 * R + 2
    total = E * (C * P * P * esz + len(scales) * (D - 1) ** 2 * P * P * esz + 2 * P * P * 4 + 16)
    off = np.arange(D) - R
    for (H, W), s in zip(shapes, scales):
        c = coords[0] * np.float32(s)
        fx = np.floor(c[:, 0]).astype(np.int64).reshape(E, -1)
        fy = np.floor(c[:, 1]).astype(np.int64).reshape(E, -1)
        xs = fx[:, :, None] + off[None, None]  # [E,9,D]
        ys = fy[:, :, None] + off[None, None]
        # union size via a small bitmap per edge (bbox-relative)
        x0 = xs.min((1, 2)); y0 = ys.min((1, 2))
        bw = xs.max((1, 2)) - x0 + 1
        n = 0
        for e0 in range(0, E, 4096):
            sl = slice(e0, min(E, e0 + 4096))
            X = xs[sl][:, :, None, :]  # [e,9,1,D]
            Y = ys[sl][:, :, :, None]  # [e,9,D,1]
            ok = (X >= 0) & (X < W) & (Y >= 0) & (Y < H)
            key = (Y - y0[sl, None, None, None]) * bw[sl, None, None, None] + (X - x0[sl, None, None, None])
            key = np.where(ok, key, -1).reshape(key.shape[0], -1)
            key.sort(axis=1)
            uniq = (key[:, 1:] != key[:, :-1]).sum(1) + 1 - (key[:, 0] == -1)
            n += int(uniq.sum())
        total += n * C * esz
    return total


# ----------------------------------------------------------------------------- Update operator
UPDATE_DIM = 384


def update_state_dict(seed=0, P=3, dim=UPDATE_DIM):
    """Random-initialised parameters of the reference's Update module (net.py:23-66), keyed like its
    state_dict, as float32 numpy arrays from a seeded numpy generator (weights ~ U(-1,1)/sqrt(fan_in)
    like nn.Linear's default scale; LayerNorm gains near 1 and small biases so every term matters)."""
    rng = np.random.default_rng(seed)
    sd = {}

    def linear(name, fin, fout):
        k = 1.0 / np.sqrt(fin)
        sd[name + ".weight"] = rng.uniform(-k, k, (fout, fin)).astype(np.float32)
        sd[name + ".bias"] = rng.uniform(-k, k, (fout,)).astype(np.float32)

    def lnorm(name):
        sd[name + ".weight"] = rng.uniform(0.8, 1.2, (dim,)).astype(np.float32)
        sd[name + ".bias"] = rng.uniform(-0.1, 0.1, (dim,)).astype(np.float32)

    linear("c1.0", dim, dim); linear("c1.2", dim, dim)
    linear("c2.0", dim, dim); linear("c2.2", dim, dim)
    lnorm("norm")
    for a in ("agg_kk", "agg_ij"):
        for l in "fgh":
            linear(f"{a}.{l}", dim, dim)
    lnorm("gru.0")
    linear("gru.1.gate.0", dim, dim); linear("gru.1.res.0", dim, dim); linear("gru.1.res.2", dim, dim)
    lnorm("gru.2")
    linear("gru.3.gate.0", dim, dim); linear("gru.3.res.0", dim, dim); linear("gru.3.res.2", dim, dim)
    linear("corr.0", 2 * 49 * P * P, dim); linear("corr.2", dim, dim); lnorm("corr.3"); linear("corr.5", dim, dim)
    linear("d.1", dim, 2); linear("w.1", dim, 2)
    return sd


def update_inputs(n_frames=12, M=8, seed=0, P=3, dim=UPDATE_DIM, n_ctx=None):
    """Edge list of a DPVO patch graph with the operator's per-edge inputs: hidden state `net`,
    correlation features `corr` and the context table `imap` gathered through kk % len(imap)."""
    rng = np.random.default_rng(seed + 1000)
    ii, jj, kk = dpvo_edges(n_frames, M=M)
    E = len(ii)
    n_ctx = n_ctx or n_frames * M
    return dict(ii=ii.astype(np.int64), jj=jj.astype(np.int64), kk=kk.astype(np.int64),
                net=(0.5 * rng.standard_normal((E, dim))).astype(np.float32),
                corr=rng.standard_normal((E, 2 * 49 * P * P)).astype(np.float32),
                imap=rng.standard_normal((n_ctx, dim)).astype(np.float32))


def encoder_state_dict(out_dim, seed):
    """Seeded parameters with the reference BasicEncoder4's state-dict keys and shapes (extractor.py:200-244);
    kaiming-style scale (fan_out) like its initialiser, small non-zero biases so that they are exercised."""
    rng = np.random.default_rng(seed)
    shapes = [("conv1", 32, 3, 7)]
    for blk, cin, cout in (("layer1.0", 32, 32), ("layer1.1", 32, 32), ("layer2.0", 32, 64), ("layer2.1", 64, 64)):
        shapes += [(blk + ".conv1", cout, cin, 3), (blk + ".conv2", cout, cout, 3)]
        if cin != cout:
            shapes.append((blk + ".downsample.0", cout, cin, 1))
    shapes.append(("conv2", out_dim, 64, 1))
    sd = {}
    for name, co, ci, k in shapes:
        std = np.sqrt(2.0 / (co * k * k))
        sd[name + ".weight"] = (rng.standard_normal((co, ci, k, k)) * std).astype(np.float32)
        sd[name + ".bias"] = (rng.standard_normal(co) * 0.05).astype(np.float32)
    return sd


def encoder_images(n, H, W, seed):
    """[1, n, 3, H, W] float32 in the range dpvo.py:389 produces (2 * (u8 / 255) - 0.5), smooth + texture."""
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    base = np.stack([128 + 100 * np.sin(0.07 * xx + c + f) * np.cos(0.05 * yy - c) for f in range(n) for c in range(3)])
    u8 = np.clip(base.reshape(n, 3, H, W) + rng.integers(-25, 25, (n, 3, H, W)), 0, 255).astype(np.uint8)
    return (2 * (u8.astype(np.float32) / 255.0) - 0.5)[None]
