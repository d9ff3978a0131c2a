"""TEST INFRASTRUCTURE ONLY (never imported by the product path).

numpy float32 restatement of DPViewer's two CUDA helpers (modules/DPViewer/dpviewer/viewer_cuda.cu):
  pose_to_matrix_kernel :221-262, iproj_kernel :56-164, backproject_and_filter's selection :199-213.
PARITY (round 1: unpinned; since round 2 pinned on the GPU box by oracle/_ref/dpviewer_cuda.so, see tests/test_viewer_gpu.py): the reference kernels only exist as CUDA inside a torch + Pangolin/OpenGL extension that cannot be
built or run here (no GPU in this container, no graphics stack on the box), and the reference has no test or golden
vector for them.  pose_to_matrix is cross-checked against the pinned lietorch oracle (inverse pose -> 4x4 matrix).
"""
import numpy as np

f32 = np.float32


def _act_so3(q, X):
    uv = [f32(2) * (q[1] * X[2] - q[2] * X[1]), f32(2) * (q[2] * X[0] - q[0] * X[2]), f32(2) * (q[0] * X[1] - q[1] * X[0])]
    return [X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]),
            X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]),
            X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0])]


def _act_se3(t, q, X):
    Y = _act_so3(q, X)
    return [Y[0] + X[3] * t[0], Y[1] + X[3] * t[1], Y[2] + X[3] * t[2], X[3]]


def _inv_se3(t, q):
    qi = [-q[0], -q[1], -q[2], q[3]]
    ti = _act_so3(qi, t)
    return [-ti[0], -ti[1], -ti[2]], qi


def pose_to_matrix(poses):
    """viewer_cuda.cu:221-262 (vectorised over poses; every operand stays float32)."""
    p = np.asarray(poses, dtype=f32)
    t, q = _inv_se3([p[:, 0], p[:, 1], p[:, 2]], [p[:, 3], p[:, 4], p[:, 5], p[:, 6]])
    one, two, zero = f32(1), f32(2), np.zeros_like(p[:, 0])
    rows = [[one - two * q[1] * q[1] - two * q[2] * q[2], two * q[0] * q[1] - two * q[3] * q[2], two * q[0] * q[2] + two * q[3] * q[1], t[0]],
            [two * q[0] * q[1] + two * q[3] * q[2], one - two * q[0] * q[0] - two * q[2] * q[2], two * q[1] * q[2] - two * q[3] * q[0], t[1]],
            [two * q[0] * q[2] - two * q[3] * q[1], two * q[1] * q[2] + two * q[3] * q[0], one - two * q[0] * q[0] - two * q[1] * q[1], t[2]],
            [zero, zero, zero, zero + one]]
    return np.stack([np.stack(r, -1) for r in rows], 1).astype(f32)


def iproj(index, n_frames, thresh, images, poses, disps, intrinsics):
    """viewer_cuda.cu:56-164, vectorised over the pixels.  Returns points, colors, count and `margin`, the distance of
    every comparison the kernel makes from its decision boundary (tests skip pixels whose margin is tiny)."""
    images, poses = np.asarray(images), np.asarray(poses, dtype=f32)
    disps, intrinsics = np.asarray(disps, dtype=f32), np.asarray(intrinsics, dtype=f32)
    ht, wd = disps.shape[1:]
    I, J = np.meshgrid(np.arange(ht), np.arange(wd), indexing="ij")
    inner = (I < ht - 1) & (J < wd - 1)
    intr = f32(8) * intrinsics[index]
    t1, q1 = _inv_se3(list(poses[index, :3]), list(poses[index, 3:]))
    d = disps[index]
    dpad = np.pad(d, ((0, 1), (0, 1)))
    dx, dy = dpad[:-1, 1:] - d, dpad[1:, :-1] - d
    grad = np.sqrt(dx * dx + dy * dy, dtype=f32)
    count = np.where(grad.astype(np.float64) > 0.01, f32(-100), f32(0)).astype(f32)
    margin = np.abs(grad.astype(np.float64) - 0.01)
    with np.errstate(divide="ignore", invalid="ignore"):
        X0 = [(J.astype(f32) - intr[2]) / intr[0], (I.astype(f32) - intr[3]) / intr[1], np.ones_like(d), d]
        X1 = _act_se3(t1, q1, X0)
        points = np.stack([X0[0] / X0[3], X0[1] / X0[3], X0[2] / X0[3]], -1).astype(f32)
        colors = np.stack([images[index, 2], images[index, 1], images[index, 0]], -1)
        for jx in range(n_frames):
            if jx == index:
                continue
            k = f32(8) * intrinsics[jx]
            X2 = _act_se3(list(poses[jx, :3]), list(poses[jx, 3:]), X1)
            x1 = k[0] * (X2[0] / X2[2]) + k[2]
            y1 = k[1] * (X2[1] / X2[2]) + k[3]
            ok = np.isfinite(x1) & np.isfinite(y1) & (np.abs(x1) < 1e9) & (np.abs(y1) < 1e9)
            # C roundf: half away from zero
            rnd = lambda v: (np.sign(v) * np.floor(np.abs(v) + f32(0.5))).astype(np.int64)
            i1, j1 = rnd(np.where(ok, y1, -5)), rnd(np.where(ok, x1, -5))
            inb = ok & (i1 >= 0) & (i1 < ht) & (j1 >= 0) & (j1 < wd) & (d.astype(np.float64) > 0.1)
            z1 = disps[jx][np.clip(i1, 0, ht - 1), np.clip(j1, 0, wd - 1)]
            z2 = X2[3] / X2[2]
            r = f32(100) * (np.maximum(z1 / z2, z2 / z1) - f32(1))
            count = count + np.where(inb & (r < f32(thresh)), f32(1), f32(0)).astype(f32)
            frac = lambda v: np.abs(np.abs(v - np.floor(v)) - 0.5)
            near = np.minimum.reduce([np.abs(np.where(ok, r, 1e9) - thresh), frac(np.where(ok, x1, 0.25)) * 50,
                                      frac(np.where(ok, y1, 0.25)) * 50, np.abs(d.astype(np.float64) - 0.1) * 1e3])
            margin = np.minimum(margin * 1.0, np.where(np.isfinite(near), near, 0.0)) if jx >= 0 else margin
    z = ~inner
    points[z] = 0
    colors = colors.copy(); colors[z] = 0
    count[z] = 0
    return points.reshape(-1, 3), colors.reshape(-1, 3), count.reshape(-1), margin.reshape(-1), inner.reshape(-1)
