"""Multi-threaded CPU restatement of the reference's pure-PyTorch bundle adjustment, dpvo/ba.py::BA
(forward only).  TEST INFRASTRUCTURE ONLY: tests/ and bench.py's cpu_baseline / --impl reference legs.

This is the "reference's pure-PyTorch CPU path" of BASELINE.json's north star.  It is NOT the same
algorithm as the CUDA BA: projection clamps Z >= 0.1, validity is |r| < 250 plus the
caller's bounds, damping is `ep` (10 in net.py:254), depths are clamped to [1e-3, 10], `fixedp` poses are
fixed, intrinsics are per frame, a failed Cholesky gives a zero step, one Gauss-Newton step per call.

Parity status: PINNED by tests/golden/ref_cpu_ba.npz (the reference's own ba.BA on its lietorch CPU
backend, tests/golden/make_golden_cpu.py), checked in tests/test_oracle_cpu.py.

Follows /root/reference/methods/dpvo/dpvo/ba.py:86-182 (normal equations, Schur complement, block_solve
:67-76 with damping ep + 1e-4 A, retractions :47-56) and projective_ops.py:53-108 (transform + Jacobians);
the SE3 retraction Exp(dx) * T uses oracle/lietorch_oracle.py (float64, N poses only).
"""
import numpy as np
import torch

from . import lietorch_oracle as lo


def _qmul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz], -1)


def _qrot(q, p):
    qv = q[..., :3]
    uv = 2.0 * torch.cross(qv.expand_as(p), p, dim=-1)
    return p + q[..., 3:4] * uv + torch.cross(qv.expand_as(p), uv, dim=-1)


def _norm(q):
    return q / q.norm(dim=-1, keepdim=True)


def _hat(v):
    z = torch.zeros_like(v[..., 0])
    return torch.stack([torch.stack([z, -v[..., 2], v[..., 1]], -1), torch.stack([v[..., 2], z, -v[..., 0]], -1),
                        torch.stack([-v[..., 1], v[..., 0], z], -1)], -2)


def _qmat(q):
    x, y, z, w = q.unbind(-1)
    return torch.stack([torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
                        torch.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
                        torch.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], -2)


def transform_jac(poses, patches, intrinsics, ii, jj, kk):
    """projective_ops.transform(..., jacobian=True) in torch float32.  poses [N,7], patches [K,3,P,P],
    intrinsics [N,4] -> coords [E,P,P,2], valid [E], Ji [E,2,6], Jj [E,2,6], Jz [E,2,1]."""
    ti, qi = poses[ii, :3], _norm(poses[ii, 3:])
    tj, qj = poses[jj, :3], _norm(poses[jj, 3:])
    qinv = qi * torch.tensor([-1.0, -1.0, -1.0, 1.0])
    tinv = -_qrot(qinv, ti)
    q = _norm(_qmul(qj, qinv))
    t = tj + _qrot(qj, tinv)
    pt = patches[kk]
    Ki, Kj = intrinsics[ii], intrinsics[jj]
    e = lambda v: v[:, None, None]
    X0 = torch.stack([(pt[:, 0] - e(Ki[:, 2])) / e(Ki[:, 0]), (pt[:, 1] - e(Ki[:, 3])) / e(Ki[:, 1]),
                      torch.ones_like(pt[:, 2])], -1)                                   # [E,P,P,3]
    X1 = _qrot(q[:, None, None, :], X0) + t[:, None, None, :] * pt[:, 2][..., None]
    H1 = pt[:, 2]
    d = 1.0 / X1[..., 2].clamp(min=0.1)
    coords = torch.stack([e(Kj[:, 0]) * (d * X1[..., 0]) + e(Kj[:, 2]), e(Kj[:, 1]) * (d * X1[..., 1]) + e(Kj[:, 3])], -1)
    c = pt.shape[-1] // 2
    X, Y, Z, H = X1[:, c, c, 0], X1[:, c, c, 1], X1[:, c, c, 2], H1[:, c, c]
    fx, fy = Kj[:, 0], Kj[:, 1]
    dj = torch.where(Z.abs() > 0.2, 1.0 / torch.where(Z.abs() > 0.2, Z, torch.ones_like(Z)), torch.zeros_like(Z))
    o = torch.zeros_like(H)
    Ja = torch.stack([H, o, o, o, Z, -Y, o, H, o, -Z, o, X, o, o, H, Y, -X, o, o, o, o, o, o, o], -1).view(-1, 4, 6)
    Jp = torch.stack([fx * dj, o, -fx * X * dj * dj, o, o, fy * dj, -fy * Y * dj * dj, o], -1).view(-1, 2, 4)
    Jj = Jp @ Ja
    R = _qmat(q)
    Ad = torch.zeros(len(ii), 6, 6)
    Ad[:, :3, :3] = R; Ad[:, 3:, 3:] = R; Ad[:, :3, 3:] = _hat(t) @ R
    Ji = -(Ad.transpose(1, 2)[:, None] @ Jj[..., None])[..., 0]
    T4 = torch.cat([t, torch.ones_like(t[:, :1])], -1)[..., None]
    Jz = Jp @ T4
    return coords, (Z > 0.2).float(), Ji, Jj, Jz


def BA(poses, patches, intrinsics, targets, weights, lmbda, ii, jj, kk, bounds, ep=100.0, fixedp=1, structure_only=False):
    """One Gauss-Newton step of ba.py:86-182.  numpy / torch inputs; returns (poses [N,7], patches [K,3,P,P])
    as float32 numpy arrays."""
    as_t = lambda x, dt=torch.float32: torch.as_tensor(np.asarray(x)).to(dt)
    poses_t, patches_t, K = as_t(poses).view(-1, 7), as_t(patches), as_t(intrinsics).view(-1, 4)
    patches_t = patches_t.view(-1, 3, patches_t.shape[-1], patches_t.shape[-1])
    targets, weights = as_t(targets).view(-1, 2), as_t(weights).view(-1, 2)
    ii, jj, kk = (as_t(x, torch.int64) for x in (ii, jj, kk))
    n = int(max(ii.max(), jj.max())) + 1
    coords, v, Ji, Jj, Jz = transform_jac(poses_t, patches_t, K, ii, jj, kk)
    p = coords.shape[2]
    cc = coords[:, p // 2, p // 2]
    r = targets - cc
    v = v * (r.norm(dim=-1) < 250).float()
    v = v * ((cc[:, 0] > bounds[0]) & (cc[:, 1] > bounds[1]) & (cc[:, 0] < bounds[2]) & (cc[:, 1] < bounds[3])).float()
    r = (v[:, None] * r)[..., None]                       # [E,2,1]
    w = (v[:, None] * weights)[..., None]                 # [E,2,1]
    wJiT, wJjT, wJzT = (w * Ji).transpose(1, 2), (w * Jj).transpose(1, 2), (w * Jz).transpose(1, 2)
    Bii, Bij, Bji, Bjj = wJiT @ Ji, wJiT @ Jj, wJjT @ Ji, wJjT @ Jj
    Eik, Ejk = wJiT @ Jz, wJjT @ Jz
    vi, vj = wJiT @ r, wJjT @ r
    n = n - fixedp
    ix, jx = ii - fixedp, jj - fixedp
    kx, ku = torch.unique(kk, return_inverse=True, sorted=True)
    m = len(kx)

    def scat_mat(A, a, b, na, nb):
        ok = (a >= 0) & (b >= 0) & (a < na) & (b < nb)
        out = torch.zeros((na * nb,) + A.shape[1:])
        out.index_add_(0, a[ok] * nb + b[ok], A[ok])
        return out.view((na, nb) + A.shape[1:])

    def scat_vec(A, a, na):
        ok = (a >= 0) & (a < na)
        out = torch.zeros((na,) + A.shape[1:])
        out.index_add_(0, a[ok], A[ok])
        return out

    C = scat_vec(wJzT @ Jz, ku, m).view(m)
    wv = scat_vec(wJzT @ r, ku, m).view(m)
    lm = float(np.asarray(lmbda).reshape(-1)[0])
    Q = 1.0 / (C + lm)
    disps = patches_t[:, 2].clone()
    new_poses = np.asarray(poses, np.float32).reshape(-1, 7).copy()
    if structure_only or n <= 0:
        dZ = Q * wv
    else:
        B = scat_mat(Bii, ix, ix, n, n) + scat_mat(Bij, ix, jx, n, n) + scat_mat(Bji, jx, ix, n, n) + scat_mat(Bjj, jx, jx, n, n)
        E = scat_mat(Eik, ix, ku, n, m) + scat_mat(Ejk, jx, ku, n, m)                      # [n,m,6,1]
        vv = scat_vec(vi, ix, n) + scat_vec(vj, jx, n)                                     # [n,6,1]
        Bm = B.permute(0, 2, 1, 3).reshape(6 * n, 6 * n)
        Em = E[..., 0].permute(0, 2, 1).reshape(6 * n, m)
        EQ = Em * Q[None]
        S = Bm - EQ @ Em.t()
        y = vv.reshape(6 * n) - EQ @ wv
        S = S + (ep + 1e-4 * S) * torch.eye(6 * n)
        U, info = torch.linalg.cholesky_ex(S)
        dX = torch.zeros(6 * n) if bool(info.any()) else torch.cholesky_solve(y[:, None], U)[:, 0]
        dZ = Q * (wv - Em.t() @ dX)
        xi = dX.view(n, 6).double().numpy()
        idx = fixedp + np.arange(n)
        new_poses[idx] = lo.mul(lo.exp(xi), new_poses[idx].astype(np.float64)).astype(np.float32)
    disps[kx] = disps[kx] + dZ[:, None, None]
    disps = disps.clamp(min=1e-3, max=10.0)
    out_patches = torch.stack([patches_t[:, 0], patches_t[:, 1], disps], 1)
    return new_poses, out_patches.numpy()
