"""CPU restatement (numpy, float64) of the reference's fastba operators.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module, and only as the checker.  The product path never does.

Parity status: the reference has NO test of any kind for fastba.  This
restatement is pinned against the reference's own CUDA extension built from the
untouched sources (oracle/build_ref.py -> oracle/_ref/cuda_ba.so) executed on the GPU
box on identical seeded inputs (tests/test_ba_gpu.py::test_*_vs_reference_ext), and the
integer operators (neighbors, unique/inverse) bit-exactly against it.

Follows (paths under /root/reference/methods/dpvo/dpvo/fastba/):
  ba()          ba_cuda.cu:36-174 (SE3 helpers), :232-376 (residual, Jacobians, accumulation),
                :447-582 (solve sequence, both eff_impl modes), :178-229 (retractions);
                block_e.cu:43-300 (block-E: algebraically the same Schur complement as the
                dense E when kk == ii*PPF + slot, which is how DPVO numbers patches)
  neighbors()   ba.cpp:59-97
  reproject()   ba_cuda.cu:379-429,585-617
All arithmetic is float64 on the float32 inputs; accumulation order is not defined in
the reference (unordered fp32 atomics), so parity is judged by tolerance.
"""
import numpy as np


def _act_so3(q, X):
    qv = q[..., :3]
    uv = 2.0 * np.cross(qv, X)
    return X + q[..., 3:4] * uv + np.cross(qv, uv)


def _rel_se3(ti, qi, tj, qj):
    """Gij = Gj * Gi^-1 (ba_cuda.cu:74-85)."""
    qij = np.stack([
        -qj[:, 3] * qi[:, 0] + qj[:, 0] * qi[:, 3] - qj[:, 1] * qi[:, 2] + qj[:, 2] * qi[:, 1],
        -qj[:, 3] * qi[:, 1] + qj[:, 1] * qi[:, 3] - qj[:, 2] * qi[:, 0] + qj[:, 0] * qi[:, 2],
        -qj[:, 3] * qi[:, 2] + qj[:, 2] * qi[:, 3] - qj[:, 0] * qi[:, 1] + qj[:, 1] * qi[:, 0],
        qj[:, 3] * qi[:, 3] + qj[:, 0] * qi[:, 0] + qj[:, 1] * qi[:, 1] + qj[:, 2] * qi[:, 2]], -1)
    tij = tj - _act_so3(qij, ti)
    return tij, qij


def _adj_se3(t, q, X):
    """Y = Adj^T-style pull-back used for Ji (ba_cuda.cu:57-72)."""
    qinv = q * np.array([-1.0, -1.0, -1.0, 1.0])
    Y0 = _act_so3(qinv, X[:, 0:3])
    Y1 = _act_so3(qinv, X[:, 3:6])
    u = np.cross(X[:, 0:3], t)  # (t2*X1 - t1*X2, ...) == X x t
    Y1 = Y1 + _act_so3(qinv, u)
    return np.concatenate([Y0, Y1], -1)


def _exp_so3(phi):
    th2 = (phi * phi).sum()
    th4 = th2 * th2
    th = np.sqrt(th2)
    if th2 < 1e-8:
        imag = 0.5 - (1.0 / 48.0) * th2 + (1.0 / 3840.0) * th4
        real = 1.0 - (1.0 / 8.0) * th2 + (1.0 / 384.0) * th4
    else:
        imag = np.sin(0.5 * th) / th
        real = np.cos(0.5 * th)
    return np.array([imag * phi[0], imag * phi[1], imag * phi[2], real])


def _exp_se3(xi):
    q = _exp_so3(xi[3:])
    tau = xi[:3].copy()
    phi = xi[3:]
    th2 = (phi * phi).sum()
    th = np.sqrt(th2)
    t = tau.copy()
    if th > 1e-4:
        a = (1 - np.cos(th)) / th2
        tau = np.cross(phi, tau)
        t = t + a * tau
        b = (th - np.sin(th)) / (th * th2)
        tau = np.cross(phi, tau)
        t = t + b * tau
    return t, q


def _retr_se3(xi, t, q):
    dt, dq = _exp_se3(xi)
    q1 = np.array([
        dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1],
        dq[3] * q[1] + dq[1] * q[3] + dq[2] * q[0] - dq[0] * q[2],
        dq[3] * q[2] + dq[2] * q[3] + dq[0] * q[1] - dq[1] * q[0],
        dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2]])
    t1 = _act_so3(dq, t) + dt
    return t1, q1


def edge_terms(poses, patches, intrinsics, target, weight, ii, jj, kk):
    """Per-edge residuals/Jacobians: returns r[E,2], w[E,2], Ji[E,2,6], Jj[E,2,6], Jz[E,2]."""
    fx, fy, cx, cy = [float(v) for v in intrinsics[0]]
    ti, qi = poses[ii, :3], poses[ii, 3:]
    tj, qj = poses[jj, :3], poses[jj, 3:]
    Xi = np.stack([(patches[kk, 0, 1, 1] - cx) / fx, (patches[kk, 1, 1, 1] - cy) / fy,
                   np.ones(len(kk)), patches[kk, 2, 1, 1]], -1)
    tij, qij = _rel_se3(ti, qi, tj, qj)
    Y3 = _act_so3(qij, Xi[:, :3]) + Xi[:, 3:4] * tij
    X, Y, Z, W = Y3[:, 0], Y3[:, 1], Y3[:, 2], Xi[:, 3]
    with np.errstate(divide="ignore", invalid="ignore"):
        d = np.where(Z >= 0.2, 1.0 / Z, 0.0)
        d2 = d * d
        x1 = fx * (X / Z) + cx
        y1 = fy * (Y / Z) + cy
        rx = target[:, 0] - x1
        ry = target[:, 1] - y1
        inb = (np.sqrt(rx * rx + ry * ry) < 128) & (Z > 0.2) & (x1 > -64) & (y1 > -64) & \
              (x1 < 2 * cx + 64) & (y1 < 2 * cy + 64)
    mask = inb.astype(np.float64)
    z = np.zeros_like(X)
    Jj0 = np.stack([fx * W * d, z, fx * -X * W * d2, fx * -X * Y * d2, fx * (1 + X * X * d2), fx * -Y * d], -1)
    Jj1 = np.stack([z, fy * W * d, fy * -Y * W * d2, fy * (-1 - Y * Y * d2), fy * (X * Y * d2), fy * X * d], -1)
    Jz0 = fx * (tij[:, 0] * d - tij[:, 2] * (X * d2))
    Jz1 = fy * (tij[:, 1] * d - tij[:, 2] * (Y * d2))
    Ji0 = _adj_se3(tij, qij, Jj0)
    Ji1 = _adj_se3(tij, qij, Jj1)
    r = np.stack([rx, ry], -1)
    w = mask[:, None] * weight
    return r, w, np.stack([Ji0, Ji1], 1), np.stack([Jj0, Jj1], 1), np.stack([Jz0, Jz1], -1)


def ba(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, PPF, t0, t1, iterations, eff_impl=False,
       return_updates=False):
    """cuda_ba.forward.  poses [Nb,7], patches [Np,3,P,P], intrinsics [Nb,4], target/weight [E,2];
    returns updated float64 copies (poses, patches) (+ list of (dX, dZ, kx) per iteration)."""
    poses = np.array(poses, dtype=np.float64).reshape(-1, 7).copy()
    P = patches.shape[-1]
    patches = np.array(patches, dtype=np.float64).reshape(-1, 3, P, P).copy()
    intrinsics = np.asarray(intrinsics, np.float64).reshape(-1, 4)
    target = np.asarray(target, np.float64).reshape(-1, 2)
    weight = np.asarray(weight, np.float64).reshape(-1, 2)
    lm = float(np.asarray(lmbda).reshape(-1)[0])
    ii = np.asarray(ii, np.int64); jj = np.asarray(jj, np.int64); kk = np.asarray(kk, np.int64)
    kx, ku = np.unique(kk, return_inverse=True)
    N = t1 - t0
    M = len(kx)
    ups = []
    for _ in range(iterations):
        r, w, Ji, Jj, Jz = edge_terms(poses, patches, intrinsics, target, weight, ii, jj, kk)
        ix = ii - t0
        jx = jj - t0
        vi = (ix >= 0) & (ix < N)
        vj = (jx >= 0) & (jx < N)
        B = np.zeros((max(N, 0), 6, max(N, 0), 6))
        E = np.zeros((max(N, 0), 6, M))
        v = np.zeros((max(N, 0), 6))
        C = np.zeros(M)
        u = np.zeros(M)
        for row in range(2):
            wr = w[:, row]
            a, b_, jz, rr = Ji[:, row], Jj[:, row], Jz[:, row], r[:, row]
            if N > 0:
                s = vi
                np.add.at(B, (ix[s], slice(None), ix[s], slice(None)), (wr[s, None, None] * a[s, :, None] * a[s, None, :]))
                np.add.at(E, (ix[s], slice(None), ku[s]), -(wr[s] * jz[s])[:, None] * a[s])
                np.add.at(v, (ix[s],), -(wr[s] * rr[s])[:, None] * a[s])
                s = vj
                np.add.at(B, (jx[s], slice(None), jx[s], slice(None)), (wr[s, None, None] * b_[s, :, None] * b_[s, None, :]))
                np.add.at(E, (jx[s], slice(None), ku[s]), (wr[s] * jz[s])[:, None] * b_[s])
                np.add.at(v, (jx[s],), (wr[s] * rr[s])[:, None] * b_[s])
                s = vi & vj
                np.add.at(B, (ix[s], slice(None), jx[s], slice(None)), -(wr[s, None, None] * a[s, :, None] * b_[s, None, :]))
                np.add.at(B, (jx[s], slice(None), ix[s], slice(None)), -(wr[s, None, None] * b_[s, :, None] * a[s, None, :]))
            np.add.at(C, ku, wr * jz * jz)
            np.add.at(u, ku, wr * rr * jz)
        Q = 1.0 / (C + lm)
        if N <= 0:
            dX = np.zeros((0, 6))
            dZ = Q * u
        else:
            Bm = B.reshape(6 * N, 6 * N)
            Em = E.reshape(6 * N, M)
            vm = v.reshape(6 * N)
            EQ = Em * Q[None]
            S = Bm - EQ @ Em.T
            y = vm - EQ @ u
            S = S + np.eye(6 * N) * (1e-4 * S + 1.0)
            L = np.linalg.cholesky(S)
            dX = np.linalg.solve(L.T, np.linalg.solve(L, y))
            dZ = Q * (u - Em.T @ dX)
            dX = dX.reshape(N, 6)
            for i in range(N):
                t_new, q_new = _retr_se3(dX[i], poses[t0 + i, :3], poses[t0 + i, 3:])
                poses[t0 + i, :3] = t_new
                poses[t0 + i, 3:] = q_new
        d = patches[kx, 2, 0, 0] + dZ
        d = np.where(d > 20, 1.0, d)
        d = np.maximum(d, 1e-4)
        patches[kx, 2] = d[:, None, None]
        ups.append((dX.copy(), dZ.copy(), kx.copy()))
    if return_updates:
        return poses, patches, ups
    return poses, patches


def _group_sum(keys, values):
    """sum `values` rows over equal `keys`: returns (unique keys ascending, sums).  Sorting + reduceat (np.add.at is
    unbuffered and far too slow for millions of edges)."""
    order = np.argsort(keys, kind="stable")
    ks = keys[order]
    if len(ks) == 0:
        return ks, values[:0]
    starts = np.flatnonzero(np.concatenate([[True], ks[1:] != ks[:-1]]))
    return ks[starts], np.add.reduceat(values[order], starts, axis=0)


def ba_sparse(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, PPF, t0, t1, iterations):
    """Same normal equations and update rule as `ba()` (ba_cuda.cu:232-376, :447-582, :178-229), assembled as
    scipy.sparse matrices (6x6 blocks summed per frame pair, E entries per (frame, patch)) and solved with a sparse
    direct solver in float64 — for global-BA sizes (BASELINE config 5: 1,000+ keyframes, 2.4 M edges) where the dense
    E of `ba()` (6N x #patches) does not fit.  Checked against `ba()` on a small window in tests/test_oracle_cpu.py.
    Returns (poses, patches, [(dX, dZ, kx) per iteration])."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    poses = np.array(poses, dtype=np.float64).reshape(-1, 7).copy()
    P = patches.shape[-1]
    patches = np.array(patches, dtype=np.float64).reshape(-1, 3, P, P).copy()
    intrinsics = np.asarray(intrinsics, np.float64).reshape(-1, 4)
    target = np.asarray(target, np.float64).reshape(-1, 2)
    weight = np.asarray(weight, np.float64).reshape(-1, 2)
    lm = float(np.asarray(lmbda).reshape(-1)[0])
    ii = np.asarray(ii, np.int64); jj = np.asarray(jj, np.int64); kk = np.asarray(kk, np.int64)
    kx, ku = np.unique(kk, return_inverse=True)
    N, M = t1 - t0, len(kx)
    n = 6 * N
    a6 = np.arange(6)
    ups = []
    for _ in range(iterations):
        r, w, Ji, Jj, Jz = edge_terms(poses, patches, intrinsics, target, weight, ii, jj, kk)
        ix, jx = ii - t0, jj - t0
        vi = (ix >= 0) & (ix < N)
        vj = (jx >= 0) & (jx < N)
        vij = vi & vj
        # per-edge 6x6 blocks / 6-vectors summed over the two residual rows
        Bii = np.einsum("er,era,erb->eab", w, Ji, Ji)
        Bjj = np.einsum("er,era,erb->eab", w, Jj, Jj)
        Bij = -np.einsum("er,era,erb->eab", w, Ji, Jj)
        Ei = -np.einsum("er,er,era->ea", w, Jz, Ji)
        Ej = np.einsum("er,er,era->ea", w, Jz, Jj)
        vi6 = -np.einsum("er,er,era->ea", w, r, Ji)
        vj6 = np.einsum("er,er,era->ea", w, r, Jj)
        C = np.bincount(ku, weights=(w * Jz * Jz).sum(1), minlength=M)
        u = np.bincount(ku, weights=(w * r * Jz).sum(1), minlength=M)
        Q = 1.0 / (C + lm)
        if N > 0:
            keys = np.concatenate([ix[vi] * N + ix[vi], jx[vj] * N + jx[vj], ix[vij] * N + jx[vij], jx[vij] * N + ix[vij]])
            blks = np.concatenate([Bii[vi], Bjj[vj], Bij[vij], np.swapaxes(Bij[vij], 1, 2)])
            pk, pb = _group_sum(keys, blks.reshape(len(keys), 36))
            fa, fb = pk // N, pk % N
            rows = (6 * fa)[:, None, None] + a6[None, :, None] + 0 * a6[None, None, :]
            cols = (6 * fb)[:, None, None] + 0 * a6[None, :, None] + a6[None, None, :]
            Bm = sp.coo_matrix((pb.ravel(), (rows.ravel(), cols.ravel())), shape=(n, n)).tocsr()
            ekeys = np.concatenate([ix[vi] * M + ku[vi], jx[vj] * M + ku[vj]])
            ek, eb = _group_sum(ekeys, np.concatenate([Ei[vi], Ej[vj]]))
            ef, ep = ek // M, ek % M
            Em = sp.coo_matrix((eb.ravel(), (((6 * ef)[:, None] + a6[None]).ravel(), np.repeat(ep, 6))), shape=(n, M)).tocsr()
            fk, fv = _group_sum(np.concatenate([ix[vi], jx[vj]]), np.concatenate([vi6[vi], vj6[vj]]))
            v = np.zeros((N, 6)); v[fk] = fv
            v = v.reshape(n)
            EQ = Em.multiply(Q[None, :]).tocsr()
            S = (Bm - EQ @ Em.T).tocsc()
            y = v - EQ @ u
            S = S + sp.diags(1e-4 * S.diagonal() + 1.0)           # ba_cuda.cu:546,560: S += I * (1e-4 * S + 1)
            dX = spl.splu(S.tocsc()).solve(y)
            dZ = Q * (u - Em.T @ dX)
            dXr = dX.reshape(N, 6)
            for i in range(N):
                t_new, q_new = _retr_se3(dXr[i], poses[t0 + i, :3], poses[t0 + i, 3:])
                poses[t0 + i, :3] = t_new
                poses[t0 + i, 3:] = q_new
        else:
            dXr = np.zeros((0, 6))
            dZ = Q * u
        d = patches[kx, 2, 0, 0] + dZ
        d = np.where(d > 20, 1.0, d)
        d = np.maximum(d, 1e-4)
        patches[kx, 2] = d[:, None, None]
        ups.append((dXr.copy(), dZ.copy(), kx.copy()))
    return poses, patches, ups


def neighbors(ii, jj):
    """cuda_ba.neighbors (ba.cpp:59-97): for every edge, the index of the previous / next edge
    of the same ii-group ordered by jj (stable), -1 at the ends."""
    ii = np.asarray(ii, np.int64); jj = np.asarray(jj, np.int64)
    n = len(ii)
    ix = np.full(n, -1, np.int64)
    jx = np.full(n, -1, np.int64)
    order = np.lexsort((np.arange(n), jj, ii))  # by ii, then jj, then original index (stable)
    same = ii[order][1:] == ii[order][:-1]
    ix[order[1:][same]] = order[:-1][same]
    jx[order[:-1][same]] = order[1:][same]
    return ix, jx


def reproject(poses, patches, intrinsics, ii, jj, kk):
    """cuda_ba.reproject: coords [1,E,2,P,P] of all patch pixels of k seen from frame j."""
    poses = np.asarray(poses, np.float64).reshape(-1, 7)
    P = patches.shape[-1]
    patches = np.asarray(patches, np.float64).reshape(-1, 3, P, P)
    fx, fy, cx, cy = [float(v) for v in np.asarray(intrinsics, np.float64).reshape(-1, 4)[0]]
    tij, qij = _rel_se3(poses[ii, :3], poses[ii, 3:], poses[jj, :3], poses[jj, 3:])
    E = len(ii)
    out = np.zeros((1, E, 2, P, P))
    for i in range(P):
        for j in range(P):
            X = np.stack([(patches[kk, 0, i, j] - cx) / fx, (patches[kk, 1, i, j] - cy) / fy, np.ones(E)], -1)
            Y = _act_so3(qij, X) + patches[kk, 2, i, j][:, None] * tij
            with np.errstate(divide="ignore", invalid="ignore"):
                out[0, :, 0, i, j] = fx * (Y[:, 0] / Y[:, 2]) + cx
                out[0, :, 1, i, j] = fy * (Y[:, 1] / Y[:, 2]) + cy
    return out


def unique_inverse(kk):
    """torch::_unique(kk, sorted=True, return_inverse=True) (ba_cuda.cu:447-449)."""
    return np.unique(np.asarray(kk, np.int64), return_inverse=True)
