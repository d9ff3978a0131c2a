"""CPU restatement of cuda_ba.solve_system — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows fastba/ba.cpp:99-180 in numpy float64 with a dense solve: J from the two 7x7 blocks of every
edge, A = J^T J, A.diag += A.diag * lm, A.diag += ep, b = -J^T res, solve the leading 7*freen block
(everything when freen < 0), zeros for the fixed poses.  Pinned by tests/golden/ref_cpu_pgo.npz
(outputs of the reference extension's own solve_system, tests/golden/make_golden_pgo.py).
"""
import numpy as np


def solve_system(J_Ginv_i, J_Ginv_j, ii, jj, res, ep, lm, freen):
    Ji = np.asarray(J_Ginv_i, np.float64); Jj = np.asarray(J_Ginv_j, np.float64)
    ii = np.asarray(ii, np.int64); jj = np.asarray(jj, np.int64)
    r = len(ii)
    n = int(max(ii.max(), jj.max())) + 1
    if np.any(ii == jj):
        raise ValueError("self edge (the reference exits, ba.cpp:151-152)")
    J = np.zeros((r * 7, n * 7))
    for x in range(r):
        J[x * 7:(x + 1) * 7, ii[x] * 7:(ii[x] + 1) * 7] = Ji[x]
        J[x * 7:(x + 1) * 7, jj[x] * 7:(jj[x] + 1) * 7] = Jj[x]
    v = np.asarray(res, np.float32).astype(np.float64).reshape(-1)
    b = -(J.T @ v)
    A = J.T @ J
    d = np.einsum("ii->i", A)
    d += d * float(np.float32(lm))          # the reference takes float arguments (ba.cpp:120)
    d += float(np.float32(ep))
    F = n * 7 if freen < 0 else min(freen, n) * 7
    delta = np.zeros(n * 7)
    if F:
        delta[:F] = np.linalg.solve(A[:F, :F], b[:F])
    return delta.astype(np.float32).reshape(n, 7)
