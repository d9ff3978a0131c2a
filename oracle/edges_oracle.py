"""TEST INFRASTRUCTURE ONLY (never imported by the product path).

numpy restatement of the reference's patch-graph edge bookkeeping:
  append_factors / remove_factors      methods/dpvo/dpvo/dpvo.py:215-239
  keyframe (edge part + slot shifts)   dpvo.py:262-310
  __edges_forw / __edges_back          dpvo.py:362-375
  reduce_edges                         methods/dpvo/dpvo/loop_closure/optim_utils.py:27-64
Pinned: tests/test_oracle_cpu.py replays tests/golden/ref_cpu_edges.npz, which
tests/golden/make_golden_edges.py produced by running the reference's own method bodies on CPU.
"""
import numpy as np


def flatmeshgrid_ij(a, b):
    """utils.py:85-87 with indexing='ij': a varies slowest."""
    A, B = np.meshgrid(a, b, indexing="ij")
    return A.reshape(-1), B.reshape(-1)


def edges_forw(n, M, r):
    """dpvo.py:362-368 -> (patch indices, target frames)."""
    return flatmeshgrid_ij(np.arange(M * max(n - r, 0), M * max(n - 1, 0)), np.arange(n - 1, n))


def edges_back(n, M, r):
    """dpvo.py:370-375."""
    return flatmeshgrid_ij(np.arange(M * max(n - 1, 0), M * max(n, 0)), np.arange(max(n - r, 0), n))


class Edges:
    """The edge fields of PatchGraph (patchgraph.py:41-53) as numpy arrays."""

    def __init__(self, ix, M, dim):
        self.ix, self.M, self.dim = np.asarray(ix, dtype=np.int64), int(M), int(dim)
        z = lambda: np.zeros(0, dtype=np.int64)
        self.ii, self.jj, self.kk = z(), z(), z()
        self.net = np.zeros((0, dim), dtype=np.float16)
        self.target = np.zeros((0, 2), dtype=np.float32)
        self.weight = np.zeros((0, 2), dtype=np.float32)
        self.ii_inac, self.jj_inac, self.kk_inac = z(), z(), z()
        self.target_inac = np.zeros((0, 2), dtype=np.float32)
        self.weight_inac = np.zeros((0, 2), dtype=np.float32)

    def append_factors(self, ii, jj):
        """dpvo.py:215-221 (`ii` are patch indices)."""
        ii, jj = np.asarray(ii, dtype=np.int64), np.asarray(jj, dtype=np.int64)
        self.jj = np.concatenate([self.jj, jj])
        self.kk = np.concatenate([self.kk, ii])
        self.ii = np.concatenate([self.ii, self.ix[ii]])
        self.net = np.concatenate([self.net, np.zeros((len(ii), self.dim), dtype=np.float16)])

    def remove_factors(self, m, store):
        """dpvo.py:223-239."""
        m = np.asarray(m, dtype=bool)
        assert self.ii.size == self.weight.shape[0]
        if store:
            self.ii_inac = np.concatenate([self.ii_inac, self.ii[m]])
            self.jj_inac = np.concatenate([self.jj_inac, self.jj[m]])
            self.kk_inac = np.concatenate([self.kk_inac, self.kk[m]])
            self.weight_inac = np.concatenate([self.weight_inac, self.weight[m]])
            self.target_inac = np.concatenate([self.target_inac, self.target[m]])
        self.weight, self.target = self.weight[~m], self.target[~m]
        self.ii, self.jj, self.kk, self.net = self.ii[~m], self.jj[~m], self.kk[~m], self.net[~m]

    def remove_frame(self, k):
        """dpvo.py:280-285."""
        self.remove_factors((self.ii == k) | (self.jj == k), store=False)
        self.kk[self.ii > k] -= self.M
        self.ii[self.ii > k] -= 1
        self.jj[self.jj > k] -= 1

    def remove_old(self, n, removal_window, loop_closure=False, optimization_window=0):
        """dpvo.py:304-310."""
        to_remove = self.ix[self.kk] < n - removal_window
        if loop_closure:
            lc = ((self.jj - self.ii) > 30) & (self.jj > (n - optimization_window))
            to_remove = to_remove & ~lc
        self.remove_factors(to_remove, store=True)


def keyframe_plan_counts(ix, M, ii, jj, kk, k, n, removal_window, loop_closure=False, optimization_window=0):
    """What DPVO.keyframe() (dpvo.py:272-310) does to the edge list under BOTH outcomes of its test, as a formula
    evaluated directly on the list (this is what the product's one-read `dpvo_keyframe_plan` must return; the test
    suite checks it against the methods above, which are pinned to the reference's own run):
      no removal:       moved_a  = edges the window rule (:304-309) moves to the inactive list at frame count n,
                        min_a    = smallest source frame that stays (None if nothing stays)
      frame k removed:  dropped_b = edges touching k (:280-281); the others are renumbered (:283-285) and
                        moved_b  = those the window rule moves at frame count n - 1,  min_b as above.
    Returns (moved_a, min_a, dropped_b, moved_b, min_b)."""
    ix = np.asarray(ix, dtype=np.int64)
    ii, jj, kk = (np.asarray(x, dtype=np.int64) for x in (ii, jj, kk))

    def rule(src, a, b, nn):
        rem = src < nn - removal_window
        if loop_closure:
            rem = rem & ~(((b - a) > 30) & (b > (nn - optimization_window)))
        return rem

    rem_a = rule(ix[kk], ii, jj, n)
    min_a = int(ii[~rem_a].min()) if (~rem_a).any() else None
    drop = (ii == k) | (jj == k)
    i2, j2, k2 = ii[~drop].copy(), jj[~drop].copy(), kk[~drop].copy()
    k2[i2 > k] -= M
    i2[i2 > k] -= 1
    j2[j2 > k] -= 1
    rem_b = rule(ix[k2], i2, j2, n - 1)
    min_b = int(i2[~rem_b].min()) if (~rem_b).any() else None
    return int(rem_a.sum()), min_a, int(drop.sum()), int(rem_b.sum()), min_b


def frames_shift(buf, k, n, mod=None):
    """dpvo.py:287-297: for i in range(k, n-1): buf[i % mod] = buf[(i+1) % mod] (in place, in order)."""
    mod = buf.shape[0] if mod is None else mod
    for i in range(k, n - 1):
        buf[i % mod] = buf[(i + 1) % mod]


def reduce_edges(flow_mag, ii, jj, max_num_edges, nms):
    """optim_utils.py:27-64.  Candidates in ascending magnitude (stable), greedy with +-nms suppression."""
    es = []
    if len(ii) == 0:
        return np.zeros((0, 2), dtype=np.int64)
    Ni, Nj = int(ii.max()) + 1, int(jj.max()) + 1
    ignore = np.zeros((Ni, Nj), dtype=bool)
    for idx in np.argsort(flow_mag, kind="stable"):
        if len(es) + 1 > max_num_edges:
            break
        i, j = int(ii[idx]), int(jj[idx])
        if j - i < 30 or flow_mag[idx] >= 1000 or ignore[i, j]:
            continue
        es.append((i, j))
        for di in range(-nms, nms + 1):
            if 0 <= i + di < Ni:
                ignore[i + di, j] = True
    return np.asarray(es, dtype=np.int64).reshape(-1, 2)
