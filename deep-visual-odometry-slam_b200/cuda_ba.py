"""`cuda_ba` — host side of the bundle-adjustment operators.

Same function names, argument order and conventions as the reference's pybind module
(/root/reference/methods/dpvo/dpvo/fastba/ba.cpp:32-97,183-189):

    forward(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk,
            PPF, t0, t1, iterations, eff_impl) -> []      (poses / patches updated in place)
    neighbors(ii, jj) -> [ix, jx]
    reproject(poses, patches, intrinsics, ii, jj, kk) -> coords[1,E,2,P,P]

Unlike the reference nothing here synchronises with the host: `neighbors` runs on the
device, `forward` derives the sorted-unique patch list on the device, and all launches go
to torch's current stream.
"""
import torch

from ._lib import lib, ptr, stream_of, check, require_cuda, c_i32, c_i64, c_f32

lib.dpvo_ba_workspace_bytes.restype = c_i64
lib.dpvo_neighbors_workspace_bytes.restype = c_i64
lib.dpvo_unique_workspace_bytes.restype = c_i64

_ws = {}  # device index -> grow-only uint8 workspace (reused on the same stream order)


def _workspace(device, nbytes):
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    t = _ws.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=device)
        _ws[key] = t
    return t


def _f32c(t, name):
    if t.dtype != torch.float32:
        raise RuntimeError(f"cuda_ba: {name} must be float32")
    if not t.is_contiguous():
        raise RuntimeError(f"cuda_ba: {name} must be contiguous")
    return t


def _i64c(t, name):
    if t.dtype != torch.int64 or t.dim() != 1:
        raise RuntimeError(f"cuda_ba: {name} must be 1-D int64")
    return t.contiguous()


def forward(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, PPF, t0, t1, iterations, eff_impl,
            prepared=None):
    """cuda_ba.forward (ba.cpp:32-57).  `prepared`: the workspace returned by `prepare(...)` for the same
    edge list and window — the edge-list-only part of the call is then skipped (it may have run on
    another stream while the inputs of this call were still being produced)."""
    require_cuda(poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk)
    _f32c(poses, "poses"); _f32c(patches, "patches")
    intrinsics = _f32c(intrinsics, "intrinsics")
    target = _f32c(target.contiguous(), "target")
    weight = _f32c(weight.contiguous(), "weight")
    lmbda = _f32c(lmbda.reshape(-1), "lmbda")
    ii = _i64c(ii, "ii"); jj = _i64c(jj, "jj"); kk = _i64c(kk, "kk")
    E = ii.numel()
    if jj.numel() != E or kk.numel() != E or target.numel() != 2 * E or weight.numel() != 2 * E:
        raise RuntimeError("cuda_ba.forward: ii/jj/kk/target/weight lengths differ")
    P = patches.shape[-1]
    if patches.shape[-2] != P or patches.shape[-3] != 3:
        raise RuntimeError("cuda_ba.forward: patches must be [...,3,P,P]")
    n_frames = poses.numel() // 7
    n_patches = patches.numel() // (3 * P * P)
    t0, t1, iterations, PPF = int(t0), int(t1), int(iterations), int(PPF)
    nbytes = lib.dpvo_ba_workspace_bytes(c_i64(E), c_i64(n_patches), c_i32(max(t1 - t0, 0)), c_i32(bool(eff_impl)),
                                         c_i32(PPF), c_i64(n_frames))
    if prepared is not None:
        ws, fn = prepared, lib.dpvo_ba_forward_prepared
        if ws.numel() < nbytes or ws.device != poses.device:
            raise RuntimeError("cuda_ba.forward: `prepared` workspace does not belong to this problem")
    else:
        ws, fn = _workspace(poses.device, nbytes), lib.dpvo_ba_forward
    check(fn(ptr(poses), ptr(patches), ptr(intrinsics), ptr(target), ptr(weight), ptr(lmbda),
             ptr(ii), ptr(jj), ptr(kk), c_i64(E), c_i64(n_frames), c_i64(n_patches), c_i32(P),
             c_i32(PPF), c_i32(t0), c_i32(t1), c_i32(iterations), c_i32(bool(eff_impl)),
             ptr(ws), c_i64(ws.numel()), stream_of(poses)), "cuda_ba.forward")
    return []


def prepare(ii, jj, kk, n_frames, n_patches, P, PPF, t0, t1, eff_impl=False, workspace=None):
    """The part of `forward` that only depends on the edge list and the window [t0, t1): the set of
    patches seen, their compact numbering, the per-patch edge lists and the cleared accumulators.
    Runs on torch's current stream; returns the workspace to hand to forward(..., prepared=ws).
    Pass the previous call's workspace back in to reuse the allocation."""
    require_cuda(ii, jj, kk)
    ii = _i64c(ii, "ii"); jj = _i64c(jj, "jj"); kk = _i64c(kk, "kk")
    E = ii.numel()
    if jj.numel() != E or kk.numel() != E:
        raise RuntimeError("cuda_ba.prepare: ii/jj/kk lengths differ")
    t0, t1, PPF = int(t0), int(t1), int(PPF)
    nbytes = lib.dpvo_ba_workspace_bytes(c_i64(E), c_i64(int(n_patches)), c_i32(max(t1 - t0, 0)), c_i32(bool(eff_impl)),
                                         c_i32(PPF), c_i64(int(n_frames)))
    ws = workspace
    if ws is None or ws.numel() < nbytes or ws.device != ii.device:
        ws = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=ii.device)
    check(lib.dpvo_ba_prepare(ptr(ii), ptr(jj), ptr(kk), c_i64(E), c_i64(int(n_frames)), c_i64(int(n_patches)),
                              c_i32(int(P)), c_i32(PPF), c_i32(t0), c_i32(t1), c_i32(bool(eff_impl)), ptr(ws),
                              c_i64(ws.numel()), stream_of(ii)), "cuda_ba.prepare")
    return ws


def neighbors(ii, jj):
    require_cuda(ii, jj)
    ii = _i64c(ii, "ii"); jj = _i64c(jj, "jj")
    n = ii.numel()
    if jj.numel() != n:
        raise RuntimeError("cuda_ba.neighbors: ii and jj lengths differ")
    ix = torch.empty(n, dtype=torch.int64, device=ii.device)
    jx = torch.empty(n, dtype=torch.int64, device=ii.device)
    if n:
        ws = _workspace(ii.device, lib.dpvo_neighbors_workspace_bytes(c_i64(n)))
        check(lib.dpvo_neighbors(ptr(ii), ptr(jj), c_i64(n), ptr(ix), ptr(jx), ptr(ws), c_i64(ws.numel()),
                                 stream_of(ii)), "cuda_ba.neighbors")
    return [ix, jx]


def reproject(poses, patches, intrinsics, ii, jj, kk):
    require_cuda(poses, patches, intrinsics, ii, jj, kk)
    _f32c(poses, "poses"); _f32c(patches, "patches"); _f32c(intrinsics, "intrinsics")
    ii = _i64c(ii, "ii"); jj = _i64c(jj, "jj"); kk = _i64c(kk, "kk")
    P = patches.shape[-1]
    E = ii.numel()
    coords = torch.empty((1, E, 2, P, P), dtype=torch.float32, device=poses.device)
    check(lib.dpvo_reproject(ptr(poses), ptr(patches), ptr(intrinsics), ptr(ii), ptr(jj), ptr(kk), c_i64(E),
                             c_i32(P), ptr(coords), stream_of(poses)), "cuda_ba.reproject")
    return coords


def unique_inverse(keys, key_bound):
    """Sorted unique + inverse of bounded non-negative int64 keys, no host sync.
    Returns (uniq[padded to len(keys)], inverse, n_uniq device int32 tensor)."""
    require_cuda(keys)
    keys = _i64c(keys, "keys")
    n = keys.numel()
    uniq = torch.empty(min(n, int(key_bound)), dtype=torch.int64, device=keys.device)
    inv = torch.empty(n, dtype=torch.int64, device=keys.device)
    cnt = torch.zeros(1, dtype=torch.int32, device=keys.device)
    ws = _workspace(keys.device, lib.dpvo_unique_workspace_bytes(c_i64(int(key_bound))))
    check(lib.dpvo_unique_inverse(ptr(keys), c_i64(n), c_i64(int(key_bound)), ptr(uniq), ptr(inv), ptr(cnt),
                                  ptr(ws), c_i64(ws.numel()), stream_of(keys)), "cuda_ba.unique_inverse")
    return uniq, inv, cnt


def cholesky_solve(S, y):
    """x = S^-1 y for a dense SPD float32 system with the BA's blocked Cholesky.  Returns (x, info);
    info is a device int32 tensor (0 = ok, 1 = non-positive pivot, x zeroed).  No host sync."""
    require_cuda(S, y)
    S = _f32c(S, "S"); y = _f32c(y.reshape(-1), "y")
    n = y.numel()
    if S.dim() != 2 or S.shape[0] != n or S.shape[1] != n:
        raise RuntimeError("cuda_ba.cholesky_solve: S must be [n,n] and y [n]")
    x = torch.zeros(n, dtype=torch.float32, device=S.device)
    info = torch.zeros(1, dtype=torch.int32, device=S.device)
    lib.dpvo_chol_workspace_bytes.restype = c_i64
    ws = _workspace(S.device, lib.dpvo_chol_workspace_bytes(c_i32(n)))
    check(lib.dpvo_chol_solve(ptr(S), ptr(y), c_i32(n), ptr(x), ptr(info), ptr(ws), c_i64(ws.numel()), stream_of(S)),
          "cuda_ba.cholesky_solve")
    return x, info


def chol_bulk_mode(mode):
    """0: split-TF32 tensor-core tiles in the blocked Cholesky's bulk trailing update (default), 1: fp32 FMA tiles
    (include/dpvo_b200.h: dpvo_chol_bulk_mode).  Process-wide; for A/B measurements and the accuracy test."""
    check(lib.dpvo_chol_bulk_mode(c_i32(int(mode))), "cuda_ba.chol_bulk_mode")


def solve_system(J_Ginv_i, J_Ginv_j, ii, jj, res, ep, lm, freen, n_poses=None, refinements=2, return_info=False):
    """cuda_ba.solve_system (fastba/ba.cpp:120-180; caller loop_closure/optim_utils.py:229): one damped
    Gauss-Newton step of the Sim(3) pose graph, on the device.  Returns [delta [n,7]] like the reference
    ([delta, info] with return_info=True; info is a device int32 tensor: 0 ok, 1 factorisation failed,
    2 self edge).  n = max(ii, jj) + 1 is read back from the index tensors like the reference does
    (one host sync) unless `n_poses` is given."""
    require_cuda(J_Ginv_i, J_Ginv_j, ii, jj, res)
    Ji = _f32c(J_Ginv_i, "J_Ginv_i"); Jj = _f32c(J_Ginv_j, "J_Ginv_j"); res = _f32c(res, "res")
    ii = _i64c(ii, "ii"); jj = _i64c(jj, "jj")
    r = ii.numel()
    if jj.numel() != r or Ji.shape != (r, 7, 7) or Jj.shape != (r, 7, 7) or res.numel() != r * 7:
        raise RuntimeError("cuda_ba.solve_system: expected J_Ginv_i/j [r,7,7], ii/jj [r], res [r,7]")
    if n_poses is None:
        n_poses = (int(torch.maximum(ii.max(), jj.max()).item()) + 1) if r else 0
    n_poses = int(n_poses)
    delta = torch.empty((n_poses, 7), dtype=torch.float32, device=res.device)
    info = torch.zeros(1, dtype=torch.int32, device=res.device)
    lib.dpvo_pgo_workspace_bytes.restype = c_i64
    nbytes = lib.dpvo_pgo_workspace_bytes(c_i32(n_poses), c_i32(int(freen)))
    ws = torch.empty(int(nbytes), dtype=torch.uint8, device=res.device)   # large and rarely used: not cached
    check(lib.dpvo_pgo_solve(ptr(Ji), ptr(Jj), ptr(ii), ptr(jj), ptr(res), c_i64(r), c_i32(n_poses), c_f32(float(ep)),
                             c_f32(float(lm)), c_i32(int(freen)), c_i32(int(refinements)), ptr(delta), ptr(info),
                             ptr(ws), c_i64(ws.numel()), stream_of(res)), "cuda_ba.solve_system")
    return [delta, info] if return_info else [delta]
