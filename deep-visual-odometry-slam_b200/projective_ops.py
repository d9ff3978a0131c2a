"""`projective_ops` — fused inference path of the reference's dpvo/projective_ops.py (same function
names, argument order and output layouts; projective_ops.py:19-130).

`poses` may be a lietorch `SE3` object (its `.data` is used) or a `[1,N,7]` tensor.  One launch per
call instead of index_select + iproj + lietorch Inv/Mul/Act4 + proj (+ the Jacobian products).
Inference only: there is no autograd through these kernels (the reference's training path keeps
using its differentiable torch ops); calling them with grad-requiring inputs under grad mode raises.
"""
import torch

from ._lib import lib, ptr, stream_of, check, require_cuda, c_i32, c_i64, c_f32, c_p

MIN_DEPTH = 0.2


def _prep(poses, patches, intrinsics, *idx):
    if not torch.is_tensor(poses):
        poses = poses.data            # lietorch SE3 object
    require_cuda(poses, patches, intrinsics, *idx)
    if torch.is_grad_enabled() and any(t.requires_grad for t in (poses, patches, intrinsics)):
        raise RuntimeError("projective_ops (dpvo_b200): inference-only kernels, no autograd; use torch.no_grad()")
    for t, name in ((poses, "poses"), (patches, "patches"), (intrinsics, "intrinsics")):
        if t.dtype != torch.float32:
            raise RuntimeError(f"projective_ops: {name} must be float32")
    poses = poses.detach().contiguous(); patches = patches.detach().contiguous(); intrinsics = intrinsics.detach().contiguous()
    if poses.dim() != 3 or poses.shape[0] != 1 or poses.shape[-1] != 7:
        raise RuntimeError("projective_ops: poses must be [1,N,7]")
    if patches.dim() != 5 or patches.shape[0] != 1 or patches.shape[2] != 3 or patches.shape[3] != patches.shape[4]:
        raise RuntimeError("projective_ops: patches must be [1,K,3,P,P]")
    if intrinsics.dim() != 3 or intrinsics.shape[1] != poses.shape[1] or intrinsics.shape[2] != 4:
        raise RuntimeError("projective_ops: intrinsics must be [1,N,4]")
    out = []
    for t in idx:
        if t.dtype != torch.int64 or t.dim() != 1:
            raise RuntimeError("projective_ops: index tensors must be 1-D int64")
        out.append(t.contiguous())
    return (poses, patches, intrinsics) + tuple(out)


def transform(poses, patches, intrinsics, ii, jj, kk, depth=False, valid=False, jacobian=False, tonly=False):
    poses, patches, intrinsics, ii, jj, kk = _prep(poses, patches, intrinsics, ii, jj, kk)
    E, P = ii.numel(), patches.shape[-1]
    if jj.numel() != E or kk.numel() != E:
        raise RuntimeError("projective_ops.transform: ii/jj/kk lengths differ")
    dev = poses.device
    coords = torch.empty((1, E, P, P, 3 if depth else 2), dtype=torch.float32, device=dev)
    val = torch.empty((1, E, P, P), dtype=torch.float32, device=dev) if (valid and not jacobian) else None
    Ji = Jj = Jz = jv = None
    if jacobian:
        Ji = torch.empty((1, E, 2, 6), dtype=torch.float32, device=dev)
        Jj = torch.empty((1, E, 2, 6), dtype=torch.float32, device=dev)
        Jz = torch.empty((1, E, 2, 1), dtype=torch.float32, device=dev)
        jv = torch.empty((1, E), dtype=torch.float32, device=dev)
    check(lib.dpvo_transform(ptr(poses), ptr(patches), ptr(intrinsics), ptr(ii), ptr(jj), ptr(kk), c_i64(E),
                             c_i64(poses.shape[1]), c_i64(patches.shape[1]), c_i32(P), c_i32(bool(depth)),
                             c_i32(bool(tonly)), ptr(coords), ptr(val), ptr(Ji), ptr(Jj), ptr(Jz), ptr(jv),
                             stream_of(poses)), "projective_ops.transform")
    if jacobian:
        return coords, jv, (Ji, Jj, Jz)
    if valid:
        return coords, val
    return coords


def flow_mag(poses, patches, intrinsics, ii, jj, kk, beta=0.3):
    poses, patches, intrinsics, ii, jj, kk = _prep(poses, patches, intrinsics, ii, jj, kk)
    E, P = ii.numel(), patches.shape[-1]
    flow = torch.empty((1, E, P, P), dtype=torch.float32, device=poses.device)
    val = torch.empty((1, E, P, P), dtype=torch.bool, device=poses.device)
    check(lib.dpvo_flow_mag(ptr(poses), ptr(patches), ptr(intrinsics), ptr(ii), ptr(jj), ptr(kk), c_i64(E),
                            c_i64(poses.shape[1]), c_i64(patches.shape[1]), c_i32(P), c_f32(float(beta)),
                            ptr(flow), ptr(val), stream_of(poses)), "projective_ops.flow_mag")
    return flow, val


def motionmag_pair(poses, patches, intrinsics, ii, jj, kk, i, j, beta=0.5, out=None):
    """`DPVO.motionmag(i, j) + DPVO.motionmag(j, i)` (dpvo.py:257-264, 266-272) as a 0-d device tensor: one launch over
    the live edge list, fixed reduction order, NaN when a direction has no edge.  `out`: optional [1] float32 buffer."""
    poses, patches, intrinsics, ii, jj, kk = _prep(poses, patches, intrinsics, ii, jj, kk)
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=poses.device)
    check(lib.dpvo_motionmag_pair(ptr(poses), ptr(patches), ptr(intrinsics), ptr(ii), ptr(jj), ptr(kk),
                                  c_i64(ii.numel()), c_i64(poses.shape[1]), c_i64(patches.shape[1]),
                                  c_i32(patches.shape[-1]), c_f32(float(beta)), c_i64(int(i)), c_i64(int(j)), ptr(out),
                                  stream_of(poses)), "projective_ops.motionmag_pair")
    return out[0]


PLAN_NONE = 0x7FFFFFFF


def keyframe_plan(poses, patches, intrinsics, ii, jj, kk, ix, i, j, k, n, M, removal_window, loop_closure,
                  optimization_window, beta=0.5, out=None, live_count=None, info=None):
    """`motionmag_pair` plus a dry run of keyframe()'s edge edits under both outcomes of the test (dpvo.py:266-310):
    int32 [8] device tensor, word 0 = the float32 bits of the motion magnitude, words 1..5 = (moved, min ii) without
    removal and (dropped, moved, min ii) with frame k removed; words 6 / 7 echo the edge store's device count (int64 [1]
    `live_count`) and error flag (int32 [1] `info`) when given — see include/dpvo_b200.h `dpvo_keyframe_plan`."""
    poses, patches, intrinsics, ii, jj, kk, ix = _prep(poses, patches, intrinsics, ii, jj, kk, ix)
    if ix.numel() != patches.shape[1]:
        raise RuntimeError("projective_ops.keyframe_plan: ix must have one entry per patch")
    if live_count is not None and (live_count.dtype != torch.int64 or not live_count.is_cuda):
        raise RuntimeError("projective_ops.keyframe_plan: live_count must be a CUDA int64 tensor")
    if info is not None and (info.dtype != torch.int32 or not info.is_cuda):
        raise RuntimeError("projective_ops.keyframe_plan: info must be a CUDA int32 tensor")
    if out is None:
        out = torch.zeros(8, dtype=torch.int32, device=poses.device)
    check(lib.dpvo_keyframe_plan(ptr(poses), ptr(patches), ptr(intrinsics), ptr(ii), ptr(jj), ptr(kk), ptr(ix),
                                 c_i64(ii.numel()), c_i64(poses.shape[1]), c_i64(patches.shape[1]),
                                 c_i32(patches.shape[-1]), c_f32(float(beta)), c_i64(int(i)), c_i64(int(j)),
                                 c_i64(int(k)), c_i64(int(n)), c_i64(int(M)), c_i64(int(removal_window)),
                                 c_i32(1 if loop_closure else 0), c_i64(int(optimization_window)), ptr(live_count),
                                 ptr(info), ptr(out), stream_of(poses)), "projective_ops.keyframe_plan")
    return out


def point_cloud(poses, patches, intrinsics, ix):
    poses, patches, intrinsics, ix = _prep(poses, patches, intrinsics, ix)
    n, P = patches.shape[1], patches.shape[-1]
    if ix.numel() != n:
        raise RuntimeError("projective_ops.point_cloud: ix must have one entry per patch")
    pts = torch.empty((1, n, P, P, 4), dtype=torch.float32, device=poses.device)
    check(lib.dpvo_point_cloud(ptr(poses), ptr(patches), ptr(intrinsics), ptr(ix), c_i64(n), c_i64(poses.shape[1]),
                               c_i32(P), ptr(pts), stream_of(poses)), "projective_ops.point_cloud")
    return pts
