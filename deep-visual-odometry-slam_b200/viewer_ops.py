"""Headless mirror of DPViewer's CUDA helpers (SURVEY.md §8 row N5;
modules/DPViewer/dpviewer/viewer_cuda.cu:166-219 `backproject_and_filter`, :268-277 `poseToMatrix`).
The window, the GL buffers and the CUDA-GL interop of viewer.cpp are not built (unavailable on this box):
results are ordinary device tensors.  No CPU fallback."""
import torch

from ._lib import lib, ptr, stream_of, check, require_cuda, c_i32, c_i64, c_f32


def poseToMatrix(poses):
    """[n,7] float32 poses -> [n,4,4] matrices of the inverse poses (what viewer.cpp:271 draws cameras with)."""
    require_cuda(poses)
    if poses.dtype != torch.float32 or poses.dim() != 2 or poses.shape[1] != 7:
        raise RuntimeError("viewer_ops.poseToMatrix: poses must be a float32 [n,7] tensor")
    poses = poses.contiguous()
    out = torch.empty((poses.shape[0], 4, 4), dtype=torch.float32, device=poses.device)
    check(lib.dpvo_viewer_pose_to_matrix(ptr(poses), ptr(out), c_i64(poses.shape[0]), stream_of(poses)),
          "viewer_ops.poseToMatrix")
    return out


def iproj(index, nFrames, thresh, images, poses, disps, intrinsics):
    """The kernel alone: (points [ht*wd,3] f32, colors [ht*wd,3] u8, count [ht*wd] f32)."""
    require_cuda(images, poses, disps, intrinsics)
    if images.dtype != torch.uint8 or images.dim() != 4 or images.shape[1] != 3:
        raise RuntimeError("viewer_ops.iproj: images must be uint8 [N,3,ht,wd]")
    if disps.dtype != torch.float32 or disps.dim() != 3 or tuple(images.shape[2:]) != tuple(disps.shape[1:]):
        raise RuntimeError("viewer_ops.iproj: disps must be float32 [N,ht,wd] matching images")
    if poses.dtype != torch.float32 or intrinsics.dtype != torch.float32:
        raise RuntimeError("viewer_ops.iproj: poses / intrinsics must be float32")
    N = min(images.shape[0], poses.shape[0], disps.shape[0], intrinsics.shape[0])
    if not (0 <= index < nFrames <= N):
        raise RuntimeError("viewer_ops.iproj: need 0 <= index < nFrames <= number of stored frames")
    images, poses, disps, intrinsics = (t.contiguous() for t in (images, poses, disps, intrinsics))
    ht, wd = disps.shape[1:]
    dev = disps.device
    points = torch.empty((ht * wd, 3), dtype=torch.float32, device=dev)
    colors = torch.empty((ht * wd, 3), dtype=torch.uint8, device=dev)
    count = torch.empty((ht * wd,), dtype=torch.float32, device=dev)
    check(lib.dpvo_viewer_iproj(c_i32(index), c_i32(nFrames), c_f32(float(thresh)), ptr(images), ptr(poses), ptr(disps),
                                ptr(intrinsics), c_i32(ht), c_i32(wd), ptr(points), ptr(colors), ptr(count),
                                stream_of(disps)), "viewer_ops.iproj")
    return points, colors, count


def backproject_and_filter(index, nFrames, thresh, showForeground, showBackground, images, poses, disps, masks,
                           intrinsics):
    """viewer_cuda.cu:166-219 -> (mPoints, points, colors)."""
    points, colors, count = iproj(index, nFrames, thresh, images, poses, disps, intrinsics)
    m = masks[index].reshape(-1)
    sel = []
    if showForeground:
        sel.append((count >= 0) & (m < 0.5))
    if showBackground:
        sel.append((count >= 2.0) & (m > 0.5))
    pts = torch.cat([points[s] for s in sel], 0) if sel else points[:0]
    col = torch.cat([colors[s] for s in sel], 0) if sel else colors[:0]
    return pts.shape[0], pts, col
