"""`cuda_corr` — host side of the correlation / patchify operators.

Same function names, argument order and return conventions as the reference's
pybind module (/root/reference/methods/dpvo/dpvo/altcorr/correlation.cpp:28-63):

    forward(fmap1, fmap2, coords, ii, jj, radius)            -> [corr]
    backward(fmap1, fmap2, coords, ii, jj, corr_grad, radius) -> [fmap1_grad, fmap2_grad]
    patchify_forward(net, coords, radius)                    -> [patches]
    patchify_backward(net, coords, gradient, radius)         -> [net_grad]

plus the fused two-level lookup `pyramid_forward` used by the B200-native update
loop (one launch for DPVO.corr, dpvo.py:200-207).

Everything here is plumbing: shape/dtype validation, output allocation through
torch's caching allocator and one ctypes call into libdpvo_b200.so on torch's
current stream.
"""
import collections
import ctypes
import os
import weakref

import torch

from . import _lib
from ._lib import lib, CorrLevel, ptr, stream_of, check, dtype_code, require_cuda, c_i32, c_i64, c_p

# ------------------------------------------------------------------ layout cache
# The reference DPVO keeps its feature ring NCHW (dpvo.py:72-73).  The sm_100a
# gather wants channels-last.  A caller that allocates the ring channels-last
# (e.g. torch.zeros(1,N,H,W,C).permute(0,1,4,2,3)) is used zero-copy; for an
# NCHW ring a channels-last shadow is kept per tensor object.  When the tensor's
# version counter has moved (the reference writes one slot per frame and shifts a
# few on keyframe removal, dpvo.py:435-438,296-299), `dpvo_ring_sync` hashes every
# slot of the source in one read pass and re-transposes only the slots whose
# content changed — no host synchronisation, ~1/3 of the traffic of a full refresh.
# Writes that do not move the version counter (raw-pointer kernels, CUDA-graph
# replays, `.data` writes) need `invalidate(fmap)`; tensors without a version
# counter (inference mode) are re-hashed on every call.
_shadow = collections.OrderedDict()  # id(tensor) -> _Shadow  (LRU)
_SHADOW_MAX = 16
shadow_refreshes = 0  # number of sync passes issued (instrumentation for tests / bench)
shadow_hits = 0


class _Shadow:
    __slots__ = ("ref", "version", "data_ptr", "view", "hashes", "parity")


def _version_of(t):
    try:
        return t._version
    except Exception:          # inference-mode tensors do not track a version
        return None


def invalidate(fmap=None):
    """Force the next lookup to re-examine the ring (all rings if `fmap` is None).  Needed only after
    writes torch does not see: raw-pointer kernels, CUDA-graph replays, `.data` assignments."""
    if fmap is None:
        for ent in _shadow.values():
            ent.version = -1
    else:
        ent = _shadow.get(id(fmap))
        if ent is not None:
            ent.version = -1


def shadow_stats():
    return {"entries": len(_shadow), "syncs": shadow_refreshes, "hits": shadow_hits,
            "bytes": sum(e.view.numel() * e.view.element_size() for e in _shadow.values())}


def _channels_last_view(fmap):
    """Return a tensor with the same logical [B,N,C,H,W] shape whose channel stride is 1."""
    global shadow_refreshes, shadow_hits
    if fmap.stride(2) == 1 or fmap.shape[2] == 1:
        return fmap
    key = id(fmap)
    ent = _shadow.get(key)
    ver = _version_of(fmap)
    if ent is not None:
        if ent.ref() is fmap and ent.data_ptr == fmap.data_ptr() and ent.view.shape == fmap.shape:
            _shadow.move_to_end(key)
            if ver is not None and ent.version == ver:
                shadow_hits += 1
                return ent.view
        else:
            _shadow.pop(key, None)
            ent = None
    B, N, C, H, W = fmap.shape
    if ent is None:
        while len(_shadow) >= _SHADOW_MAX:
            _shadow.popitem(last=False)
        ent = _Shadow()
        store = torch.zeros((B, N, H, W, C), dtype=fmap.dtype, device=fmap.device)
        ent.view = store.permute(0, 1, 4, 2, 3)
        ent.ref = weakref.ref(fmap, lambda _r, k=key: _shadow.pop(k, None))
        ent.hashes = torch.full((2 * B * N,), -1, dtype=torch.int64, device=fmap.device)   # 0xFF..: matches nothing
        ent.parity = 0
        _shadow[key] = ent
    src = fmap.detach()
    if not src.is_contiguous():
        src = src.contiguous()
    slot_bytes = C * H * W * fmap.element_size()
    with _lib.device_guard(fmap):
        if slot_bytes % 16 == 0 and src.data_ptr() % 16 == 0:
            ent.parity ^= 1
            check(lib.dpvo_ring_sync(dtype_code(fmap), ptr(src), ptr(ent.view), c_i64(B * N), c_i32(C), c_i64(H * W),
                                     ptr(ent.hashes), c_i32(ent.parity), stream_of(fmap)), "ring_sync")
        else:
            check(lib.dpvo_nchw_to_nhwc(dtype_code(fmap), ptr(src), ptr(ent.view), c_i64(B * N), c_i32(C), c_i64(H * W),
                                        stream_of(fmap)), "nchw_to_nhwc")
    shadow_refreshes += 1
    ent.version = ver if ver is not None else -1
    ent.data_ptr = fmap.data_ptr()
    return ent.view


def _check_corr_args(fmap1, fmap2, coords, ii, jj):
    require_cuda(fmap1, fmap2, coords, ii, jj)
    if fmap1.dim() != 5 or fmap2.dim() != 5 or coords.dim() != 5:
        raise RuntimeError("cuda_corr: fmap1, fmap2 and coords must be 5-D ([B,N,C,H,W] / [B,M,2,H,W])")
    if fmap1.dtype != fmap2.dtype:
        raise RuntimeError("cuda_corr: fmap1 and fmap2 must have the same dtype")
    if coords.dtype != torch.float32:
        raise RuntimeError("cuda_corr: coords must be float32")
    if coords.shape[2] != 2:
        raise RuntimeError("cuda_corr: coords must be [B,M,2,H,W]")
    if ii.dtype != torch.int64 or jj.dtype != torch.int64 or ii.dim() != 1 or jj.dim() != 1:
        raise RuntimeError("cuda_corr: ii and jj must be 1-D int64")
    if ii.numel() != coords.shape[1] or jj.numel() != coords.shape[1]:
        raise RuntimeError("cuda_corr: ii/jj length must equal coords.shape[1]")
    if fmap1.shape[0] != coords.shape[0] or fmap2.shape[0] != coords.shape[0]:
        raise RuntimeError("cuda_corr: batch sizes differ")
    if fmap1.shape[2] != fmap2.shape[2]:
        raise RuntimeError("cuda_corr: channel counts differ")
    if tuple(fmap1.shape[3:]) != tuple(coords.shape[3:]):
        raise RuntimeError("cuda_corr: fmap1 spatial size must equal the patch size of coords")


def _wants_fast(fmap1, radius):
    return (fmap1.dtype == torch.float16 and fmap1.shape[2] == 128 and tuple(fmap1.shape[3:]) == (3, 3)
            and radius == 3)


def _level(fmap2, scale, out, out_edge_stride, out_elem_stride):
    lv = CorrLevel()
    lv.fmap = fmap2.data_ptr()
    lv.sb, lv.sn, lv.sc, lv.sy, lv.sx = fmap2.stride()
    lv.N2, lv.H2, lv.W2 = fmap2.shape[1], fmap2.shape[3], fmap2.shape[4]
    lv.coord_scale = scale
    lv.out = out
    lv.out_edge_stride = out_edge_stride
    lv.out_elem_stride = out_elem_stride
    return lv


# Locality order (csrc/corr_order.cu).  Measured on B200 (profiles/r02_time_corr_locality_order.jsonl,
# r02_time_corr_bwd_locality_order.jsonl): the BACKWARD gains 20 % at 18k edges including the sort (the fp32 gradient
# lines it reduces into stay in L2), so it is on; the FORWARD lookup kernel gains 8 % (0.267 -> 0.246 ms at 45,312 edges)
# which the 0.026 ms sort gives back — the lookup is bound by bytes in flight, not by where they come from — so it is
# off unless a caller already has a permutation to pass.  The output is bit-identical either way.
ORDER_FORWARD = os.environ.get("DPVO_CORR_ORDER_FWD", "0") != "0"
ORDER_EDGES = os.environ.get("DPVO_CORR_ORDER", "1") != "0"      # backward
ORDER_MIN_EDGES = 4096
lib.dpvo_corr_order_workspace_bytes.restype = c_i64


def edge_order(coords, jj, n_slots, h, w, tile):
    """int32 permutation of the B = 1 edge list grouping edges by (target-frame slot, tile x tile pixel tile of the patch
    centre); coords in the pixel units of the (h, w) map.  Device side, no synchronisation."""
    M = coords.shape[1]
    perm = torch.empty((M,), dtype=torch.int32, device=coords.device)
    nbytes = int(lib.dpvo_corr_order_workspace_bytes(c_i64(M), c_i32(n_slots), c_i32(h), c_i32(w), c_i32(tile)))
    ws = torch.empty((max(nbytes, 1),), dtype=torch.uint8, device=coords.device)
    with _lib.device_guard(coords):
        check(lib.dpvo_corr_order(ptr(coords), ptr(jj), c_i64(M), c_i32(n_slots), c_i32(h), c_i32(w), c_i32(tile),
                                  ptr(perm), ptr(ws), c_i64(nbytes), stream_of(coords)), "cuda_corr.edge_order")
    return perm


def _launch(fmap1, coords, ii, jj, radius, levels, perm=None):
    B, M = coords.shape[0], coords.shape[1]
    arr = (CorrLevel * len(levels))(*levels)
    g = fmap1.stride()
    with _lib.device_guard(fmap1):
        check(lib.dpvo_corr_forward_ordered(
            c_i32(dtype_code(fmap1)), ptr(fmap1), c_i64(g[0]), c_i64(g[1]), c_i64(g[2]), c_i64(g[3]), c_i64(g[4]),
            c_i32(fmap1.shape[1]), c_i32(fmap1.shape[2]), c_i32(fmap1.shape[3]), c_i32(fmap1.shape[4]),
            ptr(coords), ptr(ii), ptr(jj), c_i32(B), c_i32(M), c_i32(radius),
            arr, c_i32(len(levels)), ptr(perm), stream_of(fmap1)), "cuda_corr.forward")


def forward(fmap1, fmap2, coords, ii, jj, radius):
    """corr[B,M,7,7,P,P] (x-offset major, as the reference's permuted view) in a list."""
    _check_corr_args(fmap1, fmap2, coords, ii, jj)
    radius = int(radius)
    coords = coords.contiguous()
    ii = ii.contiguous()
    jj = jj.contiguous()
    B, M, _, PH, PW = coords.shape
    DO = 2 * radius + 1
    out = torch.empty((B, M, DO, DO, PH, PW), dtype=fmap1.dtype, device=fmap1.device)
    if B * M == 0:
        return [out]
    perm = None
    if _wants_fast(fmap1, radius):
        fmap2 = _channels_last_view(fmap2)
        if not _dense_patch_blob(fmap1):
            fmap1 = fmap1.contiguous()
        if ORDER_FORWARD and B == 1 and M >= ORDER_MIN_EDGES:
            H2, W2 = fmap2.shape[3], fmap2.shape[4]
            perm = edge_order(coords, jj, fmap2.shape[1], H2, W2, 32 if max(H2, W2) > 64 else 8)
    lv = _level(fmap2, 1.0, out.data_ptr(), DO * DO * PH * PW, 1)
    _launch(fmap1, coords, ii, jj, radius, [lv], perm)
    return [out]


def _dense_patch_blob(fmap1):
    _, _, C, PH, PW = fmap1.shape
    s = fmap1.stride()
    if min(s[2], s[3], s[4]) <= 0:
        return False
    span = (C - 1) * s[2] + (PH - 1) * s[3] + (PW - 1) * s[4] + 1
    return span == C * PH * PW and s[1] % 8 == 0 and s[0] % 8 == 0 and fmap1.data_ptr() % 16 == 0


def pyramid_forward(fmap1, pyramid, scales, coords, ii, jj, radius, row_stride=None, out=None):
    """Fused multi-level lookup: returns [B, M, 49*P*P*L] with the level index fastest,
    i.e. torch.stack([corr_l ...], -1).view(B, M, -1) of the per-level results
    (dpvo.py:203-207, net.py:165-169).  One launch for two levels.  `row_stride` (elements, even,
    >= the row length) makes the result a view of a [B, M, row_stride] buffer — 896 gives the
    16-byte-aligned rows the Update operator's first layer loads with TMA (update.py)."""
    if len(pyramid) != len(scales) or len(pyramid) not in (1, 2):
        raise RuntimeError("pyramid_forward: 1 or 2 levels supported")
    for fm in pyramid:
        _check_corr_args(fmap1, fm, coords, ii, jj)
    radius = int(radius)
    coords = coords.contiguous()
    ii = ii.contiguous()
    jj = jj.contiguous()
    B, M, _, PH, PW = coords.shape
    DO = 2 * radius + 1
    L = len(pyramid)
    per = DO * DO * PH * PW
    stride = per * L if row_stride is None else int(row_stride)
    if stride < per * L or stride % 2:
        raise RuntimeError(f"pyramid_forward: row_stride must be even and >= {per * L}")
    if out is not None:
        # caller-owned buffer [B, >= M, stride] (a persistent one keeps the consumer's pointer signature stable)
        if out.dim() != 3 or out.shape[0] != B or out.shape[1] < M or out.shape[2] != stride or out.dtype != fmap1.dtype \
                or not out.is_contiguous() or out.device != fmap1.device:
            raise RuntimeError(f"pyramid_forward: out must be a contiguous [{B}, >={M}, {stride}] tensor of the feature dtype")
        out = out[:, :M, :per * L]
    else:
        out = torch.empty((B, M, stride), dtype=fmap1.dtype, device=fmap1.device)[:, :, :per * L]
    if B * M == 0:
        return out
    perm = None
    if _wants_fast(fmap1, radius):
        pyramid = [_channels_last_view(fm) for fm in pyramid]
        if not _dense_patch_blob(fmap1):
            fmap1 = fmap1.contiguous()
        if ORDER_FORWARD and B == 1 and M >= ORDER_MIN_EDGES and float(scales[0]) == 1.0:
            H1, W1 = pyramid[0].shape[3], pyramid[0].shape[4]
            perm = edge_order(coords, jj, pyramid[0].shape[1], H1, W1, 32)
    esz = out.element_size()
    levels = [_level(fm, float(sc), out.data_ptr() + l * esz, stride, L)
              for l, (fm, sc) in enumerate(zip(pyramid, scales))]
    _launch(fmap1, coords, ii, jj, radius, levels, perm)
    return out


def patchify_forward(net, coords, radius):
    require_cuda(net, coords)
    if net.dim() != 4 or coords.dim() != 3 or coords.shape[2] != 2:
        raise RuntimeError("cuda_corr.patchify_forward: net must be [B,C,H,W] and coords [B,M,2]")
    if coords.dtype != torch.float32:
        raise RuntimeError("cuda_corr.patchify_forward: coords must be float32")
    if coords.shape[0] != net.shape[0]:
        raise RuntimeError("cuda_corr.patchify_forward: batch sizes differ")
    radius = int(radius)
    coords = coords.contiguous()
    B, C, H, W = net.shape
    M = coords.shape[1]
    D = 2 * radius + 2
    out = torch.empty((B, M, C, D, D), dtype=net.dtype, device=net.device)
    s = net.stride()
    check(lib.dpvo_patchify_forward(
        c_i32(dtype_code(net)), ptr(net), c_i64(s[0]), c_i64(s[1]), c_i64(s[2]), c_i64(s[3]),
        c_i32(B), c_i32(C), c_i32(H), c_i32(W), ptr(coords), c_i32(M), c_i32(radius), ptr(out),
        stream_of(net)), "cuda_corr.patchify_forward")
    return [out]


def patchify_backward(net, coords, gradient, radius):
    require_cuda(net, coords, gradient)
    if net.dim() != 4 or coords.dim() != 3 or gradient.dim() != 5:
        raise RuntimeError("cuda_corr.patchify_backward: bad ranks")
    radius = int(radius)
    coords = coords.float().contiguous()
    gradient = gradient.to(net.dtype).contiguous()
    B, C, H, W = net.shape
    M = coords.shape[1]
    D = 2 * radius + 2
    if tuple(gradient.shape) != (B, M, C, D, D):
        raise RuntimeError("cuda_corr.patchify_backward: gradient must be [B,M,C,D,D]")
    grad = torch.zeros((B, C, H, W), dtype=net.dtype, device=net.device)
    check(lib.dpvo_patchify_backward(
        c_i32(dtype_code(net)), ptr(gradient), ptr(coords), c_i32(B), c_i32(M), c_i32(C), c_i32(H), c_i32(W),
        c_i32(radius), ptr(grad), stream_of(net)), "cuda_corr.patchify_backward")
    return [grad]


def _backward_fast_ok(fmap1, fmap2, coords, radius):
    return (fmap1.dtype in (torch.float16, torch.float32) and fmap2.dtype == fmap1.dtype and radius == 3
            and fmap1.shape[2] == 128 and tuple(fmap1.shape[3:]) == (3, 3) and tuple(coords.shape[3:]) == (3, 3))


def backward(fmap1, fmap2, coords, ii, jj, corr_grad, radius):
    """[fmap1_grad, fmap2_grad].  fp16 / C=128 / 3x3 / radius 3 (the training shape, net.py:156-169) takes the
    warp-per-edge kernel over the channels-last ring (the forward's shadow is reused) and returns fmap2_grad as a
    [B,N,C,H,W] VIEW of a channels-last buffer; everything else takes the generic kernel."""
    _check_corr_args(fmap1, fmap2, coords, ii, jj)
    radius = int(radius)
    coords = coords.contiguous()
    ii = ii.contiguous()
    jj = jj.contiguous()
    B, M, _, PH, PW = coords.shape
    DO = 2 * radius + 1
    if tuple(corr_grad.shape) != (B, M, DO, DO, PH, PW):
        raise RuntimeError("cuda_corr.backward: corr_grad must be [B,M,2r+1,2r+1,P,P]")
    corr_grad = corr_grad.to(fmap1.dtype)
    cg = corr_grad.stride()
    if _backward_fast_ok(fmap1, fmap2, coords, radius) and not os.environ.get("DPVO_CORR_BWD_GENERIC"):
        f2 = _channels_last_view(fmap2.detach())
        f1 = fmap1.detach()
        if tuple(f1.stride()[2:]) != (9, 3, 1) or f1.stride(1) % 4 or f1.stride(0) % 4 or f1.data_ptr() % 16:
            f1 = f1.contiguous()
        _, N1, C = f1.shape[:3]
        _, N2, _, H2, W2 = f2.shape
        dev = fmap1.device
        half = fmap1.dtype == torch.float16
        alloc = torch.empty if half else torch.zeros
        g1 = alloc((B, N1, C, 3, 3), dtype=fmap1.dtype, device=dev)
        g2 = alloc((B, N2, H2, W2, C), dtype=fmap1.dtype, device=dev)
        if B * M == 0:
            return [g1.zero_(), g2.zero_().permute(0, 1, 4, 2, 3)]
        acc1 = torch.zeros((B * N1 * C * 9,), dtype=torch.float32, device=dev) if half else None
        acc2 = torch.zeros((B * N2 * H2 * W2 * C,), dtype=torch.float32, device=dev) if half else None
        s1, s2 = f1.stride(), f2.stride()
        perm = None
        if ORDER_EDGES and B == 1 and M >= 2048:
            perm = edge_order(coords, jj, N2, H2, W2, 32 if max(H2, W2) > 64 else 8)
        with _lib.device_guard(fmap1):
            check(lib.dpvo_corr_backward_nhwc(
                c_i32(dtype_code(fmap1)), ptr(f1), c_i64(s1[0]), c_i64(s1[1]),
                ptr(f2), c_i64(s2[0]), c_i64(s2[1]), c_i64(s2[3]), c_i64(s2[4]),
                c_i32(N1), c_i32(N2), c_i32(H2), c_i32(W2),
                ptr(coords), ptr(ii), ptr(jj), c_i32(B), c_i32(M),
                ptr(corr_grad), c_i64(cg[0]), c_i64(cg[1]), c_i64(cg[2]), c_i64(cg[3]), c_i64(cg[4]), c_i64(cg[5]),
                ptr(g1), ptr(g2), ptr(acc1), ptr(acc2), ptr(perm), stream_of(fmap1)), "cuda_corr.backward")
        return [g1, g2.permute(0, 1, 4, 2, 3)]
    g1 = torch.zeros(fmap1.shape, dtype=fmap1.dtype, device=fmap1.device)
    g2 = torch.zeros(fmap2.shape, dtype=fmap2.dtype, device=fmap2.device)
    if B * M == 0:
        return [g1, g2]
    acc1 = acc2 = None
    if fmap1.dtype == torch.float16:
        acc1 = torch.zeros(fmap1.shape, dtype=torch.float32, device=fmap1.device)
        acc2 = torch.zeros(fmap2.shape, dtype=torch.float32, device=fmap2.device)
    s1, s2 = fmap1.stride(), fmap2.stride()
    with _lib.device_guard(fmap1):
        check(lib.dpvo_corr_backward(
            c_i32(dtype_code(fmap1)),
            ptr(fmap1), c_i64(s1[0]), c_i64(s1[1]), c_i64(s1[2]), c_i64(s1[3]), c_i64(s1[4]),
            ptr(fmap2), c_i64(s2[0]), c_i64(s2[1]), c_i64(s2[2]), c_i64(s2[3]), c_i64(s2[4]),
            c_i32(fmap1.shape[1]), c_i32(fmap2.shape[1]), c_i32(fmap1.shape[2]), c_i32(PH), c_i32(PW),
            c_i32(fmap2.shape[3]), c_i32(fmap2.shape[4]),
            ptr(coords), ptr(ii), ptr(jj), c_i32(B), c_i32(M), c_i32(radius),
            ptr(corr_grad), c_i64(cg[0]), c_i64(cg[1]), c_i64(cg[2]), c_i64(cg[3]), c_i64(cg[4]), c_i64(cg[5]),
            ptr(g1), ptr(g2), ptr(acc1), ptr(acc2), stream_of(fmap1)), "cuda_corr.backward")
    return [g1, g2]
