"""`altcorr` — Python operator surface of the correlation path.

Mirrors /root/reference/methods/dpvo/dpvo/altcorr/correlation.py (same names,
argument meaning and autograd behaviour) on top of this package's `cuda_corr`:

    corr(fmap1, fmap2, coords, ii, jj, radius=1, dropout=1)   correlation.py:77-80
    patchify(net, coords, radius, mode='bilinear')            correlation.py:55-74
    CorrLayer / PatchLayer autograd Functions                 correlation.py:4-53

`corr_pyramid` is the fused two-level form of DPVO.corr (dpvo.py:200-207).
"""
import torch

from . import cuda_corr


class CorrLayer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, fmap1, fmap2, coords, ii, jj, radius, dropout):
        coords = coords.float()
        ctx.save_for_backward(fmap1, fmap2, coords, ii, jj)
        ctx.radius = radius
        ctx.dropout = dropout
        out, = cuda_corr.forward(fmap1, fmap2, coords, ii, jj, radius)
        return out

    @staticmethod
    def backward(ctx, grad):
        fmap1, fmap2, coords, ii, jj = ctx.saved_tensors
        if ctx.dropout < 1:
            # Bernoulli edge keep-mask, as correlation.py:22-27
            keep = torch.rand(len(ii), device=ii.device) < ctx.dropout
            coords = coords[:, keep]
            grad = grad[:, keep]
            ii = ii[keep]
            jj = jj[keep]
        g1, g2 = cuda_corr.backward(fmap1, fmap2, coords, ii, jj, grad, ctx.radius)
        return g1, g2, None, None, None, None, None


class PatchLayer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, net, coords, radius):
        coords = coords.float()
        ctx.radius = radius
        ctx.save_for_backward(net, coords)
        out, = cuda_corr.patchify_forward(net, coords, radius)
        return out

    @staticmethod
    def backward(ctx, grad):
        net, coords = ctx.saved_tensors
        g, = cuda_corr.patchify_backward(net, coords, grad, ctx.radius)
        return g, None, None


def patchify(net, coords, radius, mode='bilinear'):
    """Extract (2r+1)^2 (bilinear) or (2r+2)^2 (nearest) patches around coords."""
    coords = coords.float()
    patches = PatchLayer.apply(net, coords, radius)
    if mode == 'bilinear':
        frac = (coords - coords.floor()).to(net.dtype).to(net.device)
        dx, dy = frac[:, :, None, None, None].unbind(dim=-1)
        d = 2 * radius + 1
        x00 = (1 - dy) * (1 - dx) * patches[..., :d, :d]
        x01 = (1 - dy) * (dx) * patches[..., :d, 1:]
        x10 = (dy) * (1 - dx) * patches[..., 1:, :d]
        x11 = (dy) * (dx) * patches[..., 1:, 1:]
        return x00 + x01 + x10 + x11
    return patches


def corr(fmap1, fmap2, coords, ii, jj, radius=1, dropout=1):
    coords = coords.float()
    return CorrLayer.apply(fmap1, fmap2, coords, ii, jj, radius, dropout)


def corr_pyramid(fmap1, pyramid, coords, ii, jj, radius=3, scales=(1.0, 0.25), row_stride=None, out=None):
    """Inference-only fused lookup over the 2-level pyramid: [B, M, 49*P*P*2] (a view of rows
    `row_stride` elements apart when given — see cuda_corr.pyramid_forward)."""
    return cuda_corr.pyramid_forward(fmap1, list(pyramid), list(scales), coords.float(), ii, jj, radius,
                                     row_stride=row_stride, out=out)
