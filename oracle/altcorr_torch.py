"""Multi-threaded CPU restatement of the correlation lookup in torch ops.  TEST INFRASTRUCTURE ONLY
(bench.py's cpu_baseline / --impl reference legs and tests/; never on the product path).

Same algorithm as oracle/altcorr_oracle.py (which restates altcorr/correlation_kernel.cu:83-137,
222-233 and dpvo.py:200-207), written with torch CPU tensor ops so that ATen's intra-op thread pool
uses every host core: this is the "torch-level correlation" the reference would need on a CPU (it
ships none: SURVEY.md §0.4).  fp32 accumulation, results rounded to the feature dtype where the
reference rounds.  Checked against the numpy restatement in tests/test_oracle_cpu.py.
"""
import torch


def prepare(gmap, pyramid):
    """Channels-last float32 copies of the feature rings (done once, outside the timed region —
    the GPU path's rings are resident too)."""
    g = torch.as_tensor(gmap)[0].float()                                   # [N1,C,P,P]
    fms = [torch.as_tensor(fm)[0].float().permute(0, 2, 3, 1).contiguous() for fm in pyramid]  # [N2,H,W,C]
    return g, fms


def _floor_int(v):
    f = torch.floor(v)
    f = torch.where(torch.isnan(f), torch.zeros_like(f), f)
    return f.clamp(-(1 << 24), 1 << 24).long()


def corr_pyramid(prep, scales, coords, ii, jj, radius, dtype=torch.float16, chunk=1024):
    """[1,M,(2r+1)^2*P*P*L] in `dtype`, feature order [x-off][y-off][i0][j0][level]."""
    g, fms = prep
    coords = torch.as_tensor(coords)[0].float()                            # [M,2,P,P]
    ii = torch.as_tensor(ii).long(); jj = torch.as_tensor(jj).long()
    M, _, P, _ = coords.shape
    R, D = radius, 2 * radius + 2
    off = torch.arange(D) - R
    outs = []
    for fm, s in zip(fms, scales):
        N2, H, W, C = fm.shape
        lvl = torch.empty((M, D - 1, D - 1, P, P), dtype=dtype)
        for m0 in range(0, M, chunk):
            m1 = min(M, m0 + chunk)
            c = coords[m0:m1] * float(s)
            x, y = c[:, 0], c[:, 1]                                        # [m,P,P]
            fx, fy = _floor_int(x), _floor_int(y)
            i1 = (fy[:, None, None] + off[None, :, None, None, None]).expand(-1, D, D, P, P)
            j1 = (fx[:, None, None] + off[None, None, :, None, None]).expand(-1, D, D, P, P)
            ok = (i1 >= 0) & (i1 < H) & (j1 >= 0) & (j1 < W)
            v = jj[m0:m1][:, None, None, None, None].expand_as(i1)
            win = fm[v, i1.clamp(0, H - 1), j1.clamp(0, W - 1)]            # [m,D,D,P,P,C]
            p = g[ii[m0:m1]].permute(0, 2, 3, 1)                           # [m,P,P,C]
            raw = (win * p[:, None, None]).sum(-1)
            raw = torch.where(ok, raw, torch.zeros_like(raw)).to(dtype)    # rounded like the kernel's store
            rd = lambda t: t.to(dtype)
            dx = rd(x - torch.floor(x)); dy = rd(y - torch.floor(y))
            one = torch.ones((), dtype=dtype)
            mdx, mdy = rd(one - dx), rd(one - dy)
            e = lambda w: w[:, None, None]
            o = rd(e(rd(mdx * mdy)) * raw[:, 0:D - 1, 0:D - 1])
            o = rd(o + rd(e(rd(dx * mdy)) * raw[:, 0:D - 1, 1:D]))
            o = rd(o + rd(e(rd(mdx * dy)) * raw[:, 1:D, 0:D - 1]))
            o = rd(o + rd(e(rd(dx * dy)) * raw[:, 1:D, 1:D]))
            lvl[m0:m1] = o.permute(0, 2, 1, 3, 4)                           # [m, x-off, y-off, P, P]
        outs.append(lvl)
    st = torch.stack(outs, -1)
    return st.reshape(1, M, -1)
