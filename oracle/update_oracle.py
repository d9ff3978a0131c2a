"""CPU restatement of DPVO's Update operator — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows, in plain torch functional ops on float32/float64 tensors (no nn.Module, no CUDA extension):
  Update.forward            methods/dpvo/dpvo/net.py:69-88
  GatedResidual.forward     methods/dpvo/dpvo/blocks.py:30-31
  SoftAgg.forward           methods/dpvo/dpvo/blocks.py:42-50
  scatter_softmax / _sum    methods/dpvo/dpvo/torch_scatter.py:75-132, 17-72
  fastba.neighbors          ba.cpp:59-97 (through oracle/fastba_oracle.neighbors)
Pinned by tests/golden/ref_cpu_update.npz, produced by tests/golden/make_golden_update.py from the
reference's own Update module on CPU (tests/test_oracle_cpu.py::test_update_oracle_vs_reference).
`forward(..., autocast=True)` on a CUDA device reruns the same graph the way dpvo.py:332-336 does
(torch.autocast fp16): that is the numerical path the fused kernels are compared with.
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import fastba_oracle

LN_EPS = 1e-3  # every LayerNorm of the operator (net.py:37,42,44,53)


def _lin(sd, name, x):
    return F.linear(x, sd[name + ".weight"], sd[name + ".bias"])


def _ln(sd, name, x):
    return F.layer_norm(x, (x.shape[-1],), sd[name + ".weight"], sd[name + ".bias"], LN_EPS)


def _mlp2(sd, name, x):  # Linear, ReLU, Linear (net.py:27-35)
    return _lin(sd, name + ".2", torch.relu(_lin(sd, name + ".0", x)))


def _gated_residual(sd, name, x):  # blocks.py:30-31
    gate = torch.sigmoid(_lin(sd, name + ".gate.0", x))
    return x + gate * _mlp2(sd, name + ".res", x)


def _soft_agg(sd, name, x, keys):  # blocks.py:42-50 with torch_scatter.py's softmax / sum
    _, jx = torch.unique(keys, return_inverse=True)
    G = int(jx.max().item()) + 1
    g = _lin(sd, name + ".g", x)
    f = _lin(sd, name + ".f", x)
    idx = jx.view(-1, 1).expand_as(g)
    gmax = torch.full((G, g.shape[1]), float("-inf"), dtype=g.dtype, device=g.device)
    gmax = gmax.scatter_reduce(0, idx, g, reduce="amax", include_self=False)
    ex = torch.exp(g - gmax[jx])
    den = torch.zeros((G, g.shape[1]), dtype=ex.dtype, device=g.device).scatter_add_(0, idx, ex)
    w = ex / (den[jx] + 1e-12)
    y = torch.zeros((G, g.shape[1]), dtype=f.dtype, device=g.device).scatter_add_(0, idx, (f * w).to(f.dtype))
    return _lin(sd, name + ".h", y)[jx]


def neighbors(kk, jj):
    ix, jx = fastba_oracle.neighbors(kk.cpu().numpy(), jj.cpu().numpy())
    return torch.from_numpy(ix).to(kk.device), torch.from_numpy(jx).to(kk.device)


def forward(sd, net, inp, corr, ii, jj, kk, autocast=False, neigh=None):
    """sd: dict name -> tensor (reference state_dict keys).  net, inp: [E, 384]; corr: [E, 882];
    ii, jj, kk: int64 [E]; neigh: precomputed neighbors(kk, jj) (timing runs keep the numpy
    restatement out of the timed region).  Returns (net [E,384], delta [E,2], weight [E,2])."""
    ix, jx = neigh if neigh is not None else neighbors(kk, jj)
    with torch.autocast(net.device.type, dtype=torch.float16, enabled=autocast):
        c = _lin(sd, "corr.0", corr).relu()
        c = _ln(sd, "corr.3", _lin(sd, "corr.2", c)).relu()
        net = net + inp + _lin(sd, "corr.5", c)
        net = _ln(sd, "norm", net)
        mask_ix = (ix >= 0).to(net.dtype).view(-1, 1)
        mask_jx = (jx >= 0).to(net.dtype).view(-1, 1)
        net = net + _mlp2(sd, "c1", mask_ix * net[ix])
        net = net + _mlp2(sd, "c2", mask_jx * net[jx])
        net = net + _soft_agg(sd, "agg_kk", net, kk)
        net = net + _soft_agg(sd, "agg_ij", net, ii * 12345 + jj)
        net = _gated_residual(sd, "gru.1", _ln(sd, "gru.0", net))
        net = _gated_residual(sd, "gru.3", _ln(sd, "gru.2", net))
        d = _lin(sd, "d.1", torch.relu(net))
        w = torch.sigmoid(_lin(sd, "w.1", torch.relu(net)))
    return net, d, w


def state_dict_tensors(sd_np, dtype=torch.float32, device="cpu"):
    return {k: torch.from_numpy(np.asarray(v)).to(device=device, dtype=dtype) for k, v in sd_np.items()}
