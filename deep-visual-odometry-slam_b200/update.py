"""DPVO's Update operator (net.py:23-88) on the B200 kernels — inference forward.

Same call as the reference module:

    net, (delta, weight, _) = update(net, inp, corr, flow, ii, jj, kk)

with net / inp [1, E, 384], corr [1, E, 882], ii / jj / kk int64 [E]; built from a reference
state_dict (`Update.from_state_dict`) or module (`Update.from_reference(network.update)`), so a
maintainer swaps it in with one line after the checkpoint is loaded.

What runs (19 tensor-core layer launches + row kernels, no host synchronisation; the index-only work —
neighbours, group labels, bucket sorts — is forked onto a side stream under the first three layers):
  * every nn.Linear(·, 384) over the E edge rows is `update_ops.linear384` (tcgen05 + TMEM + TMA,
    csrc/update_gemm.cu) with the row-local work that follows it in the reference fused into its
    epilogue: bias, ReLU / sigmoid, LayerNorm, residual adds, the GatedResidual product and the
    `imap[kk % n]` context gather;
  * the neighbour gathers, SoftAgg's grouped softmax / sum and its expand, and the two 384->2 heads
    are the row kernels of csrc/update_agg.cu; group labels come from `cuda_ba.unique_inverse`
    (sorted unique, like torch.unique in blocks.py:43) without the host sync torch.unique has.

Numerics: activations are stored as float16 between launches, every layer accumulates in float32 and
LayerNorm / softmax statistics are float32.  The reference under autocast (dpvo.py:332) keeps the
residual stream in float32 between its float16 GEMMs; the difference is bounded in
tests/test_update_gpu.py against the float32 oracle next to the autocast graph's own error.
CUDA only — there is no CPU fallback.
"""
import collections
import ctypes
import os

import torch

from . import _lib, cuda_ba, update_ops

DIM = update_ops.DIM
LN_EPS = 1e-3   # net.py:37,42,44,53
_L = _lib.lib
_L.dpvo_softagg_workspace_bytes.restype = _lib.c_i64

class _UpdateWeights(ctypes.Structure):
    """Mirror of dpvo_update_weights."""
    _fields_ = [("W", ctypes.c_void_p * 19), ("ldw", ctypes.c_int64 * 19), ("bias", ctypes.c_void_p * 19),
                ("ln_gamma", ctypes.c_void_p * 4), ("ln_beta", ctypes.c_void_p * 4), ("ln_eps", ctypes.c_float),
                ("corr_features", ctypes.c_int32),
                ("head_w", ctypes.c_void_p), ("head_bias_d", ctypes.c_void_p), ("head_bias_w", ctypes.c_void_p)]


_L.dpvo_update_workspace_bytes.restype = _lib.c_i64
_L.dpvo_update_workspace_bytes.argtypes = [ctypes.c_int64] * 3
_L.dpvo_update_forward.restype = ctypes.c_int
_L.dpvo_update_forward.argtypes = [ctypes.POINTER(_UpdateWeights)] + [ctypes.c_void_p] * 4 + [ctypes.c_int64] + \
    [ctypes.c_void_p] * 3 + [ctypes.c_int64] * 3 + [ctypes.c_void_p] * 4 + [ctypes.c_int64, ctypes.c_void_p]

_LINEAR = ("c1.0", "c1.2", "c2.0", "c2.2", "agg_kk.f", "agg_kk.g", "agg_kk.h", "agg_ij.f", "agg_ij.g", "agg_ij.h",
           "gru.1.gate.0", "gru.1.res.0", "gru.1.res.2", "gru.3.gate.0", "gru.3.res.0", "gru.3.res.2",
           "corr.0", "corr.2", "corr.5")
_NORMS = ("norm", "gru.0", "gru.2", "corr.3")
_HEADS = ("d.1", "w.1")


def _rows(t, name, width):
    """[1, E, width] or [E, width] -> [E, width] view."""
    if t.dim() == 3:
        if t.shape[0] != 1:
            raise RuntimeError(f"Update: {name} must have batch size 1, got {tuple(t.shape)}")
        t = t[0]
    if t.dim() != 2 or t.shape[1] != width:
        raise RuntimeError(f"Update: {name} must be [1, E, {width}], got {tuple(t.shape)}")
    return t


class Update(torch.nn.Module):
    """An nn.Module (no registered parameters: inference only) so that it can be assigned over the
    reference's `network.update` child module (dpvo.py:100-112 keeps `self.network` an nn.Module)."""

    def __init__(self, state_dict, device=None):
        super().__init__()
        dev = torch.device(device if device is not None else "cuda")
        if dev.type != "cuda":
            raise RuntimeError("Update: CUDA device required (no CPU fallback)")
        missing = [k + s for k in _LINEAR + _NORMS + _HEADS for s in (".weight", ".bias") if k + s not in state_dict]
        if missing:
            raise RuntimeError(f"Update: state_dict is missing {missing[:4]}{'...' if len(missing) > 4 else ''}")
        self.device = dev
        self.W, self.b, self.ln, self.head = {}, {}, {}, {}
        for k in _LINEAR:
            w = torch.as_tensor(state_dict[k + ".weight"]).detach().to(dev, torch.float32)
            if w.shape[0] != DIM:
                raise RuntimeError(f"Update: {k}.weight must have {DIM} output rows, got {tuple(w.shape)}")
            K = w.shape[1]
            ld = (K + 7) // 8 * 8                              # 16-byte rows for TMA
            buf = torch.zeros((DIM, ld), dtype=torch.float16, device=dev)
            buf[:, :K] = w.half()
            self.W[k] = buf[:, :K]
            self.b[k] = torch.as_tensor(state_dict[k + ".bias"]).detach().to(dev, torch.float32).contiguous()
        for k in _NORMS:
            self.ln[k] = (torch.as_tensor(state_dict[k + ".weight"]).detach().to(dev, torch.float32).contiguous(),
                          torch.as_tensor(state_dict[k + ".bias"]).detach().to(dev, torch.float32).contiguous(), LN_EPS)
        for k in _HEADS:
            self.head[k] = (torch.as_tensor(state_dict[k + ".weight"]).detach().to(dev, torch.float32).contiguous(),
                            torch.as_tensor(state_dict[k + ".bias"]).detach().to(dev, torch.float32).contiguous())
        self.head_w = torch.cat([self.head["d.1"][0], self.head["w.1"][0]], 0).contiguous()   # [4, 384]
        self.corr_features = self.W["corr.0"].shape[1]
        self._side = None
        self._cw = self._c_weights()
        self._ws = None


    def _c_weights(self):
        """dpvo_update_weights (include/dpvo_b200.h) over this instance's tensors, built once."""
        w = _UpdateWeights()
        for i, k in enumerate(_LINEAR):
            w.W[i] = self.W[k].data_ptr()
            w.ldw[i] = self.W[k].stride(0)
            w.bias[i] = self.b[k].data_ptr()
        for i, k in enumerate(_NORMS):
            w.ln_gamma[i] = self.ln[k][0].data_ptr()
            w.ln_beta[i] = self.ln[k][1].data_ptr()
        w.ln_eps = LN_EPS
        w.corr_features = self.corr_features
        w.head_w = self.head_w.data_ptr()
        w.head_bias_d = self.head["d.1"][1].data_ptr()
        w.head_bias_w = self.head["w.1"][1].data_ptr()
        return w

    # ------------------------------------------------------------------ construction helpers
    @classmethod
    def from_state_dict(cls, state_dict, device=None, prefix=""):
        if prefix:
            state_dict = {k[len(prefix):]: v for k, v in state_dict.items() if k.startswith(prefix)}
        return cls(state_dict, device)

    @classmethod
    def from_reference(cls, module, device=None):
        """`module`: the reference's dpvo.net.Update instance (any nn.Module with its state_dict keys)."""
        sd = module.state_dict()
        if device is None:
            device = next(iter(sd.values())).device
        return cls(sd, device)

    def cuda(self, device=None):
        return self

    # ------------------------------------------------------------------ pieces
    def _lin(self, name, A, **kw):
        # operands are produced / validated by forward(); the device is current and the stream fixed for the call
        return update_ops.linear384_trusted(A, self.W[name], self.b[name], self._st, **kw)

    def _gather(self, x, index):
        out = torch.empty_like(x)
        _lib.check(_L.dpvo_gather_rows_f16(_lib.ptr(x), _lib.ptr(index), _lib.c_i64(x.shape[0]), _lib.c_i32(DIM),
                                           _lib.ptr(out), _lib.stream_of(x)), "gather_rows")
        return out

    def _group(self, keys, key_bound, E):
        """Edge groups of one SoftAgg: dense labels (cuda_ba.unique_inverse — torch.unique's inverse,
        blocks.py:43, without its host sync) and the bucket sort of the edges by label."""
        _, ids, _ = cuda_ba.unique_inverse(keys, key_bound)
        G = int(min(E, key_bound))
        ws = torch.empty(int(_L.dpvo_softagg_workspace_bytes(_lib.c_i64(E), _lib.c_i64(G))), dtype=torch.uint8,
                         device=keys.device)
        _lib.check(_L.dpvo_softagg_prepare(_lib.ptr(ids), _lib.c_i64(E), _lib.c_i64(G), _lib.ptr(ws),
                                           _lib.c_i64(ws.numel()), _lib.stream_of(keys)), "softagg_prepare")
        return ids, G, ws

    def _index_work(self, ii, jj, kk, n_frames, n_patches):
        """Everything that depends on the edge list only (not on the activations): neighbours along the
        patch tracks (fastba.neighbors, net.py:74) and the two SoftAgg groupings (net.py:82-83)."""
        E = kk.numel()
        ix, jx = cuda_ba.neighbors(kk, jj)
        return dict(ix=ix, jx=jx, kk=self._group(kk, n_patches, E),
                    ij=self._group(ii * n_frames + jj, n_frames * n_frames, E))

    def _soft_agg(self, name, x, group, ln=None):
        """x + h(softagg(f(x), g(x)))[group], optionally followed by LayerNorm `ln`."""
        E = x.shape[0]
        ids, G, ws = group
        f = self._lin(name + ".f", x)
        g = self._lin(name + ".g", x)
        y = torch.zeros((G, DIM), dtype=torch.float16, device=x.device)
        _lib.check(_L.dpvo_softagg_f16(_lib.ptr(f), _lib.ptr(g), _lib.ptr(ids), _lib.c_i64(E), _lib.c_i64(G),
                                       _lib.ptr(y), _lib.c_i32(1), _lib.ptr(ws), _lib.c_i64(ws.numel()),
                                       _lib.stream_of(x)), "softagg")
        hy = self._lin(name + ".h", y)
        out = torch.empty_like(x)
        gam, bet, eps = ln if ln is not None else (None, None, 0.0)
        _lib.check(_L.dpvo_expand_add_f16(_lib.ptr(x), _lib.ptr(hy), _lib.ptr(ids), _lib.c_i64(E), _lib.ptr(gam),
                                          _lib.ptr(bet), _lib.c_f32(eps), _lib.ptr(out), _lib.stream_of(x)),
                   "expand_add")
        return out

    def _gated_residual(self, name, x, ln=None, head=None):
        gate = self._lin(name + ".gate.0", x, act=2)
        r = self._lin(name + ".res.0", x, act=1)
        return self._lin(name + ".res.2", r, gate=gate, res1=x, ln=ln, head=head)

    # ------------------------------------------------------------------ forward
    def forward(self, net, inp, corr, flow, ii, jj, kk, n_frames=None, n_patches=None):
        """inp: [1, E, 384] tensor, or a (table [n, 384], index int64 [E]) pair — the context gather
        `imap[:, kk % n]` of dpvo.py:333 then happens inside the layer that consumes it.
        n_frames / n_patches: exclusive upper bounds of jj / ii and kk (size the group tables);
        read from the index tensors (one host sync) when not given."""
        _lib.require_cuda(net, corr, ii, jj, kk)
        net = _rows(net, "net", DIM)
        corr = _rows(corr, "corr", self.corr_features)
        E = net.shape[0]
        if corr.shape[0] != E or ii.numel() != E or jj.numel() != E or kk.numel() != E:
            raise RuntimeError("Update: net, corr, ii, jj, kk disagree on the number of edges")
        if isinstance(inp, (tuple, list)):
            table, tidx = inp
            table = _rows(table, "inp table", DIM)
        else:
            table, tidx = _rows(inp, "inp", DIM), None
            if table.shape[0] != E:
                raise RuntimeError("Update: inp must have one row per edge")
        if E == 0:
            z = net.new_zeros((1, 0, 2), dtype=torch.float32)
            return net.half()[None], (z, z.clone(), None)
        if n_frames is None:
            n_frames = int(torch.maximum(ii.max(), jj.max()).item()) + 1
        if n_patches is None:
            n_patches = int(kk.max().item()) + 1
        with torch.cuda.device(net.device):
            net = net.half().contiguous()
            table = table.half().contiguous()
            if corr.dtype != torch.float16 or corr.stride(1) != 1 or corr.stride(0) % 8 or corr.data_ptr() % 16:
                buf = torch.empty((E, (self.corr_features + 7) // 8 * 8), dtype=torch.float16, device=corr.device)
                buf[:, :self.corr_features].copy_(corr)
                corr = buf[:, :self.corr_features]
            ii = ii.contiguous(); jj = jj.contiguous(); kk = kk.contiguous()
            if tidx is not None:
                if tidx.dtype != torch.int64 or tidx.numel() != E:
                    raise RuntimeError("Update: the context index must be int64 with one entry per edge")
                tidx = tidx.contiguous()
            if ii.dtype != torch.int64 or jj.dtype != torch.int64 or kk.dtype != torch.int64:
                raise RuntimeError("Update: ii, jj, kk must be int64")
            args = (net, table, tidx, corr, ii, jj, kk, int(n_frames), int(n_patches))
            ent = self._graph_lookup(args) if self.use_graphs and E >= self.graph_min_edges else None
            if ent is not None:
                ent[0].replay()
                return ent[1]
            net, delta, weight = self._forward_impl(*args)
        return net[None], (delta[None], weight[None], None)

    # ------------------------------------------------------------------ CUDA-graph cache
    # Issued eagerly the operator is host-bound: ~45 ctypes launches take 1.46 ms against 0.72 ms of GPU time at
    # E = 45,312 (profiles/r02_time_update.json).  A caller that keeps its operands in persistent buffers (this repo's
    # DPVO class: edge store, context index and correlation buffers) presents the same (pointers, E, bounds) frame
    # after frame; the second time a signature is seen the operator is captured and from then on replayed.  Outputs
    # live in the graph's buffers and are valid until the next call with the same signature.  OFF unless the caller opts
    # in (`update.use_graphs = True`): it pays only for a caller whose edge list keeps its size for many consecutive
    # calls (fixed-window replay, the 12 updates of initialisation / terminate); in DPVO's steady state keyframe
    # removal changes the edge count almost every frame and every new signature costs a capture (measured: 146 -> 93
    # frames/s on the 1,000-frame sequence), so this repo's DPVO class leaves it off.
    use_graphs = False
    graph_min_edges = 4096
    graph_cache_size = 3

    def _graph_lookup(self, args):
        if torch.cuda.is_current_stream_capturing():
            return None
        net, table, tidx, corr, ii, jj, kk, n_frames, n_patches = args
        key = (net.shape[0], net.data_ptr(), table.data_ptr(), table.shape[0], 0 if tidx is None else tidx.data_ptr(),
               corr.data_ptr(), corr.stride(0), ii.data_ptr(), jj.data_ptr(), kk.data_ptr(), n_frames, n_patches,
               net.device.index)
        if not hasattr(self, "_graphs"):
            self._graphs, self._seen = collections.OrderedDict(), {}
        ent = self._graphs.get(key)
        if ent is not None:
            self._graphs.move_to_end(key)
            return ent
        n = self._seen.get(key, 0) + 1
        if len(self._seen) > 256:
            self._seen.clear()
        self._seen[key] = n
        if n < 2:
            return None
        while len(self._graphs) >= self.graph_cache_size:
            self._graphs.popitem(last=False)
        self._forward_impl(*args)                                   # warm-up outside the capture
        torch.cuda.synchronize(net.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            o_net, o_delta, o_weight = self._forward_impl(*args)
        ent = (g, (o_net[None], (o_delta[None], o_weight[None], None)), args)   # args kept alive: the graph reads them
        self._graphs[key] = ent
        return ent

    # DPVO_UPDATE_PYTHON=1: issue the launches one ctypes call at a time (the round-1 path; A/B and debugging)
    launch_by_launch = os.environ.get("DPVO_UPDATE_PYTHON", "0") == "1"

    def _forward_impl(self, net, table, tidx, corr, ii, jj, kk, n_frames, n_patches):
        """One C call (csrc/update_forward.cu): same launches, same order, same operands as _forward_python."""
        if self.launch_by_launch:
            return self._forward_python(net, table, tidx, corr, ii, jj, kk, n_frames, n_patches)
        E = net.shape[0]
        dev = net.device
        nbytes = int(_L.dpvo_update_workspace_bytes(E, n_frames, n_patches))
        if self._ws is None or self._ws.numel() < nbytes or self._ws.device != dev \
                or torch.cuda.is_current_stream_capturing():
            ws = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8, device=dev)
            if not torch.cuda.is_current_stream_capturing():
                self._ws = ws
        else:
            ws = self._ws
        out = torch.empty((E, DIM), dtype=torch.float16, device=dev)
        delta = torch.empty((E, 2), dtype=torch.float32, device=dev)
        weight = torch.empty((E, 2), dtype=torch.float32, device=dev)
        st = torch.cuda.current_stream(dev)
        rc = _L.dpvo_update_forward(ctypes.byref(self._cw), net.data_ptr(), table.data_ptr(),
                                    0 if tidx is None else tidx.data_ptr(), corr.data_ptr(), corr.stride(0),
                                    ii.data_ptr(), jj.data_ptr(), kk.data_ptr(), E, n_frames, n_patches,
                                    out.data_ptr(), delta.data_ptr(), weight.data_ptr(), ws.data_ptr(), ws.numel(),
                                    st.cuda_stream)
        if rc:
            _lib.check(rc, "update_forward")
        ws.record_stream(st)
        return out, delta, weight

    def _forward_python(self, net, table, tidx, corr, ii, jj, kk, n_frames, n_patches):
        E = net.shape[0]
        if True:
            # the index-only work runs on a side stream underneath the correlation encoder
            main = torch.cuda.current_stream(net.device)
            self._st = _lib.c_p(main.cuda_stream)
            if self._side is None:
                self._side = torch.cuda.Stream(net.device)
                self._fork, self._join = torch.cuda.Event(), torch.cuda.Event()
            self._fork.record(main)
            with torch.cuda.stream(self._side):
                self._side.wait_event(self._fork)
                idx = self._index_work(ii, jj, kk, n_frames, n_patches)
                self._join.record(self._side)
            # correlation encoder + input injection + norm (net.py:71-72)
            c = self._lin("corr.0", corr, act=1)
            c = self._lin("corr.2", c, ln=self.ln["corr.3"], act=1)
            net = self._lin("corr.5", c, res1=net, res2=table, idx=tidx, ln=self.ln["norm"])
            main.wait_event(self._join)
            for v in (idx["ix"], idx["jx"], *idx["kk"][::2], *idx["ij"][::2]):
                v.record_stream(main)
            # temporal neighbours along each patch track (net.py:74-80)
            t = self._lin("c1.0", self._gather(net, idx["ix"]), act=1)
            net = self._lin("c1.2", t, res1=net)
            t = self._lin("c2.0", self._gather(net, idx["jx"]), act=1)
            net = self._lin("c2.2", t, res1=net)
            # soft aggregation over patches and over frame pairs (net.py:82-83); gru.0's LayerNorm rides on the second
            net = self._soft_agg("agg_kk", net, idx["kk"])
            net = self._soft_agg("agg_ij", net, idx["ij"], ln=self.ln["gru.0"])
            # gated residual blocks (net.py:41-46,85)
            net = self._gated_residual("gru.1", net, ln=self.ln["gru.2"])
            # ... the second one also accumulates the two 384->2 heads on ReLU(net) (net.py:57-66,88)
            acc = torch.zeros((E, 4), dtype=torch.float32, device=net.device)
            net = self._gated_residual("gru.3", net, head=(self.head_w, acc))
            delta = torch.empty((E, 2), dtype=torch.float32, device=net.device)
            weight = torch.empty((E, 2), dtype=torch.float32, device=net.device)
            _lib.check(_L.dpvo_update_heads_finish(_lib.ptr(acc), _lib.ptr(self.head["d.1"][1]),
                                                   _lib.ptr(self.head["w.1"][1]), _lib.c_i64(E), _lib.ptr(delta),
                                                   _lib.ptr(weight), _lib.stream_of(net)), "update_heads_finish")
        return net, delta, weight
