"""Golden vectors for the Update operator (N1) from the UNMODIFIED reference module on CPU.

    python tests/golden/make_golden_update.py     (needs /root/reference and oracle/_ref/*.so)

Instantiates dpvo/net.py::Update(3), loads the seeded parameters of synth.update_state_dict(7) and
runs Update.forward in float32 on a small patch graph.  The only substitution: fastba.neighbors is a
CUDA-only extension call (ba.cpp:59-97), so the module attribute is pointed at the numpy restatement
oracle/fastba_oracle.neighbors (itself pinned against the reference extension on the GPU).
The parameters are NOT stored (12 MB): tests regenerate them from the same seeded numpy generator.
Output (committed): tests/golden/ref_cpu_update.npz
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
sys.path.insert(0, "/root/reference/methods/dpvo")

from dpvo import net as refnet                                   # noqa: E402
from oracle import fastba_oracle                                 # noqa: E402

synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")

SEED_W, SEED_X = 7, 3


def _neighbors(ii, jj):
    ix, jx = fastba_oracle.neighbors(ii.numpy(), jj.numpy())
    return torch.from_numpy(ix), torch.from_numpy(jx)


def main():
    torch.set_num_threads(4)
    refnet.fastba.neighbors = _neighbors
    upd = refnet.Update(3).eval()
    sd = {k: torch.from_numpy(v) for k, v in synth.update_state_dict(SEED_W).items()}
    upd.load_state_dict(sd, strict=True)
    d = synth.update_inputs(n_frames=10, M=6, seed=SEED_X)
    ii, jj, kk = (torch.from_numpy(d[k]) for k in ("ii", "jj", "kk"))
    net = torch.from_numpy(d["net"])[None]
    corr = torch.from_numpy(d["corr"])[None]
    imap = torch.from_numpy(d["imap"])
    inp = imap[kk % imap.shape[0]][None]                          # ctx gather, dpvo.py:333
    with torch.no_grad():
        net_out, (delta, weight, _) = upd(net, inp, corr, None, ii, jj, kk)
    out = dict(seed_w=SEED_W, seed_x=SEED_X, n_frames=10, M=6,
               net=net_out[0].numpy(), delta=delta[0].numpy(), weight=weight[0].numpy())
    p = os.path.join(HERE, "ref_cpu_update.npz")
    np.savez_compressed(p, **out)
    print(p, os.path.getsize(p), "bytes", {k: np.shape(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
