"""Golden vectors for the per-frame encoders (N3) from the UNMODIFIED reference module on CPU.

    python tests/golden/make_golden_encoder.py     (needs /root/reference)

Instantiates dpvo/extractor.py::BasicEncoder4 as Patchifier does (net.py:95-96: fnet = instance norm / 128 outputs,
inet = no norm / 384 outputs), loads the seeded parameters of synth.encoder_state_dict and runs forward in float32 on
two small seeded images.  Parameters and images are NOT stored: tests regenerate them from the same seeds.
Output (committed): tests/golden/ref_cpu_encoder.npz (float16 storage of the float32 outputs would lose the reference's
precision, so the maps are kept float32 but small: 2 x C x 14 x 18).
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

CASES = dict(fnet=dict(out_dim=128, norm_fn="instance", seed=21), inet=dict(out_dim=384, norm_fn="none", seed=22))
N_IMG, H, W, IMG_SEED = 2, 56, 72, 5


def main():
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
    sys.path.insert(0, "/root/reference/methods/dpvo")
    from dpvo import extractor
    synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
    torch.set_num_threads(4)
    images = torch.from_numpy(synth.encoder_images(N_IMG, H, W, IMG_SEED))
    out = {}
    for name, c in CASES.items():
        m = extractor.BasicEncoder4(output_dim=c["out_dim"], norm_fn=c["norm_fn"]).eval()
        m.load_state_dict({k: torch.from_numpy(v) for k, v in synth.encoder_state_dict(c["out_dim"], c["seed"]).items()}, strict=True)
        with torch.no_grad():
            out[name] = m(images)[0].numpy()
    p = os.path.join(HERE, "ref_cpu_encoder.npz")
    np.savez_compressed(p, **out)
    print(p, os.path.getsize(p), "bytes", {k: (v.shape, float(np.abs(v).mean())) for k, v in out.items()})


if __name__ == "__main__":
    main()
