"""One pass of both encoders at 480x640 (for an ncu launch list)."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
enc = importlib.import_module("deep-visual-odometry-slam_b200.encoder")
synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
img = torch.from_numpy(synth.encoder_images(1, 480, 640, 0)).cuda()
for name, dim, norm in (("fnet", 128, "instance"), ("inet", 384, "none")):
    e = enc.BasicEncoder4({k: torch.from_numpy(v) for k, v in synth.encoder_state_dict(dim, 1).items()}, norm, output_scale=0.25)
    for _ in range(2):
        e(img)
torch.cuda.synchronize()
print("ok")
