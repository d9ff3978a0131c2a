"""A few replays of hotpath.FullFrame's frame as STREAM launches (ncu lists kernels, not graph nodes)."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
hotpath = importlib.import_module("deep-visual-odometry-slam_b200.hotpath")
ff = hotpath.FullFrame(torch.device("cuda", 0), n_frames=40, seed=0, inactive_capacity=400_000)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    ff.step_frame()
torch.cuda.synchronize()
print("ok", ff.steady_edges)
