"""Which stage of hotpath.FullFrame leaves unjoined work in a capture: captures stages one at a time."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
hotpath = importlib.import_module("deep-visual-odometry-slam_b200.hotpath")
ff = hotpath.FullFrame(torch.device("cuda", 0), n_frames=12, seed=1, inactive_capacity=400_000)
ff.step_frame(); torch.cuda.synchronize()


def side(fn):
    def run():
        main = torch.cuda.current_stream()
        s = torch.cuda.Stream()
        f, j = torch.cuda.Event(), torch.cuda.Event()
        f.record(main); s.wait_event(f)
        with torch.cuda.stream(s):
            fn(); j.record(s)
        main.wait_event(j)
    return run


stages = dict(normalise=ff.normalise, encode=ff.encode, step_full=ff.step_full, point_cloud=ff.point_cloud,
              keyframe=ff.keyframe, edges_append=ff._edges_append, motion_model=ff.motion_model,
              patchify=ff.patchify, side_edges=side(ff._edges_append), side_motion=side(ff.motion_model),
              side_patchify=side(ff.patchify), step_frame=ff.step_frame)
for name, fn in stages.items():
    try:
        fn(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        g.replay(); torch.cuda.synchronize()
        print(name, "ok", flush=True)
    except Exception as ex:
        print(name, "FAILED", str(ex).split("\n")[0], flush=True)
        torch.cuda.synchronize()
