"""How busy the GPU is while this repo's DPVO class tracks the synthetic sequence (development tool):
frames [start, start + count) under torch.profiler — busy union of all kernels / copies against the wall span, the
number of launches per frame and the largest kernels.   python tools/seq_gpu_busy.py [--start 500] [--count 60]"""
import argparse, importlib, json, os, sys, time
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def host_profile(slam, hf, intr, a, frames):
    """where the HOST time of a frame goes (the tracker is host-bound): cProfile over the same frames, unsynchronised"""
    import cProfile, io, pstats
    pr = cProfile.Profile()
    t0 = time.perf_counter()
    pr.enable()
    for t in range(a.start, frames):
        slam(t, hf[t].cuda(non_blocking=True), intr)
    pr.disable()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    out = {}
    for key in ("tottime", "cumulative"):
        buf = io.StringIO()
        pstats.Stats(pr, stream=buf).strip_dirs().sort_stats(key).print_stats(45)
        out[key] = buf.getvalue()
    tag = "nolc" if a.no_loop_closure else "lc"
    with open(os.path.join(ROOT, "gpurun_out", f"seq_host_profile_{tag}.txt"), "w") as f:
        f.write(f"frames {a.count} from {a.start}; wall {1e3 * wall / a.count:.3f} ms per frame under cProfile\n")
        for key, v in out.items():
            f.write(f"==== sorted by {key}\n{v}\n")
    print(f"host profile: {1e3 * wall / a.count:.3f} ms per frame under cProfile")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--start", type=int, default=500)
    ap.add_argument("--count", type=int, default=60)
    ap.add_argument("--no-loop-closure", action="store_true")
    ap.add_argument("--top", type=int, default=14)
    ap.add_argument("--host", action="store_true", help="cProfile of the host side instead of the CUPTI trace")
    a = ap.parse_args()
    import refdpvo
    from torch.profiler import profile, ProfilerActivity
    refdpvo.select_backend("dropin")
    frames = a.start + a.count
    seq = refdpvo.SynthSequence(frames, seed=1234)
    over = {"LOOP_CLOSURE": False} if a.no_loop_closure else {}
    slam = refdpvo.build_slam(seed=1234, cfg=refdpvo.make_cfg(**over), seq=seq, pipeline="b200")
    hf = seq.host_frames(pin=True)
    intr = torch.as_tensor(seq.intrinsics).cuda()
    torch.set_grad_enabled(False)
    for t in range(a.start):
        slam(t, hf[t].cuda(non_blocking=True), intr)
    torch.cuda.synchronize()
    if a.host:
        return host_profile(slam, hf, intr, a, frames)
    t0 = time.perf_counter()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for t in range(a.start, frames):
            slam(t, hf[t].cuda(non_blocking=True), intr)
        torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start]
    ivs = sorted((e.time_range.start, e.time_range.end) for e in evs)
    busy, cs, ce = 0.0, None, None
    for s, e in ivs:
        if ce is None or s > ce:
            if ce is not None:
                busy += ce - cs
            cs, ce = s, e
        else:
            ce = max(ce, e)
    if ce is not None:
        busy += ce - cs
    span = ivs[-1][1] - ivs[0][0]
    agg = {}
    for e in evs:
        name = e.name.replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        name = name.split("(")[0].split("<")[0].replace("void ", "").replace("dpvo::", "").replace("at::native::", "")[-44:]
        c = agg.setdefault(name, [0, 0.0]); c[0] += 1; c[1] += e.time_range.end - e.time_range.start
    top = sorted(agg.items(), key=lambda kv: -kv[1][1])[:a.top]
    row = {"frames": a.count, "start": a.start, "loop_closure": not a.no_loop_closure,
           "wall_ms_per_frame_under_profiler": 1e3 * wall / a.count, "gpu_span_ms_per_frame": span / 1e3 / a.count,
           "gpu_busy_ms_per_frame": busy / 1e3 / a.count, "gpu_busy_fraction": busy / span,
           "launches_per_frame": len(evs) / a.count,
           "top_kernels_ms_per_frame": {k: [c, round(t / 1e3 / a.count, 4)] for k, (c, t) in top}}
    print(json.dumps(row), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    tag = "nolc" if a.no_loop_closure else "lc"
    with open(os.path.join(ROOT, "gpurun_out", f"seq_gpu_busy_{tag}.json"), "w") as f:
        json.dump(row, f, indent=1)


if __name__ == "__main__":
    main()
