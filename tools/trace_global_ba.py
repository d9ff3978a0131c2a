"""Where a global bundle adjustment's time goes (development / measurement tool).

    python tools/trace_global_ba.py --sequence 1000          # the 1,000-frame synthetic sequence's LAST global BA
    python tools/trace_global_ba.py --synthetic 1000         # tools/time_global_ba.py's problem (random loop pairs)

For either problem: (1) ms per `fastba.BA(..., iterations=2, eff_impl=True)` call with CUDA events, (2) one call under
torch.profiler (CUPTI kernel records, real concurrency between the caller's stream and the solver's internal
critical-path stream): per-kernel count / total / mean, busy union and span, (3) the frame-pair occupancy of the
reduced camera system S (block (a, b) is non-zero when one patch is seen from both frames).
"""
import argparse, importlib, json, os, sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def kernel_profile(fn):
    from torch.profiler import profile, ProfilerActivity
    fn(); torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        fn()
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start]
    agg, ivs = {}, []
    for e in evs:
        name = e.name.split("(")[0].split("<")[0].replace("dpvo::", "")[-48:] or e.name[:48]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += e.time_range.end - e.time_range.start
        ivs.append((e.time_range.start, e.time_range.end))
    ivs.sort()
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
    span = (ivs[-1][1] - ivs[0][0]) if ivs else 0.0
    rows = {k: {"n": c, "total_us": round(t, 1), "mean_us": round(t / c, 2)}
            for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])}
    return {"span_us": round(span, 1), "busy_union_us": round(busy, 1), "kernels": rows}


def occupancy(ii, jj, kk, n):
    npatch = int(kk.max().item()) + 1
    inc = torch.zeros(npatch, n, device=ii.device)
    inc[kk, jj] = 1
    inc[kk, ii] = 1
    return ((inc.t() @ inc) > 0).cpu().numpy()


def time_call(fn, reps=5):
    ts = []
    for _ in range(reps + 1):
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts[1:])), [round(t, 3) for t in ts[1:]]


def sequence_problem(frames):
    """run this repo's DPVO class over the synthetic sequence and keep the arguments of its last global BA"""
    import refdpvo
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    refdpvo.select_backend("dropin")
    seq = refdpvo.SynthSequence(frames, seed=1234)
    slam = refdpvo.build_slam(seed=1234, cfg=refdpvo.make_cfg(), seq=seq, pipeline="b200")
    hf = seq.host_frames(pin=True)
    intr = torch.as_tensor(seq.intrinsics).cuda()
    captured = {}
    dmod = importlib.import_module("deep-visual-odometry-slam_b200.dpvo")
    orig = dmod.fastba.BA

    def spy(*args, **kw):
        if kw.get("eff_impl"):
            captured["args"] = [x.clone() if torch.is_tensor(x) else x for x in args]
            captured["kw"] = dict(kw)
        return orig(*args, **kw)
    dmod.fastba.BA = spy
    for t in range(frames):
        slam(t, hf[t].cuda(non_blocking=True), intr)
    torch.cuda.synchronize()
    dmod.fastba.BA = orig
    return captured["args"], captured["kw"], orig


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sequence", type=int, default=0)
    ap.add_argument("--synthetic", type=int, default=0)
    ap.add_argument("--loops", type=int, default=200)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    pkg = importlib.import_module("deep-visual-odometry-slam_b200")
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    torch.set_grad_enabled(False)
    if a.sequence:
        args, kw, BA = sequence_problem(a.sequence)
        poses, patches = args[0], args[1]
        ii, jj, kk, t0, t1 = args[6], args[7], args[8], args[9], args[10]
        what = f"sequence_{a.sequence}"

        def call():
            BA(poses.clone(), patches.clone(), *args[2:], **kw)
    else:
        synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
        nf, M = a.synthetic, 96
        rng = np.random.default_rng(0)
        ii, jj, kk = synth.dpvo_edges(nf, M, removal=10 ** 9)
        for _ in range(a.loops):
            i, j = int(rng.integers(0, nf)), int(rng.integers(0, nf))
            if abs(i - j) >= 30:
                ii = np.concatenate([ii, np.full(M, i)]); jj = np.concatenate([jj, np.full(M, j)])
                kk = np.concatenate([kk, i * M + np.arange(M)])
        d = synth.ba_inputs(nf, seed=1, M=M, n_buf=nf + 2, edges=(ii.astype(np.int64), jj.astype(np.int64), kk.astype(np.int64)))
        dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
        P0 = dev(d["poses"])[None]; Z0 = dev(d["patches"])[None]
        rest = (dev(d["intrinsics"])[None], dev(d["target"])[None], dev(d["weight"])[None],
                torch.tensor([1e-4], device="cuda"), dev(d["ii"]), dev(d["jj"]), dev(d["kk"]))
        ii, jj, kk = rest[4], rest[5], rest[6]
        t0, t1 = 1, nf
        what = f"synthetic_{nf}"

        def call():
            pkg.fastba.BA(P0.clone(), Z0.clone(), *rest, t0, t1, M, 2, True)
    occ = occupancy(ii.long(), jj.long(), kk.long(), int(t1))
    # the blocked Cholesky's bulk update in both arithmetics (include/dpvo_b200.h: dpvo_chol_bulk_mode)
    for mode, name in ((1, "fp32"), (0, "split_tf32")):
        pkg.cuda_ba.chol_bulk_mode(mode)
        ms, reps = time_call(call)
        prof = kernel_profile(call)
        tag = what + "_" + name + (("_" + a.tag) if a.tag else "")
        row = {"what": what, "tag": a.tag, "chol_bulk": name, "t0": int(t0), "t1": int(t1), "free_poses": int(t1 - t0),
               "system": 6 * int(t1 - t0), "edges": int(kk.numel()), "ms_per_call": ms, "reps_ms": reps,
               "occupied_frame_pairs": float(occ[t0:, t0:].mean()), **prof}
        with open(os.path.join(ROOT, "gpurun_out", f"gba_trace_{tag}.json"), "w") as f:
            json.dump(row, f, indent=1)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
