"""Where the host-buffer (e2e) frame goes: PCIe copy alone, host issue time alone, pipelined step
(development tool).   python tools/time_e2e.py"""
import importlib, json, os, sys, time
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    hotpath = importlib.import_module("deep-visual-odometry-slam_b200.hotpath")
    dev = torch.device("cuda", 0)
    hp = hotpath.UpdateHotPath(dev, n_frames=40, seed=0)
    hin, hout = hp.make_host_inputs(), hp.make_host_outputs()
    ev = lambda: torch.cuda.Event(enable_timing=True)
    K = 50
    # (a) the H2D copy alone
    for _ in range(3):
        hp._stage[0].copy_(hin, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = ev(), ev(); e0.record()
    for i in range(K):
        hp._stage[i & 1].copy_(hin, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    copy_ms = e0.elapsed_time(e1) / K
    # (b) host time to issue one device-resident step (no sync inside) and its GPU time
    for _ in range(5):
        hp.step()
    torch.cuda.synchronize()
    w0 = time.perf_counter(); e0.record()
    for _ in range(K):
        hp.step()
    issue_ms = (time.perf_counter() - w0) * 1e3 / K
    e1.record(); torch.cuda.synchronize()
    gpu_ms = e0.elapsed_time(e1) / K
    # (c) pipelined host-buffer step
    for _ in range(5):
        hp.step_host(hin, hout)
    torch.cuda.synchronize()
    w0 = time.perf_counter(); e0.record()
    for _ in range(K):
        hp.step_host(hin, hout)
    issue2_ms = (time.perf_counter() - w0) * 1e3 / K
    e1.record(); torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1) / K
    print(json.dumps({"h2d_bytes": hin.numel(), "h2d_copy_ms": copy_ms, "h2d_GBs": hin.numel() / copy_ms / 1e6,
                      "host_issue_ms_per_step": issue_ms, "gpu_ms_per_step": gpu_ms,
                      "host_issue_ms_per_e2e_step": issue2_ms, "e2e_ms_per_step": e2e_ms}))


if __name__ == "__main__":
    main()
