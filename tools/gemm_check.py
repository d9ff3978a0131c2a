"""tcgen05 edge-MLP layer against torch on the GPU (development tool).   python tools/gemm_check.py"""
import ctypes, importlib, os, sys, time
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("deep-visual-odometry-slam_b200")
upd = importlib.import_module("deep-visual-odometry-slam_b200.update_ops")


def ref(A, W, b, gate=None, res1=None, res2=None, idx=None, ln=None, act=0):
    v = A.float() @ W.float().t() + b
    if gate is not None: v = v * gate.float()
    if res1 is not None: v = v + res1.float()
    if res2 is not None: v = v + (res2[idx] if idx is not None else res2).float()
    if ln is not None: v = torch.nn.functional.layer_norm(v, (384,), ln[0], ln[1], ln[2])
    if act == 1: v = v.relu()
    if act == 2: v = v.sigmoid()
    return v


def main():
    torch.manual_seed(0)
    dev = "cuda"
    bad = 0
    prof = "--profile" in sys.argv
    for (M, K, ld) in [] if prof else [(128, 64, 64), (128, 384, 384), (300, 384, 384), (45312, 384, 384), (1000, 882, 896), (45312, 882, 896)]:
        Abuf = torch.full((M, ld), float("nan"), dtype=torch.float16, device=dev)
        Abuf[:, :K] = torch.randn(M, K, device=dev).half()
        A = Abuf[:, :K]
        Wp = torch.zeros(384, ld, dtype=torch.float16, device=dev)
        Wp[:, :K] = (torch.randn(384, K, device=dev) / K ** 0.5).half()
        W = Wp[:, :K]
        b = torch.randn(384, device=dev)
        gate = torch.rand(M, 384, device=dev).half()
        res1 = torch.randn(M, 384, device=dev).half()
        table = torch.randn(77, 384, device=dev).half()
        idx = torch.randint(0, 77, (M,), device=dev)
        gam, bet = torch.randn(384, device=dev), torch.randn(384, device=dev)
        cases = [dict(), dict(act=1), dict(act=2), dict(res1=res1), dict(gate=gate, res1=res1),
                 dict(res1=res1, res2=table, idx=idx, ln=(gam, bet, 1e-3)), dict(ln=(gam, bet, 1e-3), act=1)]
        for kw in cases:
            want = ref(A, W, b, **kw)
            got = upd.linear384(A, W, b, **kw).float()
            torch.cuda.synchronize()
            err = (got - want).abs().max().item()
            scale = want.abs().max().item()
            flag = "" if err <= 2e-3 * scale + 1e-3 else "  <-- BAD"
            bad += bool(flag)
            print(f"M={M} K={K} {sorted(kw)}: max err {err:.3e} (scale {scale:.2f}){flag}", flush=True)
    # timing
    M = 45312
    for K, ld in ((384, 384), (882, 896)):
        A = torch.randn(M, ld, device=dev).half()[:, :K]
        W = torch.randn(384, ld, device=dev).half()[:, :K]
        b = torch.randn(384, device=dev)
        for _ in range(3): upd.linear384(A, W, b, act=1)
        torch.cuda.synchronize()
        if prof:
            continue
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): upd.linear384(A, W, b, act=1)
        e1.record(); torch.cuda.synchronize()
        ms_stream = e0.elapsed_time(e1) / 20
        g = torch.cuda.CUDAGraph()                      # host launch cost (tensor-map encode + ctypes) out of the picture
        with torch.cuda.graph(g):
            for _ in range(20): upd.linear384(A, W, b, act=1)
        g.replay(); torch.cuda.synchronize()
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(f"  stream-launched {ms_stream*1e3:.1f} us/launch, graph {ms*1e3:.1f} us/launch")
        if os.environ.get("DPVO_LINEAR_PROFILE") and K == 384:
            buf = (ctypes.c_longlong * 16)()
            pkg._lib.lib.dpvo_linear_profile_read(buf)
            names = ["prod_wait_empty", "prod_total", "mma_wait_A", "mma_wait_acc", "mma_wait_W", "mma_total",
                     "epi_wait_acc", "epi_total", "tiles"]
            print("profile (cycles, CTA 0):", {n: int(buf[i]) for i, n in enumerate(names)})
        Ac = A.contiguous(); Wc = W.contiguous()
        for _ in range(3): torch.relu(torch.nn.functional.linear(Ac, Wc, b.half()))
        e0.record()
        for _ in range(20): torch.relu(torch.nn.functional.linear(Ac, Wc, b.half()))
        e1.record(); torch.cuda.synchronize()
        ms_t = e0.elapsed_time(e1) / 20
        print(f"timing M={M} K={K}: ours {ms*1e3:.1f} us ({2*M*K*384/ms/1e9:.0f} TFLOP/s), torch linear+relu {ms_t*1e3:.1f} us")
    print("BAD" if bad else "ALL OK")
    return bad


if __name__ == "__main__":
    sys.exit(main())
