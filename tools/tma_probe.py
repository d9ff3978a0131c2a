"""TMA ingest rate per SM vs ring depth (development tool).   python tools/tma_probe.py"""
import ctypes, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("deep-visual-odometry-slam_b200")
lib = pkg._lib
# the probe is a development micro-benchmark: it is NOT part of libdpvo_b200.so; build it on demand next to this file
import subprocess
_CS = os.path.join(ROOT, "deep-visual-odometry-slam_b200", "csrc")
_SO = os.path.join(ROOT, "tools", "libtma_probe.so")
if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(os.path.join(ROOT, "tools", "tma_probe.cu")):
    subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
                           "-Xcompiler", "-fPIC", "-shared", "-I", _CS, "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tools", "tma_probe.cu"), os.path.join(_CS, "api.cu"), "-o", _SO, "-lcudart"])
_probe = ctypes.CDLL(_SO)
import sys as _s
rows, K = (45312 * 6, 64) if "--chunked" in _s.argv else (45312, 384)
A = torch.randn(rows, K, device="cuda").half()
big = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for grid in (148, 16):
    for stages in (2, 4, 8, 12):
        cyc = torch.zeros(grid, dtype=torch.int64, device="cuda")
        big.zero_()                                   # flush A out of L2 (126 MB)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        lib.check(_probe.dpvo_debug_tma_probe(lib.ptr(A), lib.c_i64(rows), lib.c_i32(K), lib.c_i32(stages),
                                               lib.c_i32(4), lib.c_i32(grid), lib.ptr(cyc), lib.stream_of(A)), "probe")
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        tiles = rows // 128
        per_cta = -(-tiles // grid) * 4 * (K // 64) * 16384     # bytes pulled by the busiest CTA
        c = cyc.max().item()
        print(f"grid {grid:3d} stages {stages:2d}: {per_cta / c:6.1f} B/clk/SM  ({rows * K * 2 * 4 / ms / 1e6:7.1f} GB/s total, {ms*1e3:.0f} us)")
