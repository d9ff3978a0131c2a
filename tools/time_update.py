"""Update operator at the bench size (E = 45,312 edges): fused kernels vs the torch graph the
reference runs (oracle restatement under autocast fp16, on the same GPU), with errors against the
float32 oracle (development tool).   python tools/time_update.py [--profile]"""
import importlib, json, os, sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("deep-visual-odometry-slam_b200")
synth = importlib.import_module("deep-visual-odometry-slam_b200.synth")
upd = importlib.import_module("deep-visual-odometry-slam_b200.update")
from oracle import update_oracle as uo


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def main():
    dev = "cuda"
    sd_np = synth.update_state_dict(11)
    d = synth.update_inputs(n_frames=40, M=96, seed=1)
    E = len(d["ii"])
    ii, jj, kk = (torch.from_numpy(d[k]).to(dev) for k in ("ii", "jj", "kk"))
    net = torch.from_numpy(d["net"]).to(dev).half()
    imap = torch.from_numpy(d["imap"]).to(dev).half()
    tidx = kk % imap.shape[0]
    corr_buf = torch.zeros(E, 896, dtype=torch.float16, device=dev)
    corr_buf[:, :882] = torch.from_numpy(d["corr"]).to(dev).half()
    corr = corr_buf[:, :882]
    op = upd.Update.from_state_dict({k: torch.from_numpy(v) for k, v in sd_np.items()}, device=dev)
    nf, npatch = 40, 40 * 96
    run = lambda: op(net[None], (imap, tidx), corr[None], None, ii, jj, kk, n_frames=nf, n_patches=npatch)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    for _ in range(3): out = run()
    torch.cuda.synchronize()
    K = 20
    e0, e1 = ev(), ev(); e0.record()
    for _ in range(K): out = run()
    e1.record(); torch.cuda.synchronize()
    ms_ours = e0.elapsed_time(e1) / K
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        out_g = run()
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K): g.replay()
    e1.record(); torch.cuda.synchronize()
    ms_graph = e0.elapsed_time(e1) / K
    if "--profile" in sys.argv:
        return
    sd = uo.state_dict_tensors(sd_np, device=dev)
    neigh = uo.neighbors(kk, jj)
    inp32 = imap.float()[tidx]
    ref = lambda ac: uo.forward(sd, net.float(), inp32, corr.float().contiguous(), ii, jj, kk, autocast=ac, neigh=neigh)
    with torch.no_grad():
        for _ in range(2): a = ref(True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5): a = ref(True)
        e1.record(); torch.cuda.synchronize()
        ms_auto = e0.elapsed_time(e1) / 5
        w = ref(False)
    o_net, (o_d, o_w, _) = out
    res = dict(edges=E, ms_fused=round(ms_ours, 4), ms_fused_cuda_graph=round(ms_graph, 4),
               ms_torch_autocast_graph=round(ms_auto, 4),
               err_fused=dict(net=rel(o_net[0].float(), w[0]), delta=rel(o_d[0], w[1]), weight=rel(o_w[0], w[2])),
               err_autocast=dict(net=rel(a[0].float(), w[0]), delta=rel(a[1].float(), w[1]), weight=rel(a[2].float(), w[2])))
    print(json.dumps(res))


if __name__ == "__main__":
    main()
