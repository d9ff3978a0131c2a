"""Training-side exchange on the box's NVLink: parallel.FlatGradAllReducer over a flat fp32
gradient buffer the size of DPVO's network (3,384,324 parameters + the found-inf slot = 13.54 MB), one NCCL all-reduce
per step.  Launch under torchrun; rank 0 prints one JSON line.  Device time = max over ranks.  Checks the result too:
every rank's gradient = rank + 1 -> the mean must be (world + 1) / 2, and an overflow on the last rank must set the
flag everywhere."""
import importlib, json, os, sys
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
par = importlib.import_module("deep-visual-odometry-slam_b200.parallel")

N_PARAMS = 3_384_324


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    rank, world = par.init(backend="nccl", device=dev)
    p = torch.nn.Parameter(torch.zeros(N_PARAMS, device=dev))
    red = par.FlatGradAllReducer([p])
    red.flat[:red.numel] = float(rank + 1)
    flag = red.allreduce()
    torch.cuda.synchronize()
    ok = bool(torch.allclose(p.grad, torch.full_like(p.grad, (world + 1) / 2))) and not bool(flag)
    red.flat[:red.numel] = 1.0
    if rank == world - 1:
        red.flat[7] = float("inf")
    ok = ok and bool(red.allreduce()) and bool(torch.isfinite(red.flat).all())
    for _ in range(10):
        red.allreduce(found_inf=False)
    K = 100
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    par.barrier(dev)
    e0.record()
    for _ in range(K):
        red.allreduce(found_inf=False)
    e1.record()
    torch.cuda.synchronize()
    ms = par.max_over_ranks(e0.elapsed_time(e1) / K, device=dev)
    # the collective alone, without the flag write / averaging kernels around it
    e0.record()
    for _ in range(K):
        dist.all_reduce(red.flat)
    e1.record()
    torch.cuda.synchronize()
    ms_nccl = par.max_over_ranks(e0.elapsed_time(e1) / K, device=dev)
    if rank == 0:
        nb = red.nbytes()
        print(json.dumps({"world": world, "bytes": nb, "correct": ok, "ms_per_step": round(ms, 4),
                          "ms_all_reduce_only": round(ms_nccl, 4),
                          "bus_GBps": round(2 * (world - 1) / world * nb / ms_nccl / 1e6, 1)}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
