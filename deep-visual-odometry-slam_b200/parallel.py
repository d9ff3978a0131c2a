"""Multi-GPU plumbing: one process per GPU over torch.distributed (NCCL on the box, gloo in CPU tests).

SURVEY.md §8(e):
  * inference shards by independent sequence — no data-path collective; the only communication is
    the barrier and the max-over-ranks of the device time (`barrier`, `max_over_ranks`,
    `shard_sequences`);
  * training is data parallel over clips (the reference trains with batch size 1 on one GPU,
    train.py:~150-200): one all-reduce per step of a single flat fp32 gradient buffer that the
    parameters' `.grad` tensors are views of (no pack / unpack copies), with the AMP GradScaler's
    found-inf flag folded into the same buffer so the skip decision is identical on every rank
    (`FlatGradAllReducer`);
  * global BA does not shard (one Cholesky factorisation): replicas only.
"""
import os

import torch
import torch.distributed as dist


def init(backend=None, device=None):
    """Join the process group described by RANK / WORLD_SIZE / MASTER_* (torchrun).  Returns
    (rank, world).  A single process (WORLD_SIZE unset or 1) needs no group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        kw = {"device_id": device} if (backend == "nccl" and device is not None) else {}
        dist.init_process_group(backend, **kw)
    return rank, world


def world_size():
    return dist.get_world_size() if dist.is_initialized() else 1


def barrier(device=None):
    if dist.is_initialized():
        dist.barrier()
    if device is not None and torch.device(device).type == "cuda":
        torch.cuda.synchronize(device)


def max_over_ranks(value, device="cpu"):
    """Max of a python float over all ranks (device time of a step: a job is as slow as its slowest rank)."""
    if not dist.is_initialized():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, device="cpu"):
    if not dist.is_initialized():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def shard_sequences(n_sequences, rank, world):
    """Sequence ids owned by `rank`: contiguous blocks, sizes differing by at most one, every
    sequence owned exactly once (inference shards by independent sequence, no collective)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(int(n_sequences), world)
    start = rank * base + min(rank, extra)
    return list(range(start, start + base + (1 if rank < extra else 0)))


class FlatGradAllReducer:
    """Data-parallel gradient exchange through ONE flat fp32 buffer.

    `params`: iterable of parameters (same order and shapes on every rank).  Their `.grad` tensors
    are (re)pointed at slices of one contiguous buffer, so backward writes straight into it and
    `allreduce()` is a single collective over `numel + 1` floats — the extra slot carries the
    found-inf flag of the AMP GradScaler (any rank's overflow makes every rank skip the step).
    """

    def __init__(self, params, device=None):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("FlatGradAllReducer: no trainable parameters")
        device = device or self.params[0].device
        self.numel = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(self.numel + 1, dtype=torch.float32, device=device)
        self.slices = []
        off = 0
        for p in self.params:
            if p.dtype != torch.float32:
                raise ValueError("FlatGradAllReducer: master parameters must be float32")
            self.slices.append(self.flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        self.rebinds = 0          # how many gradients had to be copied back into the buffer (instrumentation)
        self._bind(copy=False)

    def _bind(self, copy=True):
        """Make every `p.grad` a view of the flat buffer again.  `optimizer.zero_grad()` defaults to
        set_to_none=True (the reference calls it, train.py:295), after which backward allocates fresh
        gradient tensors OUTSIDE the buffer; those are copied into their slices here and re-pointed, so the
        collective never runs over a stale buffer.  A parameter that received no gradient gets zeros."""
        for p, sl in zip(self.params, self.slices):
            g = p.grad
            if g is not None and g.data_ptr() == sl.data_ptr() and g.shape == sl.shape and g.is_contiguous():
                continue
            if copy:
                if g is None:
                    sl.zero_()
                else:
                    sl.copy_(g)
                self.rebinds += 1
            p.grad = sl

    def zero_grad(self):
        """Use this (or `optimizer.zero_grad(set_to_none=False)`) instead of a bare `optimizer.zero_grad()` to keep
        backward writing straight into the buffer; the bare call still works, at the cost of one copy per step."""
        self.flat.zero_()
        self._bind(copy=False)

    def allreduce(self, found_inf=None, average=True, sanitize=True):
        """Sum (or mean) the gradients over all ranks in place.  `found_inf`: this rank's overflow
        flag (bool / 0-d tensor) or None to derive it from the gradients.  Returns the global flag
        as a 0-d tensor on the buffer's device (no host synchronisation).
        sanitize=True zeroes non-finite gradients before the sum (the caller skips the step on the returned flag);
        sanitize=False leaves them in, so inf/NaN reach every rank and `GradScaler.unscale_` finds them itself —
        the plain-DDP behaviour the reference's AMP loop (train.py:341-345) needs with no further change."""
        self._bind(copy=True)
        g = self.flat[:self.numel]
        if found_inf is None:
            flag = (~torch.isfinite(g).all()).to(torch.float32)
        else:
            flag = torch.as_tensor(found_inf, device=self.flat.device).to(torch.float32).reshape(())
        # keep non-finite values out of the sum: the step is skipped anyway, and NaN * 0 != 0
        if found_inf is None and sanitize:
            torch.nan_to_num_(g, nan=0.0, posinf=0.0, neginf=0.0)
        self.flat[self.numel] = flag
        if dist.is_initialized():
            if average and dist.get_backend() == "nccl":
                # NCCL averages inside the collective (no extra pass over the buffer); the flag slot becomes
                # (ranks that overflowed) / world, still > 0 exactly when any rank overflowed
                dist.all_reduce(self.flat, op=dist.ReduceOp.AVG)
            else:
                dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
                if average:
                    g.mul_(1.0 / dist.get_world_size())
        return self.flat[self.numel] > 0

    def nbytes(self):
        return self.flat.numel() * 4
