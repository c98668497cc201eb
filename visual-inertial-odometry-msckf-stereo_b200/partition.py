"""Multi-GPU plumbing: independent filters are block-partitioned over the ranks of one node (one process per
GPU); there is no collective on the data path, only a final gather of trajectories and timings
(SURVEY.md section 8e).  Works with any torch.distributed backend (nccl on the GPUs, gloo in the CPU tests)."""
import numpy as np


def filter_range(rank, world, total):
    """Contiguous block of filter indices owned by `rank`: filter k lives on rank floor(k * world / total)."""
    if world < 1 or not (0 <= rank < world) or total < 0:
        raise ValueError('bad partition arguments')
    lo = -(-rank * total // world)          # ceil(rank * total / world)
    hi = -(-(rank + 1) * total // world)
    return lo, hi


def owner_of(k, world, total):
    return (k * world) // total


def gather_blocks(local, world, rank, dst=0, device=None):
    """Gathers per-rank arrays that differ only in their leading (filter) dimension on rank `dst`, in filter order.
    `local` is a torch tensor; returns the concatenated tensor on dst, None elsewhere.  world == 1 needs no
    process group."""
    import torch
    if world == 1:
        return local
    import torch.distributed as dist
    n = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    pad = max(sizes)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[:local.shape[0]] = local
    out = [torch.zeros_like(buf) for _ in range(world)] if rank == dst else None
    dist.gather(buf, out, dst=dst)
    if rank != dst:
        return None
    return torch.cat([o[:s] for o, s in zip(out, sizes)], dim=0)


def reduce_max(x, world, device=None):
    """Max over ranks of a python float (device timings are always reported as the max over ranks)."""
    import torch
    if world == 1:
        return float(x)
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def reduce_sum(x, world, device=None):
    import torch
    if world == 1:
        return float(x)
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
