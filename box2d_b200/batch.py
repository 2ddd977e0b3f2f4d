"""Batches of independent worlds across GPUs (BASELINE.json configs[4]).

Worlds are independent units (api/World.cs:9-10), so ranks own contiguous world-index ranges and never exchange
body data; the only collective is a reduction of a few batch statistics / timings (NCCL on GPUs, gloo in CPU tests).
"""
import numpy as np


def shard_range(n_worlds, rank, world_size):
    """Contiguous [begin, end) of world indices owned by `rank`: ranges differ by at most one world."""
    base, extra = divmod(n_worlds, world_size)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def batch_stats(bodies):
    """Per-rank statistics vector from a ReadBodies() array: [n_bodies, sum |v|^2 (kinetic proxy), max |v|, min y]."""
    vx, vy = bodies["linearVelocity"]["x"].astype(np.float64), bodies["linearVelocity"]["y"].astype(np.float64)
    v2 = vx * vx + vy * vy
    return np.array([len(bodies), v2.sum(), np.sqrt(v2.max()) if len(bodies) else 0.0,
                     bodies["transform"]["p"]["y"].min() if len(bodies) else 0.0], dtype=np.float64)


def reduce_stats(stats, dist=None, device=None):
    """All-reduce of batch_stats vectors: sums for [0:2], max for [2], min for [3]. `dist` = torch.distributed or None."""
    if dist is None or not dist.is_initialized():
        return np.array(stats, dtype=np.float64)
    import torch
    t = torch.tensor(stats, dtype=torch.float64, device=device)
    s, mx, mn = t[0:2].clone(), t[2:3].clone(), t[3:4].clone()
    dist.all_reduce(s, op=dist.ReduceOp.SUM)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    return torch.cat([s, mx, mn]).cpu().numpy()


def reduce_timing(ms_values, units, dist=None, device=None):
    """Max-over-ranks of device times, sum-over-ranks of processed units (the bench.py contract)."""
    if dist is None or not dist.is_initialized():
        return [float(x) for x in ms_values], float(units)
    import torch
    t = torch.tensor(list(ms_values), dtype=torch.float64, device=device)
    u = torch.tensor([float(units)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return [float(x) for x in t.tolist()], float(u.item())
