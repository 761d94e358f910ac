"""Multi-GPU sharding of a ray fan: one process per GPU, contiguous blocks of ray indices per
rank, spline tables replicated, no data-path collective.  torch.distributed (NCCL on GPUs, gloo
in the CPU tests) is used only for the result gather and the tiny per-candidate reductions
(SURVEY.md 8(e)).  Rays are independent (the loop RT_bench.py:807 carries no state across k), so a
sharded run is bit-identical to the single-GPU run.
"""
import numpy as np


def shard_bounds(R, world, rank):
    """Contiguous block [lo, hi) of ray indices owned by `rank`; blocks differ by at most one ray."""
    base, rem = divmod(int(R), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def cyclic_indices(R, world, rank, block=4096):
    """Block-cyclic ownership: rank r owns blocks r, r+world, r+2*world, ... of `block` consecutive rays.
    Rays of a fan sorted by launch angle differ 3x in step count; contiguous shards would
    leave the ranks unevenly loaded, interleaved blocks give every rank the same mix."""
    nb = (int(R) + block - 1) // block
    mine = np.arange(rank, nb, world, dtype=np.int64)
    idx = (mine[:, None] * block + np.arange(block, dtype=np.int64)[None, :]).ravel()
    return idx[idx < R]


def gather_rays(local, R, group=None):
    """All-gather a per-ray tensor (..., R_local) from every rank into (..., R) (ragged shards ok)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    lead = local.shape[:-1]
    sizes = [shard_bounds(R, world, r)[1] - shard_bounds(R, world, r)[0] for r in range(world)]
    mx = max(sizes)
    pad = torch.zeros(*lead, mx, dtype=local.dtype, device=local.device)
    pad[..., :local.shape[-1]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad.contiguous(), group=group)
    return torch.cat([b[..., :n] for b, n in zip(bufs, sizes)], dim=-1)


def reduce_sum_max(sum_vals, max_vals, group=None):
    """All-reduce per-candidate partial sums (SUM) and maxima (MAX) in place; used for the mean/max
    Snell error of a sweep whose rays span ranks (RT_bench.py:954)."""
    import torch.distributed as dist
    dist.all_reduce(sum_vals, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(max_vals, op=dist.ReduceOp.MAX, group=group)
    return sum_vals, max_vals


def trace_sharded(engine, method, fan_fn, R, step, max_size, gamma, box, group=None, **kw):
    """Trace this rank's block of an R-ray fan (fan_fn(lo, hi) -> theta0, x0, y0 numpy) and gather
    d_ray / last_row / errors on every rank.  Returns dict of full-size torch tensors."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = shard_bounds(R, world, rank)
    th, x0, y0 = fan_fn(lo, hi)
    dev = torch.device("cuda", engine.device)
    th, x0, y0 = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (th, x0, y0))
    r = engine.trace(method, th, x0, y0, step, max_size, gamma, box, full=False, **kw)
    out = {k: gather_rays(r[k], R, group) for k in ("d_ray", "last_row", "errors", "px_stats") if r[k] is not None}
    out["stats"] = r.stats
    return out
