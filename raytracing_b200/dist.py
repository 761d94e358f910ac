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


class _DevArray:
    """A (3, R) float64 device buffer owned by librtb, exposed to torch through __cuda_array_interface__."""

    def __init__(self, ptr, R):
        self.__cuda_array_interface__ = dict(shape=(3, int(R)), typestr="<f8", data=(int(ptr), False), version=3,
                                             strides=None)


class FusedGather:
    """Result gather fused into the integration kernel: every rank's trace kernel stores a finished
    ray's d_ray column straight into the (3, R_total) array of every GPU of the node over NVLink peer mappings
    (CUDA IPC between the per-GPU processes), so no all-gather follows the kernel -- ranks only agree that they are
    done (`finish`: a one-element all-reduce).  Same bytes as `gather_rays(d_ray)`.

    Shards are contiguous (`offset` = this rank's first ray) or block-cyclic (`cyclic_block` = the block of
    `cyclic_indices`, a power of two: local ray r is global ray ((r // block) * world + rank) * block + r % block).

    The arrays are DOUBLE-BUFFERED: trace k writes set k % 2 on every GPU, so a rank that races ahead into trace k+1
    stores into the other set while slower ranks still read the view `finish` gave them for trace k; it cannot reach
    trace k+2 (which reuses set k % 2) before every rank has entered finish() of trace k+1, i.e. -- stream order --
    after every rank's work on view k queued before its trace k+1.  Consumers must therefore use a view (on the
    engine's stream) before they launch the next trace but one.

        fg = FusedGather(engine, R_total, lo, group)      # collective: exchanges IPC handles
        engine.trace(...device tensors of rays lo..hi...) # writes every peer
        d_ray_all = fg.finish()                           # torch view (3, R_total) of this rank's copy of this trace
        fg.close()                                        # collective
    """

    def __init__(self, engine, R_total, offset=0, group=None, cyclic_block=None):
        import torch
        import torch.distributed as dist
        self.engine, self.group, self.R = engine, group, int(R_total)
        engine.use_torch_stream()      # the kernels and the finishing all-reduce must be ordered on one stream
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > 8:
            raise ValueError("FusedGather: at most 8 GPUs (one node)")
        self.offset, self.cyclic_block = int(offset), cyclic_block
        self._flag = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", engine.device))
        self.own, self.bases, err = [None, None], [[], []], None
        handles_mine = [None, None]
        try:
            for b in range(2):
                self.own[b], handles_mine[b] = engine.gather_alloc(3 * 8 * self.R)
        except Exception as e:  # the outcome is agreed on below: no rank may leave the others in a collective
            handles_mine, err = None, e
        handles = [None] * self.world
        dist.all_gather_object(handles, handles_mine, group=group)
        if err is None and all(h is not None for h in handles):
            try:
                for b in range(2):
                    self.bases[b] = [self.own[b] if r == self.rank else engine.gather_open(handles[r][b])
                                     for r in range(self.world)]
            except Exception as e:
                err = e
        ok = torch.tensor([0 if err is not None or not self.bases[1] else 1], dtype=torch.int32, device=self._flag.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok.item()) == 0:     # some rank could not allocate / map (no peer access): nobody uses the fused path
            self._release()
            raise RuntimeError("FusedGather: peer mapping unavailable on at least one rank (%s)" % (err,))
        self.views = [torch.as_tensor(_DevArray(self.own[b], self.R), device=self._flag.device) for b in range(2)]
        self._cur = 0
        self._arm()
        dist.barrier(group=group)      # every peer has mapped these buffers before anyone traces

    def _arm(self):
        """Point the engine's next trace at buffer set self._cur."""
        if self.cyclic_block:
            self.engine.set_gather_cyclic(self.bases[self._cur], self.R, self.cyclic_block, self.world, self.rank)
        else:
            self.engine.set_gather(self.bases[self._cur], self.R, self.offset)

    def _release(self):
        for b in range(2):
            for r, p in enumerate(self.bases[b]):
                if r != self.rank:
                    self.engine.gather_close(p, own=False)
            if self.own[b] is not None:
                self.engine.gather_close(self.own[b], own=True)
        self.bases, self.own = [[], []], [None, None]

    def finish(self):
        """All ranks' kernels of this trace have completed (stream-ordered after this rank's own): the local array
        is whole.  Returns its view and points the next trace at the other buffer set."""
        import torch.distributed as dist
        dist.all_reduce(self._flag, group=self.group)
        view = self.views[self._cur]
        self._cur ^= 1
        self._arm()
        return view

    def close(self):
        import torch
        import torch.distributed as dist
        torch.cuda.synchronize(self.engine.device)
        self.engine.set_gather(None, 0, 0)
        dist.barrier(group=self.group)  # nobody is still writing into a buffer about to be unmapped / freed
        self._release()


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
