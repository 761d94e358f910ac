"""CPU, world_size 2, gloo: the N>1 host logic (shard bounds, ragged result gather, per-candidate
reductions).  The per-shard compute is stood in by the oracle here (test infrastructure); on the GPU
box the same gather code runs over NCCL (tests/test_gpu_trace.py::test_sharded_equals_unsharded)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raytracing_b200.dist import gather_rays, reduce_sum_max, shard_bounds


def test_shard_bounds_cover():
    for R in (0, 1, 7, 42, 1 << 20):
        for w in (1, 2, 3, 8):
            b = [shard_bounds(R, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == R
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, R, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import raytracing_b200 as rt
    from oracle import rt_oracle as O
    uc = "3"
    z, grd = rt.build_field(uc)
    c = rt.scenario_constants(uc)
    M = rt.max_size_for(uc, rt.DELTA_S, 91)
    lo, hi = shard_bounds(R, world, rank)
    th, x0, y0 = rt.make_fan(uc, R, lo, hi)
    o = O.trace(6, z, grd, rt.DELTA_S, M, c[0], c[5:9], False, th, x0, y0, full=False)
    d_ray = gather_rays(torch.from_numpy(o["d_ray"]), R)
    last = gather_rays(torch.from_numpy(o["final"]), R)
    s = torch.tensor([o["d_ray"][1].sum()], dtype=torch.float64)
    mx = torch.tensor([o["d_ray"][1].max()], dtype=torch.float64)
    reduce_sum_max(s, mx)
    if rank == 0:
        q.put((d_ray.numpy(), last.numpy(), s.item(), mx.item()))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_gather_equals_unsharded():
    import raytracing_b200 as rt
    from oracle import rt_oracle as O
    R, world = 37, 2  # ragged: 19 + 18 rays
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, R, q)) for r in range(world)]
    for p in procs:
        p.start()
    d_ray, last, ssum, smax = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    uc = "3"
    z, grd = rt.build_field(uc)
    c = rt.scenario_constants(uc)
    th, x0, y0 = rt.make_fan(uc, R)
    o = O.trace(6, z, grd, rt.DELTA_S, rt.max_size_for(uc, rt.DELTA_S, 91), c[0], c[5:9], False, th, x0, y0, full=False)
    assert np.array_equal(d_ray, o["d_ray"]) and np.array_equal(last, o["final"])  # bit-identical
    assert abs(ssum - o["d_ray"][1].sum()) < 1e-9 and smax == o["d_ray"][1].max()
