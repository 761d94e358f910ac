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


class _OracleShootEngine:
    """Stand-in for Engine in the CPU test of the sharded two-point search: same shoot_coarse / shoot_refine
    contract on torch CPU tensors, computed by the oracle (test infrastructure)."""
    torch_device = torch.device("cpu")
    device = 0

    def __init__(self, z, grd):
        self.z, self.grd = z, grd

    def set_field(self, z, grd):
        pass

    def _tracers(self, method, src, steps, max_sizes, gamma, box):
        from oracle import shoot_oracle as S
        return [S.Tracer(method, self.z, self.grd, float(s), int(m), gamma, box, src) for s, m in zip(steps, max_sizes)]

    def shoot_coarse(self, method, angles, src, steps, max_sizes, targets, gamma, box, index0=0, want_exit=False, exact=False):
        from raytracing_b200.engine import TraceResult
        tx, ty = targets[0].numpy(), targets[1].numpy()
        miss2, index = [], []
        for tr in self._tracers(method, src, steps.numpy(), max_sizes.numpy(), gamma, box):
            ex, _ = tr(angles.numpy())
            m2 = (ex[0][None, :] - tx[:, None]) ** 2 + (ex[1][None, :] - ty[:, None]) ** 2
            j = np.argmin(m2, axis=1)
            miss2.append(m2[np.arange(tx.size), j]); index.append(j + index0)
        return TraceResult(miss2=torch.from_numpy(np.array(miss2)), index=torch.from_numpy(np.array(index, dtype=np.int64)),
                           exit=None, stats=dict(kernel_ms=0.0, ray_steps=0, launches=0))

    def shoot_refine(self, method, lo, hi, src, steps, max_sizes, targets, gamma, box, exact=False):
        from oracle import shoot_oracle as S
        from raytracing_b200.engine import TraceResult
        tx, ty = targets[0].numpy(), targets[1].numpy()
        Cn, Tn = lo.shape
        theta, res, iters = np.zeros((Cn, Tn)), np.zeros((5, Cn, Tn)), 0
        for c, tr in enumerate(self._tracers(method, src, steps.numpy(), max_sizes.numpy(), gamma, box)):
            for t in range(Tn):
                f = lambda th: float(((tr([th])[0][:2, 0] - np.array([tx[t], ty[t]])) ** 2).sum())
                theta[c, t], n = S.golden(f, float(lo[c, t]), float(hi[c, t]))
                e, ns = tr([theta[c, t]])
                res[:, c, t] = np.hypot(e[0, 0] - tx[t], e[1, 0] - ty[t]), e[2, 0], e[0, 0], e[1, 0], ns[0]
                iters = max(iters, n)
        return TraceResult(theta=torch.from_numpy(theta), result=torch.from_numpy(res), iterations=iters,
                           stats=dict(kernel_ms=0.0, ray_steps=0, launches=0))


def _shoot_problem():
    import raytracing_b200 as rt
    steps = np.array([rt.SIGMA / 2.0, rt.SIGMA / 1.5])
    angles = np.linspace(0.0, np.pi / 2, 21)
    tx, ty = np.array([4.2, 2.0, 0.5]), np.array([1.0, 1.0, 1.0])  # points on the top edge y = 1 of the vert box
    return steps, angles, (tx, ty)


def _shoot_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import raytracing_b200 as rt
    z, grd = rt.build_field("3")
    steps, angles, targets = _shoot_problem()
    r = rt.shooting_search(6, z, grd, "3", targets, angles, steps=steps, engine=_OracleShootEngine(z, grd), group=True)
    if rank == 0:
        q.put({k: r[k] for k in ("theta", "miss", "T", "index", "nsteps", "iterations", "coarse_miss")})
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_shooting_search_equals_unsharded():
    """Angle-sharded coarse fan + MIN/argmin all-reduce + target-sharded refinement + gather (world 2, gloo)
    == the single-process oracle search."""
    import raytracing_b200 as rt
    from oracle import shoot_oracle as S
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_shoot_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=600)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    z, grd = rt.build_field("3")
    c = rt.scenario_constants("3")
    steps, angles, targets = _shoot_problem()
    msz = rt.candidate_rows("3", steps)
    for ci in range(2):
        tr = S.Tracer(6, z, grd, steps[ci], int(msz[ci]), c[0], c[5:9], (-2.0, -2.0))
        o = S.shooting_search(tr, targets, angles)
        assert np.array_equal(got["index"][ci], o["index"]) and np.array_equal(got["nsteps"][ci], o["nsteps"])
        assert np.array_equal(got["theta"][ci], o["theta"]) and np.array_equal(got["T"][ci], o["T"])
        assert np.allclose(got["coarse_miss"][ci], o["coarse_miss"], rtol=0, atol=1e-15)
    assert got["iterations"] >= 25


def test_reduce_argmin_ties_to_lowest_index():
    """Single-process group: identity; and the tie rule on hand-made shards."""
    from raytracing_b200.shooting import INT64_MAX
    a = (torch.tensor([[1.0, 4.0, 2.0]], dtype=torch.float64), torch.tensor([[3, 9, 5]]))
    b = (torch.tensor([[1.0, 3.0, 2.5]], dtype=torch.float64), torch.tensor([[12, 10, 11]]))
    gmin = torch.minimum(a[0], b[0])
    cand = torch.minimum(torch.where(a[0] == gmin, a[1], INT64_MAX), torch.where(b[0] == gmin, b[1], INT64_MAX))
    assert cand.tolist() == [[3, 10, 5]]
