"""torchrun script (not a pytest module; tests/test_gpu_trace.py::test_fused_gather_across_processes launches it on a
box with >= 2 GPUs):
  1. FusedGather across processes == NCCL all-gather of d_ray, bit for bit, contiguous shards;
  2. the arrays are double-buffered: with inputs that CHANGE from trace to trace and ranks of deliberately different
     speed, the view finish() returned for trace k still holds trace k after the peers have run trace k+1;
  3. block-cyclic shards (rtb_set_gather_cyclic) land at their global indices;
  4. the DELTA_S sweep sharded over the ranks (api.sweep_metrics(group=...): rtb_sweep partial sums + reduce_sum_max
     over NCCL) == the unsharded sweep."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import raytracing_b200 as rt
from raytracing_b200.dist import FusedGather, cyclic_indices, gather_rays, shard_bounds
from raytracing_b200.fans import fan_at

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
eng = rt.Engine(local)
uc, R = "3", 100_003
z, grd = rt.build_field(uc)
eng.set_field(z, grd)
c = rt.scenario_constants(uc)
M = rt.max_size_for(uc, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1)

# 1 + 2: contiguous shards, inputs that change per trace
lo, hi = shard_bounds(R, world, rank)
th0, x0, y0 = (torch.from_numpy(a).to(dev) for a in rt.make_fan(uc, R, lo, hi))
fg = FusedGather(eng, R, lo)
views, wants = [], []
for k in range(4):
    th = th0 * (1.0 - 0.01 * k)                       # a different fan every trace
    if rank == k % world:
        time.sleep(0.05)                              # ranks out of step: a fast rank runs ahead into the next trace
    r = eng.trace(6, th, x0, y0, rt.DELTA_S, M, c[0], c[5:9], full=False)
    v = fg.finish()
    want = gather_rays(r.d_ray, R)
    assert torch.equal(v, want), "rank %d: fused gather differs from the NCCL all-gather (trace %d)" % (rank, k)
    views.append(v)
    wants.append(want)
    if k >= 1:   # the previous view was not touched by this trace (it went to the other buffer set)
        assert torch.equal(views[k - 1], wants[k - 1]), "rank %d: view of trace %d overwritten by trace %d" % (rank, k - 1, k)
fg.close()
dist.barrier()

# 3: block-cyclic shards
block = 1024
idx = cyclic_indices(R, world, rank, block)
th, xx, yy = (torch.from_numpy(a).to(dev) for a in fan_at(uc, R, idx))
fg = FusedGather(eng, R, cyclic_block=block)
r = eng.trace(6, th, xx, yy, rt.DELTA_S, M, c[0], c[5:9], full=False)
v = fg.finish()
whole = eng.trace(6, *(torch.from_numpy(a).to(dev) for a in rt.make_fan(uc, R)), rt.DELTA_S, M, c[0], c[5:9], full=False)
assert torch.equal(v, whole.d_ray), "rank %d: block-cyclic fused gather differs from the unsharded trace" % rank
fg.close()
dist.barrier()

# 4: the DELTA_S sweep sharded over the ranks
for suc, m in (("1", 6), ("3", 6)):
    sz, sgrd = rt.build_field(suc)
    divisors, options, _ = rt.sweep_candidates(suc)
    sel = np.linspace(0, len(options) - 1, 9).astype(int)
    one = rt.sweep_metrics(m, sz, sgrd, suc, options[sel], divisors[sel], engine=eng)
    shard = rt.sweep_metrics(m, sz, sgrd, suc, options[sel], divisors[sel], engine=eng, group=dist.group.WORLD)
    a, b = np.asarray(one, dtype=np.float64), np.asarray(shard, dtype=np.float64)
    assert np.allclose(a, b, rtol=1e-12, atol=0), "rank %d: sharded sweep differs (scenario %s): %g" % (rank, suc, np.abs(a - b).max())
    if suc == "1":
        assert np.array_equal(a[1], b[1])             # the maxima are exact
dist.barrier()
if rank == 0:
    print("fused gather ok", world, "ranks", R, "rays; double buffer, block-cyclic map and sharded sweep ok")
dist.destroy_process_group()
