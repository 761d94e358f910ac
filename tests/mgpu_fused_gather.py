"""torchrun script (not a pytest module): FusedGather across processes == NCCL all-gather of d_ray, bit for bit."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import raytracing_b200 as rt
from raytracing_b200.dist import FusedGather, gather_rays, shard_bounds

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
eng = rt.Engine(local)
uc, R = "3", 100_003
eng.set_field(*rt.build_field(uc))
c = rt.scenario_constants(uc)
M = rt.max_size_for(uc, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1)
lo, hi = shard_bounds(R, world, rank)
th, x0, y0 = (torch.from_numpy(a).to(dev) for a in rt.make_fan(uc, R, lo, hi))
fg = FusedGather(eng, R, lo)
for _ in range(2):   # twice: the arrays are reusable
    fg.view.fill_(float("nan"))
    dist.barrier()
    r = eng.trace(6, th, x0, y0, rt.DELTA_S, M, c[0], c[5:9], full=False)
    fused = fg.finish().clone()
    want = gather_rays(r.d_ray, R)
    assert torch.equal(fused, want), "rank %d: fused gather differs from the NCCL all-gather" % rank
fg.close()
dist.barrier()
if rank == 0:
    print("fused gather ok", world, "ranks", R, "rays")
dist.destroy_process_group()
