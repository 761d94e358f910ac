"""Dev: one scenario / method / fan size, device buffers, FINAL mode, 3 launches (for ncu -k regex:trace_kernel -s 2)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import raytracing_b200 as rt

uc, m, R = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
eng = rt.get_engine()
eng.set_field(*rt.build_field(uc))
c = rt.scenario_constants(uc)
dev = torch.device("cuda", 0)
th, x0, y0 = (torch.from_numpy(a).to(dev) for a in rt.make_fan(uc, R))
M = rt.max_size_for(uc, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1)
for _ in range(3):
    r = eng.trace(m, th, x0, y0, rt.DELTA_S, M, c[0], c[5:9], full=False)
print(r.stats)
