"""Dev: one scenario / method / fan size / output mode on device buffers, a few launches -- the command ncu wraps
(ncu -k regex:<kernel> -s <skip> -c 1).  usage: prof_trace.py <scenario 1-4> <method> <log2 rays> <final|full|stats|exit> [reps]

  final : d_ray + last row                  (trace_kernel<.., TRAJ=0, EXTRA=0>)
  full  : the reference's (max_size, 6, R)  (trace_kernel<.., TRAJ=1>)
  stats : final + online p_x statistics     (X_STATS flavour, what configs[2..3] run)
  exit  : final + exit points through rtb_shoot_coarse (X_EXIT flavour, what configs[4] runs), one DELTA_S candidate
"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import raytracing_b200 as rt

uc, m, lg, mode = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
R = 1 << lg
eng = rt.get_engine()
eng.set_field(*rt.build_field(uc))
c = rt.scenario_constants(uc)
dev = torch.device("cuda", 0)
th, x0, y0 = (torch.from_numpy(a).to(dev) for a in rt.make_fan(uc, R))
M = rt.max_size_for(uc, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1)
if mode == "exit":
    import numpy as np
    far = (np.array([1e300]), np.array([1e300]))
    for _ in range(reps):
        r = eng.shoot_coarse(m, th, (float(x0[0]), float(y0[0])), torch.tensor([rt.DELTA_S], device=dev, dtype=torch.float64),
                             torch.tensor([M], device=dev), (torch.tensor(far[0], device=dev), torch.tensor(far[1], device=dev)),
                             c[0], c[5:9])
    st = r.stats
    print("s%s op%d 2^%d %s: %.3f ms, %.4e ray-steps/s ray_steps=%d" % (uc, m, lg, mode, st["kernel_ms"], st["ray_steps"] / st["kernel_ms"] * 1e3,
                                                                        st["ray_steps"]))
    sys.exit(0)
kw = dict(full=(mode == "full"), want_px_stats=(mode == "stats"))
out = {}
if mode == "full":
    out["s_ray"] = torch.empty(M, 6, R, dtype=torch.float64, device=dev)
for _ in range(reps):
    r = eng.trace(m, th, x0, y0, rt.DELTA_S, M, c[0], c[5:9], out=out, **kw)
st = r.stats
print("s%s op%d 2^%d %s: %.3f ms, %.4e ray-steps/s ray_steps=%d" % (uc, m, lg, mode, st["kernel_ms"], st["ray_steps"] / st["kernel_ms"] * 1e3,
                                                                        st["ray_steps"]))
