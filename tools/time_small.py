"""Latency of the reference's own small fans (configs[0]: 31-43 rays): kernel time and time per step of the longest
ray, FINAL and FULL output, default and faithful kernels.  Run on the GPU box."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import raytracing_b200 as rt

eng = rt.get_engine()
for uc, m in (("2", 6), ("2", 1), ("1", 6), ("3", 6), ("2", 9), ("4", 11)):
    eng.set_field(*rt.build_field(uc))
    c = rt.scenario_constants(uc)
    th, x0, y0 = rt.default_fan(uc)
    M = rt.max_size_for(uc, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1)
    for full in (False, True):
        for exact in (False, True):
            best, wall = 1e9, 1e9
            for _ in range(10):
                r = eng.trace(m, th, x0, y0, rt.DELTA_S, M, c[0], c[5:9], interface=(uc == "1"), full=full, exact=exact)
                best = min(best, r.stats["kernel_ms"])
                wall = min(wall, r.stats["total_ms"])
            steps = int(r.d_ray[2].max())
            print("s%s op%-2d R=%d full=%d exact=%d: kernel %.3f ms, call %.3f ms, longest ray %d steps -> %.2f us/step"
                  % (uc, m, len(th), full, exact, best, wall, steps, 1e3 * best / steps))
