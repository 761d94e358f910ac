"""Measured deviations of the CUDA path from the reference's own trajectories (tests/golden/ref_traj.npz, generated
by importing the unmodified RT_bench.py): for every (scenario, method) of the default fans, the largest scale-aware
difference max |a - b| / max(1, |b|) over the stored rows of s_ray, over d_ray[0:2] and the Snell errors, the number of
rays whose step count differs, and the tolerance the tests apply (1e-10, or 30 x the reference algorithm's own
sensitivity to 1-ulp trig noise where that is larger, tests/_util.py).  Both kernel families.  Run on the GPU box."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", "tests"))
import raytracing_b200 as rt
from _util import SCEN_METHODS, scenario_setup, tolerances

G = np.load(os.path.join(HERE, "..", "tests", "golden", "ref_traj.npz"))
SENS = json.load(open(os.path.join(HERE, "..", "tests", "golden", "sensitivity.json")))
eng = rt.get_engine()


def dev(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b)))) if a.size else 0.0


print("%-10s %-7s %12s %12s %12s %6s %10s" % ("case", "kernels", "s_ray rows", "d_ray", "errors", "steps", "test tol"))
for uc in "1234":
    s = scenario_setup(uc)
    eng.set_field(s["z"], s["grd"])
    for m in SCEN_METHODS[uc]:
        key = "default_s%s_op%d/" % (uc, m)
        tol = tolerances(SENS, uc, m)
        for exact in (False, True):
            r = eng.trace(m, s["theta0"], s["x0"], s["y0"], s["step"], s["M"], s["gamma"], s["box"],
                          interface=(uc == "1"), full=True, exact=exact)
            ref_d = G[key + "d_ray"]
            same = r.d_ray[2] == ref_d[2]
            stride = int(G[key + "stride"])
            e = dev(r.errors[same], G[key + "errors"][same]) if uc == "1" else 0.0
            print("s%s op%-5d %-7s %12.2e %12.2e %12.2e %6d %10.1e" % (
                uc, m, "exact" if exact else "default", dev(r.s_ray[::stride][:, :, same], G[key + "rows"][:, :, same]),
                dev(r.d_ray[:2, same], ref_d[:2, same]), e, int((~same).sum()), tol["all"]))
