import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracing_b200 as rt
from oracle import rt_oracle as O
for uc, m in (("4", 11), ("4", 10), ("3", 9)):
    z, grd = rt.build_field(uc); eng = rt.get_engine(0); eng.set_field(z, grd)
    c = rt.scenario_constants(uc); th, x0, y0 = rt.default_fan(uc)
    M = rt.max_size_for(uc, rt.DELTA_S, 91)
    r = eng.trace(m, th, x0, y0, rt.DELTA_S, M, c[0], c[5:9], full=True)
    o = O.trace(m, z, grd, rt.DELTA_S, M, c[0], c[5:9], False, th, x0, y0, nthreads=8)
    err = np.abs(r.s_ray - o["s_ray"]).max(axis=(0, 1))
    print(uc, m, "per-ray max err:", " ".join("%.1e" % e for e in err))
    k = int(np.argmax(err)); e = np.abs(r.s_ray - o["s_ray"])[:, :, k].max(axis=1)
    bad = np.nonzero(e > 1e-10)[0]
    print("  worst ray", k, "theta", th[k], "first bad row", bad[:3] if bad.size else None, "steps", r.d_ray[2, k], o["d_ray"][2, k])
    if bad.size:
        i = bad[0]
        print("  row", i, "gpu", r.s_ray[i, :, k], "\n        ora", o["s_ray"][i, :, k])
