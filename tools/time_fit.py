"""interpolacion(): scipy on the host cores (the reference's path, RT_bench.py:435-464) vs rtb_fit_field on the GPU.
Prints one JSON line per scenario grid; run on the GPU box."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import raytracing_b200 as rt
from raytracing_b200 import constants as K

eng = rt.get_engine()
for uc in ("1", "2", "3"):
    lim = K.constants(uc)[5:9]
    x, y, _X, _Y, Z = rt.genZ(lim[0], lim[1], lim[2], lim[3], K.SCENARIO_FUNCS[uc])
    host = []
    for _ in range(3):
        t = time.perf_counter()
        zs, gs = rt.interpolacion(x, y, Z)
        host.append(time.perf_counter() - t)
    t = time.perf_counter()
    eng.set_field(zs, gs)
    upload = time.perf_counter() - t
    dev, kern, wall = [], [], []
    for _ in range(5):
        t = time.perf_counter()
        z, grd = eng.fit_field(x, y, Z, K.DELTA)
        wall.append(time.perf_counter() - t)
        dev.append(z.fit["stats"]["total_ms"])
        kern.append(z.fit["stats"]["kernel_ms"])
    c = grd[0].tck[2]
    print(json.dumps(dict(scenario=uc, grid=list(Z.shape), scipy_fit_s=min(host), scipy_upload_s=upload,
                          device_fit_wall_s=min(wall), device_total_ms=min(dev), device_kernels_ms=min(kern),
                          max_coef_diff=float(np.abs(c - gs[0].tck[2]).max()),
                          coef_scale=float(np.abs(c).max()))))
