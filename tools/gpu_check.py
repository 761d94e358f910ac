"""Dev script: GPU vs oracle on the default fans of every (scenario, method), plus a timing probe."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracing_b200 as rt
from oracle import rt_oracle as O

eng = rt.get_engine(0)
print("fp64 peak TFLOP/s", eng.fp64_peak() / 1e12, flush=True)
for uc in "2314":
    z, grd = rt.build_field(uc)
    eng.set_field(z, grd)
    c = rt.scenario_constants(uc)
    gamma, box = c[0], c[5:9]
    theta0, x0, y0 = rt.default_fan(uc)
    if uc in "14":
        theta0, x0, y0 = theta0[::5], x0[::5], y0[::5]
    # field parity
    rng = np.random.default_rng(1)
    px = rng.uniform(box[0] - .1, box[1] + .1, 20000); py = rng.uniform(box[2] - .1, box[3] + .1, 20000)
    on, og = O.n_gradient(np.stack([px, py], 1), grd, z)
    for ex in (True, False):
        n, gx, gy = eng.n_gradient(px, py, exact=ex)
        print(uc, "field exact=%d" % ex, np.abs(n - on).max(), np.abs(gx - og[:, 0]).max(), np.abs(gy - og[:, 1]).max(), flush=True)
    step = rt.DELTA_S
    M = rt.max_size_for(uc, step, 91)
    for m in ([10, 11] if uc == "4" else range(1, 10)):
        o = O.trace(m, z, grd, step, M, gamma, box, uc == "1", theta0, x0, y0, nthreads=8)
        for ex in (True, False):
            r = eng.trace(m, theta0, x0, y0, step, M, gamma, box, interface=(uc == "1"), full=True, exact=ex, want_px_stats=True)
            r2 = eng.trace(m, theta0, x0, y0, step, M, gamma, box, interface=(uc == "1"), full=False, exact=ex)
            print(uc, "op%d exact=%d" % (m, ex), "steps_eq", np.array_equal(r.d_ray[2], o["d_ray"][2]),
                  "s_ray %.2e" % np.abs(r.s_ray - o["s_ray"]).max(), "d_ray %.2e" % np.abs(r.d_ray - o["d_ray"]).max(),
                  "err %.2e" % np.abs(r.errors - o["errors"]).max() if r.errors is not None else "",
                  "final-vs-full %.1e" % np.abs(r2.last_row - r.last_row).max(),
                  "err2 %.1e" % (np.abs(r2.errors - r.errors).max() if r.errors is not None else 0),
                  "ms %.2f" % r2.stats["kernel_ms"], flush=True)

# timing probe: fisheye 1M rays
import torch
uc = "2"
z, grd = rt.build_field(uc); eng.set_field(z, grd)
c = rt.scenario_constants(uc); gamma, box = c[0], c[5:9]
R = 1 << 20
th, x0, y0 = rt.make_fan(uc, R)
th, x0, y0 = (torch.from_numpy(a).cuda() for a in (th, x0, y0))
M = 910
for m in (6, 1, 2, 3, 4, 7, 8, 9, 5):
    for it in range(2):
        r = eng.trace(m, th, x0, y0, rt.DELTA_S, M, gamma, box, full=False)
    print("fisheye 1M op%d FINAL: %.2f ms, %.3e ray-steps/s" % (m, r.stats["kernel_ms"], r.stats["ray_steps"] / r.stats["kernel_ms"] * 1e3), flush=True)
s = torch.empty(M, 6, R, dtype=torch.float64, device="cuda")
for it in range(2):
    r = eng.trace(6, th, x0, y0, rt.DELTA_S, M, gamma, box, full=True, out=dict(s_ray=s))
print("fisheye 1M op6 FULL: %.2f ms, %.3e ray-steps/s, %.1f GB/s" % (r.stats["kernel_ms"], r.stats["ray_steps"] / r.stats["kernel_ms"] * 1e3, r.stats["ray_steps"] * 48 / r.stats["kernel_ms"] / 1e6))
