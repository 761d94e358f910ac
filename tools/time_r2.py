"""Dev: A/B timings of the round-2 kernel paths on one GPU (kernel ms from rtb_stats, best of 3).
  ray queue vs static grid : fisheye 2^20, vert 2^22 (+ p_x stats), interface 2^18, shuffled vert 2^20
  trajectory store         : fisheye 2^20 FULL (45.8 GB), staged + TMA vs per-lane stores
"""
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
import raytracing_b200 as rt

eng = rt.get_engine()
dev = torch.device("cuda", 0)


def best(fn, n=3):
    rs = [fn() for _ in range(n)]
    r = min(rs, key=lambda r: r.stats["kernel_ms"])
    return r.stats["kernel_ms"], r.stats["ray_steps"]


def fan(uc, lg, shuffle=False):
    th, x0, y0 = rt.make_fan(uc, 1 << lg)
    if shuffle:
        th = np.random.default_rng(1).permutation(th)
    return [torch.from_numpy(a).to(dev) for a in (th, x0, y0)]


for uc, lg, m, kw, tag in (("2", 20, 6, {}, "fisheye 2^20 op6 final"), ("2", 20, 1, {}, "fisheye 2^20 op1 final"),
                           ("3", 22, 6, {}, "vert 2^22 op6 final"), ("3", 22, 6, dict(want_px_stats=True), "vert 2^22 op6 stats"),
                           ("1", 18, 6, {}, "interface 2^18 op6 final"), ("1", 18, 6, dict(interface=True), "interface 2^18 op6 +snell"),
                           ("3", 20, 6, dict(shuffle=True), "vert 2^20 op6 SHUFFLED"),
                           ("3", 20, 6, dict(shuffle=True, want_px_stats=True), "vert 2^20 op6 SHUFFLED stats"),
                           ("4", 18, 11, dict(want_px_stats=True), "aniso 2^18 op11 stats"),
                           ("2", 20, 9, {}, "fisheye 2^20 op9 final")):
    eng.set_field(*rt.build_field(uc))
    c = rt.scenario_constants(uc)
    kw = dict(kw)
    a = fan(uc, lg, kw.pop("shuffle", False))
    M = rt.max_size_for(uc, rt.DELTA_S, 91)
    for mode, name in ((2, "static"), (1, "queue ")):
        eng.set_option("queue", mode)
        ms, steps = best(lambda: eng.trace(m, *a, rt.DELTA_S, M, c[0], c[5:9], full=False, **kw))
        print("%-32s %s %9.3f ms  %.4e ray-steps/s" % (tag, name, ms, steps / ms * 1e3), flush=True)
eng.set_option("queue", 0)

# the remedy for an unsorted fan: trace it in sorted order (Engine.trace(sort_fan=True) = RTB_F_SORT_FAN: radix sorts + gather + scatter inside rtb_trace)
uc = "3"
eng.set_field(*rt.build_field(uc))
c = rt.scenario_constants(uc)
a = fan(uc, 20, True)
M = rt.max_size_for(uc, rt.DELTA_S, 91)
for name, kw in (("as given ", {}), ("sort_fan ", dict(sort_fan=True))):
    for _ in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r = eng.trace(6, *a, rt.DELTA_S, M, c[0], c[5:9], full=False, **kw)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("%-32s %s %9.3f ms wall (kernel %.3f)  %.4e ray-steps/s" % ("vert 2^20 op6 SHUFFLED", name, dt * 1e3, r.stats["kernel_ms"],
                                                                       r.stats["ray_steps"] / dt), flush=True)

# mixed DELTA_S candidates inside every warp (per-ray step / max_size): where lanes of one warp differ widely in length
uc = "3"
eng.set_field(*rt.build_field(uc))
c = rt.scenario_constants(uc)
for lg in (18, 20):
    a = fan(uc, lg)
    R = 1 << lg
    div = np.random.default_rng(6).choice([0.5, 1.0, 2.0, 20.0], R)
    steps = rt.SIGMA / div
    msz = np.ceil(80 / steps).astype(np.int64) + 1
    d_steps, d_msz = torch.from_numpy(steps).to(dev), torch.from_numpy(msz).to(dev)
    for mode, name in ((2, "static"), (1, "queue ")):
        eng.set_option("queue", mode)
        ms, nst = best(lambda: eng.trace(6, *a, None, int(msz.max()), c[0], c[5:9], full=False, steps=d_steps, max_sizes=d_msz))
        print("%-32s %s %9.3f ms  %.4e ray-steps/s" % ("vert 2^%d op6 MIXED candidates" % lg, name, ms, nst / ms * 1e3), flush=True)
eng.set_option("queue", 0)

uc = "2"
eng.set_field(*rt.build_field(uc))
c = rt.scenario_constants(uc)
a = fan(uc, 20)
M = rt.max_size_for(uc, rt.DELTA_S, 91)
s_ray = torch.empty(M, 6, 1 << 20, dtype=torch.float64, device=dev)
for m in (6, 1):
    for mode, name in ((0, "per-lane st.global"), (1, "staged + TMA store")):
        eng.set_option("traj_store", mode)
        ms, steps = best(lambda: eng.trace(m, *a, rt.DELTA_S, M, c[0], c[5:9], full=True, out=dict(s_ray=s_ray)))
        print("fisheye 2^20 op%d FULL %s %9.3f ms  %.4e ray-steps/s  %.0f GB/s" % (m, name, ms, steps / ms * 1e3, steps * 48 / ms / 1e6),
              flush=True)
eng.set_option("traj_store", 0)
