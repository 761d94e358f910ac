"""Dev script: kernel timing of every method on the fisheye 2^20 fan (FINAL) + vert 2^18 fan."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raytracing_b200 as rt
eng = rt.get_engine(0)
tag = os.environ.get("RTB_LIB_PATH", "default")
for uc, R, methods in (("2", 1 << 20, (6, 1, 2, 3, 4, 7, 8, 9, 5)), ("3", 1 << 18, (6, 1, 3)), ("1", 1 << 18, (6, 1, 4)), ("4", 1 << 16, (10, 11))):
    z, grd = rt.build_field(uc); eng.set_field(z, grd)
    c = rt.scenario_constants(uc); gamma, box = c[0], c[5:9]
    th, x0, y0 = (torch.from_numpy(a).cuda() for a in rt.make_fan(uc, R))
    M = rt.max_size_for(uc, rt.DELTA_S, 91)
    for m in methods:
        for ex in (False, True):
            if ex and m not in (6, 1, 10): continue
            for it in range(2):
                r = eng.trace(m, th, x0, y0, rt.DELTA_S, M, gamma, box, full=False, exact=ex, interface=False)
            print("%s s%s op%d exact=%d: %.2f ms %.3e steps/s" % (tag[-12:], uc, m, ex, r.stats["kernel_ms"], r.stats["ray_steps"] / r.stats["kernel_ms"] * 1e3), flush=True)
