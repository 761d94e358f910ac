"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the two-point shooting search of
raytracing_b200/shooting.py (BASELINE.json configs[4]).

PARITY UNPINNED: the reference has no two-point search (RT_bench.py imports scipy.optimize.minimize
at line 48 and never calls it; README.md:6 "shooting method" = forward fan).  The procedure is
assembled from reference parts: every ray is a trazar ray (RT_bench.py:766-948, here through the C
port rt_oracle.c or, in tests/golden/gen_golden.py, through the real RT.trazar), the refinement is a
literal restatement of golden (RT_bench.py:175-199).  The only new definitions are the exit point
(crossing of the last segment with the box, from rows i-1 and i of s_ray) and the miss distance.
"""
import numpy as np

from . import rt_oracle as O

GOLD_RATIO = (np.sqrt(5) - 1) / 2   # RT_bench.py:65
GOLD_TOL = np.sqrt(np.finfo(float).eps)  # RT_bench.py:66


def exit_points(s_ray, d_ray, box):
    """(ex, ey, eT) per ray from a trazar result: the last segment (rows i-1 -> i, i = d_ray[2]) cut at
    the first box limit it violates (strict tests of RT_bench.py:878); a ray still inside reports row i."""
    xi, xs, yi, ys = box
    R = d_ray.shape[1]
    out = np.zeros((3, R))
    for k in range(R):
        i = int(d_ray[2, k])
        x1, y1, T1 = s_ray[i, 0, k], s_ray[i, 1, k], s_ray[i, 4, k]
        if i == 0:
            out[:, k] = x1, y1, T1
            continue
        x0, y0, T0 = s_ray[i - 1, 0, k], s_ray[i - 1, 1, k], s_ray[i - 1, 4, k]
        t = 1.0
        if x1 > xs or x1 < xi or y1 > ys or y1 < yi:
            if x1 > xs:
                t = min(t, (xs - x0) / (x1 - x0))
            if x1 < xi:
                t = min(t, (xi - x0) / (x1 - x0))
            if y1 > ys:
                t = min(t, (ys - y0) / (y1 - y0))
            if y1 < yi:
                t = min(t, (yi - y0) / (y1 - y0))
            t = max(t, 0.0)
        out[:, k] = x0 + t * (x1 - x0), y0 + t * (y1 - y0), T0 + t * (T1 - T0)
    return out


def golden(func, a, b, tol=GOLD_TOL):
    """RT_bench.py:175-199, literally."""
    c = b - (b - a) * GOLD_RATIO
    d = a + (b - a) * GOLD_RATIO
    n = 0
    while abs(c - d) > tol:
        if func(c) < func(d):
            b = d
        else:
            a = c
        c = b - (b - a) * GOLD_RATIO
        d = a + (b - a) * GOLD_RATIO
        n += 1
    return (b + a) / 2, n


class Tracer:
    """exit point + step count of rays from a common source, one DELTA_S candidate."""

    def __init__(self, method, z, grd, step, max_size, gamma, box, src, trace_fn=None):
        self.a = (method, O._as_host(z), (O._as_host(grd[0]), O._as_host(grd[1])), step, max_size, gamma, box)
        self.src, self.box = src, box
        self.trace_fn = trace_fn   # optional: the REAL reference (gen_golden.py) instead of the C port

    def __call__(self, thetas):
        th = np.atleast_1d(np.asarray(thetas, dtype=np.float64))
        if self.trace_fn is not None:
            s_ray, d_ray = self.trace_fn(th)
        else:
            m, z, grd, step, M, gamma, box = self.a
            r = O.trace(m, z, grd, step, M, gamma, box, False, th, self.src[0], self.src[1], full=True, nthreads=8)
            s_ray, d_ray = r["s_ray"], r["d_ray"]
        return exit_points(s_ray, d_ray, self.box), d_ray[2].astype(np.int64)


def shooting_search(tracer, targets, angles):
    """One candidate.  Returns dict of (Tn,) arrays: theta, miss, T, exit_x, exit_y, nsteps, index, coarse_miss."""
    tx, ty = (np.asarray(t, dtype=np.float64) for t in targets)
    angles = np.asarray(angles, dtype=np.float64)
    ex, _ = tracer(angles)
    Tn, A = tx.size, angles.size
    out = {k: np.zeros(Tn) for k in ("theta", "miss", "T", "exit_x", "exit_y", "coarse_miss")}
    out["nsteps"] = np.zeros(Tn, dtype=np.int64)
    out["index"] = np.zeros(Tn, dtype=np.int64)
    out["iterations"] = 0
    for t in range(Tn):
        dx, dy = ex[0] - tx[t], ex[1] - ty[t]
        m2 = dx * dx + dy * dy
        j = int(np.argmin(m2))
        lo, hi = angles[max(j - 1, 0)], angles[min(j + 1, A - 1)]

        def f(theta):
            e, _ = tracer([theta])
            ddx, ddy = e[0, 0] - tx[t], e[1, 0] - ty[t]
            return ddx * ddx + ddy * ddy
        theta, n = golden(f, lo, hi)
        e, ns = tracer([theta])
        ddx, ddy = e[0, 0] - tx[t], e[1, 0] - ty[t]
        out["theta"][t], out["miss"][t] = theta, np.sqrt(ddx * ddx + ddy * ddy)
        out["T"][t], out["exit_x"][t], out["exit_y"][t] = e[2, 0], e[0, 0], e[1, 0]
        out["nsteps"][t], out["index"][t], out["coarse_miss"][t] = ns[0], j, np.sqrt(m2[j])
        out["iterations"] = max(out["iterations"], n)
    return out
