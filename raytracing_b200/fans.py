"""Deterministic synthetic ray fans of the named benchmark sizes."""
import numpy as np


def make_fan(user_choice, R, lo=None, hi=None):
    """(theta0, x0, y0) float64 arrays for rays lo..hi-1 (default all) of an R-ray fan.

    fisheye: start (1,0), theta = linspace(pi/2-0.35, pi/2+0.35, R): every ray is a circle
    through (+-1,0) that stays inside the +-1.5 box, so no ray exits early and work is uniform.
    other scenarios: start (-2,-2) (RT_bench.py:261,277,812), theta = linspace(0|2pi/60, pi/2, R).
    Slices are computed from the global index so any sharding reproduces the same values.
    """
    from .constants import scenario_id
    uc = scenario_id(user_choice)
    lo = 0 if lo is None else int(lo)
    hi = R if hi is None else int(hi)
    return fan_at(uc, R, np.arange(lo, hi, dtype=np.float64))


def fan_at(user_choice, R, k):
    """Rays with global indices `k` (any order) of the R-ray fan of make_fan."""
    from .constants import scenario_id
    uc = scenario_id(user_choice)
    k = np.asarray(k, dtype=np.float64)
    if uc == "2":
        a, b = np.pi / 2 - 0.35, np.pi / 2 + 0.35
        x0, y0 = 1.0, 0.0
    else:
        a, b = (2 * (np.pi / 60) if uc == "1" else 0.0), np.pi / 2
        x0, y0 = -2.0, -2.0
    theta = a + (b - a) * (k / max(R - 1, 1)) if R > 1 else np.full(k.size, a)
    return theta, np.full(k.size, x0), np.full(k.size, y0)
