"""rtb_sweep_noalloc (include/rtb.h): the DELTA_S sweep on the library's fixed scratch array, for callers that cannot
allocate.  It has no counterpart in the reference; it is checked against rtb_sweep, the entry point it is a variant of,
on the same inputs -- through the C ABI, twice in a row (the scratch array keeps the previous call's contents), with a
different workload in between.  Collected last (file name) so that a failure here hides no other test."""
import numpy as np
import pytest

import raytracing_b200 as rt
from raytracing_b200 import _cabi
from _util import assert_close, scenario_setup

pytestmark = pytest.mark.gpu

MAX_RAYS, MAX_CAND = 131072, 256          # RTB_NOALLOC_MAX_RAYS, RTB_NOALLOC_MAX_CANDIDATES of include/rtb.h


def _case(uc, R, n_cand):
    s = scenario_setup(uc)
    divisors, options, _ = rt.sweep_candidates(uc)
    sel = np.linspace(0, len(options) - 1, n_cand).astype(int)
    options, divisors = options[sel], divisors[sel]
    msz = np.array([rt.max_size_for(uc, st, d + 1) for st, d in zip(options, divisors)], dtype=np.int64)
    th, x0, y0 = (np.ascontiguousarray(a) for a in rt.make_fan(uc, R))
    metric = "closure" if uc == "2" else "px_cv"
    return s, th, x0, y0, options, msz, metric, (1 if metric == "px_cv" else 0)


def _run_noalloc(engine, m, s, th, x0, y0, options, msz, metric, skip, **kw):
    import torch
    dev = [torch.from_numpy(a).cuda() for a in (th, x0, y0, options, msz)]
    Cn = len(options)
    # outputs are the caller's: filled with garbage so that stale values cannot pass for results
    sm = torch.full((Cn,), -7.0, dtype=torch.float64, device="cuda")
    mx = torch.full((Cn,), -7.0, dtype=torch.float64, device="cuda")
    cnt = torch.full((Cn,), -7, dtype=torch.int64, device="cuda")
    stats = engine.sweep_noalloc(m, *dev, s["gamma"], s["box"], metric, sm, mx, cnt, skip_lo=skip, skip_hi=skip, **kw)
    torch.cuda.synchronize()
    return sm.cpu().numpy(), mx.cpu().numpy(), cnt.cpu().numpy(), stats


def _check(engine, uc, m, R, n_cand):
    s, th, x0, y0, options, msz, metric, skip = _case(uc, R, n_cand)
    engine.set_field(s["z"], s["grd"])
    want = engine.sweep(m, th, x0, y0, options, msz, s["gamma"], s["box"], metric, skip_lo=skip, skip_hi=skip)
    for call in range(2):                                      # two calls in a row on the shared array
        got = _run_noalloc(engine, m, s, th, x0, y0, options, msz, metric, skip)
        assert np.array_equal(got[2], want[2]), "count (call %d)" % call
        assert_close(got[0], want[0], 1e-10, "sum (call %d)" % call)     # floats: 1e-10 (tests/_util.py policy);
        assert_close(got[1], want[1], 1e-10, "max (call %d)" % call)     # same kernels, fixed reduction order
        assert got[3]["ray_steps"] == want[3]["ray_steps"]
        assert got[3]["launches"] == 2


def test_noalloc_sweep_equals_rtb_sweep(engine):
    # vert, op6: 16 candidates x 4096 rays = 65536 rays in one launch (warp-staged cell records), p_x CV metric
    _check(engine, "3", 6, 4096, 16)
    # a different scenario and metric in between: fisheye closure reads the last rows, which start from a cleared
    # array whatever the previous call left there
    _check(engine, "2", 6, 512, 8)
    # aniso, op11 (golden-section stepper), and the first workload again after the others
    _check(engine, "4", 11, 256, 6)
    _check(engine, "3", 6, 4096, 16)


def test_noalloc_sweep_limits(engine):
    import torch
    s, th, x0, y0, options, msz, metric, skip = _case("3", 1024, 4)
    engine.set_field(s["z"], s["grd"])
    big = rt.make_fan("3", MAX_RAYS // 4 + 1024)
    with pytest.raises(Exception, match="RTB_NOALLOC_MAX_RAYS"):              # C * R beyond the array
        _run_noalloc(engine, 6, s, *(np.ascontiguousarray(a) for a in big), options, msz, metric, skip)
    many = np.resize(options, MAX_CAND + 1)
    with pytest.raises(Exception, match="RTB_NOALLOC_MAX_CANDIDATES"):
        _run_noalloc(engine, 6, s, th[:8], x0[:8], y0[:8], many, np.resize(msz, MAX_CAND + 1), metric, skip)
    with pytest.raises(Exception, match="RTB_SWEEP_CLOSURE or RTB_SWEEP_PX_CV"):
        _run_noalloc(engine, 6, s, th, x0, y0, options, msz, "snell", 0)
    with pytest.raises(Exception, match="device pointers only"):
        _run_noalloc(engine, 6, s, th, x0, y0, options, msz, metric, skip, ptr_kind=_cabi.RTB_PTR_HOST)
    # the largest problem the array holds: 131072 rays
    full = [np.ascontiguousarray(a) for a in rt.make_fan("3", MAX_RAYS // 4)]
    want = engine.sweep(6, *full, options, msz, s["gamma"], s["box"], metric, skip_lo=skip, skip_hi=skip)
    got = _run_noalloc(engine, 6, s, *full, options, msz, metric, skip)
    assert np.array_equal(got[2], want[2])
    assert_close(got[0], want[0], 1e-10, "sum at the limit")
    assert_close(got[1], want[1], 1e-10, "max at the limit")
    torch.cuda.synchronize()
