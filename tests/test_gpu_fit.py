"""GPU: interpolacion() on the device (rtb_gradient / rtb_fit_field / rtb_get_field) against the
reference's own calls -- np.gradient (RT_bench.py:449) and scipy RectBivariateSpline (RT:455-457) -- and against the
CPU restatement oracle/fit_oracle.py; then the traced rays of a device-fitted field against the golden fixtures."""
import numpy as np
import pytest
from scipy import interpolate

import raytracing_b200 as rt
from raytracing_b200 import constants as K
from raytracing_b200._cabi import RtbError
from oracle import fit_oracle as F
from _util import assert_close, tolerances

pytestmark = pytest.mark.gpu


def _grid(uc):
    lim = K.constants(uc)[5:9]
    x, y, _X, _Y, Z = rt.genZ(lim[0], lim[1], lim[2], lim[3], K.SCENARIO_FUNCS[uc])
    return x, y, Z


@pytest.mark.parametrize("uc", ["1", "2", "3"])
def test_gradient_bit_exact(uc):
    x, y, Z = _grid(uc)
    want = np.gradient(Z, K.DELTA, edge_order=2)
    got = rt.get_engine().gradient(Z, K.DELTA)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


def test_gradient_ragged_and_errors():
    eng = rt.get_engine()
    rng = np.random.default_rng(11)
    for shape in ((3, 3), (3, 17), (129, 5), (257, 300)):
        Z = rng.standard_normal(shape) * 10.0 ** rng.integers(-3, 4)
        want = np.gradient(Z, 0.0173, edge_order=2)
        got = eng.gradient(Z, 0.0173)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    with pytest.raises(RtbError):      # numpy: "Shape of array too small to calculate a numerical gradient"
        eng.gradient(np.zeros((2, 5)), 0.1)


@pytest.mark.parametrize("uc", ["1", "2", "3"])
def test_fit_coefficients_vs_scipy(uc):
    """Full-size scenario grids: knots identical, bilinear coefficients identical, bicubic coefficients within a few
    ulp of max|coefficient| of FITPACK's (same collocation system, different elimination order)."""
    x, y, Z = _grid(uc)
    zs, gs = rt.build_field(uc)                       # scipy, as the reference builds them
    z, grd = rt.interpolacion_gpu(x, y, Z)
    for mine, ref in ((z, zs), (grd[0], gs[0]), (grd[1], gs[1])):
        tu, tv, c = mine.tck
        assert mine.degrees == tuple(ref.degrees)
        assert np.array_equal(tu, ref.tck[0]) and np.array_equal(tv, ref.tck[1])
        if mine is z:
            assert np.array_equal(c, ref.tck[2])
        else:
            assert np.abs(c - ref.tck[2]).max() <= 2e-13 * np.abs(ref.tck[2]).max()


def test_fit_matches_cpu_restatement():
    x, y, Z = _grid("2")
    z, grd = rt.interpolacion_gpu(x, y, Z)
    _zo, (a, b) = F.interpolacion(x, y, Z, K.DELTA)
    for mine, want in ((grd[0], a), (grd[1], b)):
        c = mine.tck[2]
        assert np.abs(c - want[2]).max() <= 1e-14 * np.abs(want[2]).max()     # same algorithm, FMA contraction only


def test_fit_nonuniform_grid_and_evaluation():
    """Ragged sample spacing: no uniform run, so the general evaluator serves the fitted field."""
    rng = np.random.default_rng(3)
    y = np.cumsum(rng.uniform(0.05, 0.2, 61)) - 4
    x = np.cumsum(rng.uniform(0.05, 0.2, 83)) - 5
    X, Y = np.meshgrid(x, y)
    Z = 1.5 + 0.3 * np.sin(X) * np.cos(0.7 * Y)
    eng = rt.get_engine()
    z, grd = eng.fit_field(x, y, Z, 0.1)
    G = np.gradient(Z, 0.1, edge_order=2)
    zs = interpolate.RectBivariateSpline(y, x, Z, kx=1, ky=1)
    g0 = interpolate.RectBivariateSpline(y, x, G[0], kx=3, ky=3)
    g1 = interpolate.RectBivariateSpline(y, x, G[1], kx=3, ky=3)
    assert np.abs(grd[0].tck[2] - g0.tck[2]).max() <= 2e-13 * np.abs(g0.tck[2]).max()
    px = rng.uniform(x[0], x[-1], 4000)
    py = rng.uniform(y[0], y[-1], 4000)
    eng.set_field(z, grd)
    for exact in (False, True):
        n, gx, gy = eng.n_gradient(px, py, exact=exact)
        assert_close(n, zs.ev(py, px), 1e-12, "n")
        assert_close(gy, g0.ev(py, px), 1e-11, "dn/dy")
        assert_close(gx, g1.ev(py, px), 1e-11, "dn/dx")


@pytest.mark.parametrize("uc,m", [("1", 6), ("2", 6), ("3", 8), ("4", 11), ("2", 1)])
def test_rays_of_device_fitted_field_match_reference(golden_traj, sensitivity, uc, m):
    """The whole chain genZ -> rtb_fit_field -> rtb_trace against the reference's trajectories."""
    z, grd = rt.build_field_gpu(uc)
    s_ray, d_ray, _t, errors = rt.trazar(m, z, grd, False, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1, uc)
    key = "default_s%s_op%d/" % (uc, m)
    if key + "d_ray" not in golden_traj:
        pytest.skip("no fixture for " + key)
    tol = tolerances(sensitivity, uc, m)
    assert np.array_equal(d_ray[2], golden_traj[key + "d_ray"][2])
    assert_close(d_ray[:2], golden_traj[key + "d_ray"][:2], tol["d_ray"], "d_ray")
    if uc == "1":
        assert_close(errors, golden_traj[key + "errors"], tol["err"], "errors")


def test_device_pointers_and_refit_after_replacement():
    torch = pytest.importorskip("torch")
    eng = rt.get_engine()
    x, y, Z = _grid("2")
    dev = torch.device("cuda", eng.device)
    z, grd = eng.fit_field(torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev), torch.from_numpy(Z).to(dev),
                           K.DELTA)
    c_dev = grd[1].tck[2].copy()
    zs, gs = rt.build_field("2")
    assert np.abs(c_dev - gs[1].tck[2]).max() <= 2e-13 * np.abs(gs[1].tck[2]).max()
    # another field takes the context, then the handles are used again: refit from the kept samples
    rt.trazar(6, *rt.build_field("3"), False, rt.DELTA_S, 91, "3")
    a = rt.trazar(6, z, grd, False, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1, "2")
    b = rt.trazar(6, zs, gs, False, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1, "2")
    assert np.array_equal(a[1][2], b[1][2])
    assert_close(a[0], b[0], 1e-10, "s_ray device fit vs scipy fit")


def test_fit_errors():
    eng = rt.get_engine()
    x, y = np.linspace(0, 1, 8), np.linspace(0, 1, 6)
    with pytest.raises(ValueError):
        eng.fit_field(x, y, np.zeros((8, 6)), 0.1)                 # transposed samples
    with pytest.raises(RtbError):
        eng.fit_field(x[:3], y, np.zeros((6, 3)), 0.1)             # fewer samples than a cubic needs
    xb = x.copy()
    xb[3] = xb[2]
    with pytest.raises(RtbError):
        eng.fit_field(xb, y, np.zeros((6, 8)), 0.1)                # scipy: "x must be strictly increasing"
    z, grd = eng.fit_field(x, y, np.ones((6, 8)), 0.1)             # a failed fit does not poison the context
    n, gx, gy = eng.n_gradient([0.3], [0.4])
    assert n[0] == 1.0 and gx[0] == 0.0 and gy[0] == 0.0
