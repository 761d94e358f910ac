"""oracle/fit_oracle.py pinned against the reference's own calls (np.gradient RT:449, RectBivariateSpline RT:455-457)."""
import numpy as np
import pytest
from scipy import interpolate

import raytracing_b200 as rt
from raytracing_b200 import constants as K
from oracle import fit_oracle as F


def _scenario_grid(uc, coarsen=1):
    lim = K.constants(uc)[5:9]
    x, y, _X, _Y, Z = rt.genZ(lim[0], lim[1], lim[2], lim[3], K.SCENARIO_FUNCS[uc], delta=K.DELTA * coarsen)
    return x, y, Z


def _check(x, y, Z, delta, z_exact=False):
    z, (ga, gb) = F.interpolacion(x, y, Z, delta)
    G = np.gradient(Z, delta, edge_order=2)
    g0, g1 = F.gradient(Z, delta)
    assert np.array_equal(g0, G[0]) and np.array_equal(g1, G[1])                   # bit-exact
    zs = interpolate.RectBivariateSpline(y, x, Z, kx=1, ky=1)
    assert np.array_equal(z[0], zs.tck[0]) and np.array_equal(z[1], zs.tck[1])
    # FITPACK's Givens sweep leaves the k=1 coefficients within an ulp of the samples on some grids and exactly on
    # them on the reference's own (DELTA) grids; the restatement returns the samples
    assert np.abs(z[2] - zs.tck[2]).max() <= (0 if z_exact else 4e-16 * np.abs(Z).max())
    for mine, g in ((ga, G[0]), (gb, G[1])):
        s = interpolate.RectBivariateSpline(y, x, g, kx=3, ky=3)
        assert np.array_equal(mine[0], s.tck[0]) and np.array_equal(mine[1], s.tck[1])
        scale = max(np.abs(s.tck[2]).max(), 1e-300)
        assert np.abs(mine[2] - s.tck[2]).max() <= 2e-13 * scale


@pytest.mark.parametrize("uc", ["1", "2", "3"])
def test_fit_matches_scipy_on_scenario_grids(uc):
    x, y, Z = _scenario_grid(uc, coarsen=4)      # the scenario fields on a 4x coarser grid: seconds in pure Python
    _check(x, y, Z, K.DELTA * 4)


def test_fit_matches_scipy_on_the_full_fisheye_grid():
    x, y, Z = _scenario_grid("2")
    assert Z.shape == (511, 511)
    _check(x, y, Z, K.DELTA, z_exact=True)


def test_fit_matches_scipy_on_ragged_grids():
    rng = np.random.default_rng(5)
    for my, mx in ((4, 4), (5, 9), (17, 6), (40, 31)):
        y = np.cumsum(rng.uniform(0.2, 1.0, my))
        x = np.cumsum(rng.uniform(0.01, 0.03, mx)) - 3
        _check(x, y, rng.standard_normal((my, mx)), 0.0173)


def test_reference_quirk_gradient_uses_DELTA_not_grid_spacing():
    """RT:449 passes DELTA although linspace's spacing differs slightly; the restatement keeps it."""
    x, y, Z = _scenario_grid("2", coarsen=4)
    assert abs((x[1] - x[0]) - K.DELTA * 4) > 1e-6
    g0, _ = F.gradient(Z, K.DELTA * 4)
    assert np.array_equal(g0[1:-1], (Z[2:] - Z[:-2]) / (2. * K.DELTA * 4))
