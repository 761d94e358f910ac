"""CPU: the two-point shooting oracle (oracle/shoot_oracle.py over the C port) against problems solved
through the REAL reference (tests/golden/ref_shoot.npz: RT.trazar rays + RT.golden refinement), and the
host-side pieces of raytracing_b200/shooting.py.  The procedure itself is parity-unpinned (no reference
function); its parts are pinned here."""
import os

import numpy as np
import pytest

import raytracing_b200 as rt
from conftest import GOLDEN
from oracle import shoot_oracle as S

CASES = [("1", 6, 3.0), ("1", 8, 2.0), ("3", 8, 2.0)]


@pytest.fixture(scope="module")
def golden_shoot():
    return np.load(os.path.join(GOLDEN, "ref_shoot.npz"))


@pytest.mark.parametrize("uc,m,div", CASES)
def test_oracle_matches_reference_procedure(golden_shoot, uc, m, div):
    g = lambda k: golden_shoot["s%s_op%d/%s" % (uc, m, k)]
    z, grd = rt.build_field(uc)
    c = rt.scenario_constants(uc)
    step = rt.SIGMA / div
    assert step == float(g("step"))
    M = int(rt.candidate_rows(uc, [step])[0])
    tr = S.Tracer(m, z, grd, step, M, c[0], c[5:9], (-2.0, -2.0))
    ex, ns = tr(g("angles"))
    assert np.array_equal(ns, g("coarse_nsteps"))
    assert np.abs(ex - g("coarse_exit")).max() <= 1e-10 * 20
    r = S.shooting_search(tr, (g("tx"), g("ty")), g("angles"))
    assert np.array_equal(r["index"], g("index")) and np.array_equal(r["nsteps"], g("nsteps"))
    assert np.abs(r["theta"] - g("theta")).max() <= 1e-10
    assert np.abs(r["T"] - g("T")).max() <= 1e-9 and np.abs(r["miss"] - g("miss")).max() <= 1e-9
    # the search recovers the launch angles that generated the targets to golden's resolution
    assert np.abs(r["theta"] - g("true_angles")).max() < 3 * rt.GOLD_TOL


def test_exit_points_definition():
    """Hand-made trajectory: segment (1,1)->(3,2) leaves the box x<=2 at t=0.5."""
    s = np.zeros((3, 6, 2))
    s[0, 0:2, :] = 0.0
    s[1, 0, :], s[1, 1, :], s[1, 4, :] = 1.0, 1.0, 10.0
    s[2, 0, 0], s[2, 1, 0], s[2, 4, 0] = 3.0, 2.0, 14.0     # ray 0 exits through x = 2
    d = np.array([[0, 0], [0, 0], [2, 1]], dtype=float)      # ray 1 stopped inside (max_size) at row 1
    e = S.exit_points(s, d, (-5.0, 2.0, -5.0, 5.0))
    assert np.allclose(e[:, 0], [2.0, 1.5, 12.0]) and np.allclose(e[:, 1], [1.0, 1.0, 10.0])
    # corner: both limits violated -> the first one crossed wins
    s[2, 0, 0], s[2, 1, 0] = 3.0, 9.0                        # y limit 5 crossed at t=0.5 as well (x at 0.5)
    e = S.exit_points(s, d, (-5.0, 2.0, -5.0, 5.0))
    assert np.allclose(e[:2, 0], [2.0, 5.0])


def test_brackets_and_selection():
    ang = np.linspace(0.0, 1.0, 11)
    idx = np.array([[0, 5, 10]])
    lo, hi = rt.brackets_from_index(idx, ang)
    assert np.array_equal(lo, [[0.0, ang[4], ang[9]]]) and np.array_equal(hi, [[ang[1], ang[6], 1.0]])
    res = dict(T=np.array([[1.0, 2.0], [1.0002, 2.0001], [1.01, 2.0]]), miss=np.zeros((3, 2)), steps=np.array([.1, .2, .3]))
    i, dev = rt.sweep_two_point(res, tol_T=1e-3)
    assert i == 1 and dev[0] == 0.0 and abs(dev[2] - 0.01) < 1e-12
    assert rt.sweep_two_point(res, tol_T=1.0)[0] is None       # all pass -> None, like find_index
    assert rt.candidate_rows("1", [rt.DELTA_S])[0] == 30228 and rt.default_source("2") == (1.0, 0.0)
