"""CPU: pin the oracle (oracle/rt_oracle.c) against fixtures generated from the UNMODIFIED reference
(tests/golden/gen_golden.py).  This is what makes the oracle a trustworthy checker for the GPU tests."""
import numpy as np
import pytest

from _util import SCEN_METHODS, assert_close, check_steps, scenario_setup, tolerances
from oracle import rt_oracle as O

CASES = [(uc, m) for uc, ms in SCEN_METHODS.items() for m in ms]


@pytest.mark.parametrize("uc", ["1", "2", "3"])
def test_field_bit_exact(golden_field, uc):
    """FITPACK bispev restatement == scipy (through RT.n_gradient) to the last bit, incl. knots and clamping."""
    s = scenario_setup(uc)
    x, y = golden_field["s%s/x" % uc], golden_field["s%s/y" % uc]
    n, g = O.n_gradient(np.stack([x, y], 1), s["grd"], s["z"])
    assert np.array_equal(n, golden_field["s%s/n" % uc])
    assert np.array_equal(g, golden_field["s%s/g" % uc])


@pytest.mark.parametrize("uc,m", CASES)
def test_default_fan_vs_reference(golden_traj, sensitivity, uc, m):
    s = scenario_setup(uc)
    key = "default_s%s_op%d/" % (uc, m)
    tol = tolerances(sensitivity, uc, m)
    assert int(golden_traj[key + "M"]) == s["M"]
    assert np.array_equal(golden_traj[key + "theta0"], s["theta0"])
    o = O.trace(m, s["z"], s["grd"], s["step"], s["M"], s["gamma"], s["box"], uc == "1", s["theta0"], s["x0"], s["y0"],
                nthreads=8)
    same = check_steps(o["d_ray"][2], golden_traj[key + "d_ray"][2], tol, key)
    stride = int(golden_traj[key + "stride"])
    assert_close(o["s_ray"][::stride][:, :, same], golden_traj[key + "rows"][:, :, same], tol["all"], key + "rows")
    assert_close(o["final"][:, same], golden_traj[key + "last_row"][:, same], tol["all"], key + "last_row")
    assert_close(o["d_ray"][:2, same], golden_traj[key + "d_ray"][:2, same], tol["d_ray"], key + "d_ray")
    assert_close(o["errors"][same], golden_traj[key + "errors"][same], tol["err"], key + "errors")


@pytest.mark.parametrize("uc", ["2", "3", "4"])
def test_synthetic_fan_vs_reference(golden_traj, sensitivity, uc):
    import raytracing_b200 as rt
    s = scenario_setup(uc)
    R, idx = int(golden_traj["synth_s%s/R" % uc]), golden_traj["synth_s%s/idx" % uc]
    th = np.concatenate([rt.make_fan(uc, R, i, i + 1)[0] for i in idx])
    x0, y0 = rt.make_fan(uc, R, 0, len(idx))[1:]
    for m in SCEN_METHODS[uc]:
        key = "synth_s%s_op%d/" % (uc, m)
        tol = tolerances(sensitivity, uc, m)
        assert np.array_equal(golden_traj[key + "theta0"], th)
        o = O.trace(m, s["z"], s["grd"], s["step"], s["M"], s["gamma"], s["box"], False, th, x0, y0, nthreads=8)
        same = check_steps(o["d_ray"][2], golden_traj[key + "d_ray"][2], tol, key)
        stride = int(golden_traj[key + "stride"])
        assert_close(o["s_ray"][::stride][:, :, same], golden_traj[key + "rows"][:, :, same], tol["all"], key)


def test_fingerprints_baseline_md(golden_traj):
    """BASELINE.md survey fingerprints: total ray-steps of the default fans."""
    want = {"default_s1_op1": 158051, "default_s1_op6": 159561, "default_s1_op9": 159544, "default_s2_op6": 909,
            "default_s3_op6": 56567, "default_s3_op7": 56571, "default_s4_op10": 60457, "default_s4_op11": 60457}
    for k, v in want.items():
        assert int(golden_traj[k + "/d_ray"][2].sum()) == v


def test_rays_leaving_on_their_first_steps_vs_reference():
    """ref_edge.npz (gen_golden.py edge): op7 takes its two bootstrap steps before the loop that holds the box test
    (RT:838-866), so a ray that is outside after step 1 or 2 still reaches row 3; op6/op1 break at once (RT:878)."""
    import os
    from conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "ref_edge.npz"))
    s = scenario_setup("3")
    theta, pos_x = g["theta"], g["pos_x"]
    for m in (7, 6, 1):
        o = O.trace(m, s["z"], s["grd"], s["step"], s["M"], s["gamma"], s["box"], False, theta, pos_x, np.full(theta.size, -2.0))
        d = g["s3_op%d/d_ray" % m]
        assert np.array_equal(o["d_ray"][2], d[2]), m
        assert_close(o["d_ray"][:2], d[:2], 1e-10, "edge d_ray op%d" % m)
        assert_close(o["s_ray"][:8], g["s3_op%d/rows" % m], 1e-10, "edge rows op%d" % m)
    assert list(g["s3_op7/d_ray"][2][:4]) == [3, 3, 3, 3] and list(g["s3_op6/d_ray"][2][:4]) == [1, 2, 1, 2]
