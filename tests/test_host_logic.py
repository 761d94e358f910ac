"""CPU: host-side mirror of the reference's selection surface (constants, tables, rules)."""
import numpy as np
import pytest

import raytracing_b200 as rt
from raytracing_b200 import api, constants as K


def test_constants_match_reference_values():
    # SURVEY.md 8 [probe] values of RT_bench.py:59-97
    assert K.SIGMA == 0.05293304824724534
    assert K.DELTA == 0.017644349415748446
    assert K.DELTA_S == 0.002646652412362267
    assert K.GOLD_TOL == 1.4901161193847656e-08
    assert K.GOLD_RATIO == (np.sqrt(5) - 1) / 2


def test_scenario_table():
    c1 = rt.scenario_constants("1")
    assert c1[1] == 42 and len(c1[2]) == 43 and c1[4] == 80 and c1[5:9] == (-2, 20, -2, 4)
    c2 = rt.scenario_constants("fisheye")
    assert c2[1] == 1 and c2[4] == 10 * 2 * np.pi and tuple(c2[3]) == (1, 0)
    assert rt.scenario_constants("4")[0] == 3 and rt.scenario_constants("3")[0] == 1
    with pytest.raises(ValueError):
        rt.scenario_constants("5")


def test_max_size_and_methods():
    assert rt.max_size_for("1", K.DELTA_S, 91) == 30228
    assert rt.max_size_for("2", K.DELTA_S, 91) == 910
    assert rt.method_id("op7") == 7 and rt.method_id(11) == 11 and rt.method_id("3") == 3

    def op9():
        pass
    assert rt.method_id(op9) == 9
    for bad in (0, 12, "opx"):
        with pytest.raises(ValueError):
            rt.method_id(bad)


def test_calibrated_tables():
    assert rt.calibrated_step("1", 6)[0] == K.SIGMA / 2.55
    assert rt.calibrated_step("2", 1) == (2 * np.pi / 4587, 4587)
    assert rt.calibrated_step("4", 11)[0] == K.SIGMA / 2.74


def test_sweep_candidates_sizes():
    for uc, n in (("1", 200), ("2", 300), ("3", 200), ("4", 200)):
        d, o, _ = api.sweep_candidates(uc)
        assert len(d) == n == len(o)
    d, o, dev = api.sweep_candidates("2")
    assert d[0] == 303 and d[-1] == 4 and dev == 5


def test_find_index_rules():
    assert api.find_index_interface([0.1, 0.1, 0.3, 0.1], [0.5, 0.5, 0.9, 0.5], 0.2) == 1
    assert api.find_index_interface([0.1, 0.1], [0.5, 0.5], 0.2) is None          # nothing above threshold
    assert api.find_index_interface([0.1, 0.1, 0.3], [0.9, 0.5, 0.9], 0.2) is None  # max rule rejects every i
    assert api.find_index_fisheye([1, 2, 6, 3], 5) == 1
    assert api.find_index_fisheye([6, 1], 5) == -1                                  # reference quirk: i-1 = -1
    assert api.find_index_fisheye([1, 2], 5) is None
    assert api.find_index_vert([0.01, 0.02, 0.03, 0.06, 0.01], 0.05) == 2
    assert api.find_index_vert([0.06, 0.01, 0.01], 0.05) is None


def test_find_index_on_reference_metrics(golden_sweep):
    """Selection rules applied to the reference's own sweep metrics give the reference's index."""
    for key, uc in (("s2_op6", "2"), ("s2_op1", "2"), ("s3_op6", "3"), ("s3_op4", "3"), ("s4_op11", "4")):
        metric = list(golden_sweep[key + "/metric"])
        fn = api.find_index_fisheye if uc == "2" else api.find_index_vert
        idx = fn(metric, 5 if uc == "2" else 0.05)
        assert (-1 if idx is None else idx) == int(golden_sweep[key + "/index"])
    idx = api.find_index_interface(list(golden_sweep["s1_op6/mean"]), list(golden_sweep["s1_op6/max"]), 0.2)
    assert idx == int(golden_sweep["s1_op6/index"]) == 45   # divisor 2.55, the calibrated value RT:1424
    assert round(api.sweep_candidates("1")[0][idx], 2) == 2.55


def test_benchmark_statistic():
    data = np.array([1.0, 1.1, 0.9, 1.0, 50.0, 1.05])
    assert 50.0 not in api.remove_outliers_iqr(data)
    seq = iter(np.tile([1.0, 1.01, 0.99, 1.0], 1000))
    assert abs(api.benchmark_statistic(lambda: next(seq), trial=20) - 1.0) < 0.02


def test_benchmark_statistic_vs_reference_vectors():
    """ref_bench.npz (gen_golden.py bench): the reference's remove_outliers_iqr (RT:123-138) and its
    median-of-the-last-30 % / 0.5 % loop (RT:1516-1539) on recorded sample streams."""
    import os
    from conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "ref_bench.npz"))
    for case in range(3):
        samples = g["case%d/samples" % case]
        assert np.array_equal(api.remove_outliers_iqr(samples[0]), g["case%d/cleaned0" % case])
        stream = iter(samples.ravel())
        got = api.benchmark_statistic(lambda: next(stream), trial=samples.shape[1], replicas=1)
        assert got == float(g["case%d/final" % case])          # same statements, same numpy: bit-exact
        with pytest.raises(StopIteration):                       # ... after exactly as many rounds as the reference
            next(stream)


def test_make_fan_sharding_consistent():
    full = rt.make_fan("2", 1 << 12)
    part = rt.make_fan("2", 1 << 12, 1000, 1100)
    assert np.array_equal(full[0][1000:1100], part[0])
    th = rt.make_fan("3", 1 << 10)[0]
    assert th[0] == 0.0 and abs(th[-1] - np.pi / 2) < 1e-15


def test_field_builder_matches_reference_bytes(golden_field):
    """genZ + interpolacion restated in raytracing_b200.field produce the same spline the reference
    evaluates (checked through the oracle's bit-exact field test) and the documented sizes."""
    z, grd = rt.build_field("2")
    assert z.tck[2].size == 261121 and len(grd[0].tck[0]) == 515
    z, grd = rt.build_field("3")
    assert grd[0].tck[2].size == 397243 and (len(grd[0].tck[0]), len(grd[0].tck[1])) == (543, 741)


def test_pchip_knot_derivatives_match_scipy():
    """Host part of the wavefront product (RT_bench.py:1017-1019): derivative of the PCHIP interpolant at its knots."""
    from scipy.interpolate import PchipInterpolator
    import raytracing_b200 as rt
    rng = np.random.default_rng(3)
    for n in (2, 3, 4, 9, 31):
        x = np.sort(rng.uniform(-2, 1, n))
        y = np.cumsum(rng.normal(0, 1, n))
        if n == 9:
            y[4] = y[3]          # flat segment -> zero derivative rule
        d = rt.pchip_knot_derivatives(x, y)
        r = PchipInterpolator(x, y).derivative()(x)
        assert np.abs(d - r).max() <= 1e-13 * max(1.0, np.abs(r).max())


def test_cli_menu_mapping_and_results_wording(capsys):
    from raytracing_b200 import cli
    assert [cli.menu_to_method("1", k) for k in (1, 6, 9)] == [1, 6, 9]
    assert cli.menu_to_method("4", 1) == 10 and cli.menu_to_method("4", 2) == 11      # RT:1286-1290
    for bad in (("4", 3), ("2", 10), ("1", 0)):
        with pytest.raises(ValueError):
            cli.menu_to_method(*bad)
    cli.print_results("2", 1.5)
    cli.print_results("1", 0.25)
    cli.print_results("3", 0.01)
    out = capsys.readouterr().out
    assert "Closure error  1.5 %" in out and "Average ray error:  0.25 degrees" in out
    assert "Average ray Coefficient of Variation:  0.01" in out and out.count("RESULTS") == 3
    s = np.zeros((4, 6, 3)); s[-1, 0, 0] = 1.0
    assert cli.results_metric("2", s, None) == 0.0
    assert cli.results_metric("1", None, np.array([0.1, 0.3])) == pytest.approx(0.2)


@pytest.mark.parametrize("gamma", [0.6, 3.0])
def test_golden_tail_decision_rule(gamma):
    """The rule the anisotropic golden search uses for its passes >= 12 (step_fast.cuh golden_aniso_tail): the sign of
    t = e + kappa (e^2/2 + q^2/6) outside a band.  numpy restatement (tools/proto/golden_tail.py) against the evaluating
    search on random steps: no decision outside the band differs, and every pass the evaluating search would hand to the
    reference-order comparison lies inside the band."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tools", "proto"))
    import golden_tail
    differ, missed, tie_new, tie_old = golden_tail.run(K1=12, gamma=gamma, N=20000, verbose=False)
    assert differ == 0 and missed == 0
    assert tie_new <= tie_old + 0.05          # the wider band costs few extra out-of-line decisions
