"""GPU: batched two-point shooting search (rtb_shoot_coarse / rtb_shoot_refine through the C ABI) against
the CPU oracle, the fixtures solved through the real reference, and size-independent properties at the
4096-angle scale of BASELINE.json configs[4].  Parity-unpinned procedure (no reference function);
tolerances: indices / step counts / iteration counts bit-exact, angles 1e-10, miss and T 1e-9."""
import os

import numpy as np
import pytest

import raytracing_b200 as rt
from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _setup(uc):
    z, grd = rt.build_field(uc)
    c = rt.scenario_constants(uc)
    return z, grd, c[0], c[5:9]


@pytest.mark.parametrize("uc,m,div", [("1", 6, 3.0), ("1", 8, 2.0), ("3", 8, 2.0)])
def test_shooting_vs_reference_fixture(engine, uc, m, div):
    g = np.load(os.path.join(GOLDEN, "ref_shoot.npz"))
    k = lambda n: g["s%s_op%d/%s" % (uc, m, n)]
    z, grd, gamma, box = _setup(uc)
    engine.set_field(z, grd)
    step = float(k("step"))
    msz = rt.candidate_rows(uc, [step])
    co = engine.shoot_coarse(m, k("angles"), (-2.0, -2.0), [step], msz, (k("tx"), k("ty")), gamma, box, want_exit=True)
    assert np.abs(co.exit[:, 0] - k("coarse_exit")).max() <= 2e-9       # |x| up to 20
    assert np.array_equal(co.index[0], k("index"))
    r = rt.shooting_search(m, z, grd, uc, (k("tx"), k("ty")), k("angles"), steps=[step], engine=engine)
    assert np.array_equal(r["index"][0], k("index")) and np.array_equal(r["nsteps"][0], k("nsteps"))
    assert np.abs(r["theta"][0] - k("theta")).max() <= 1e-10
    assert np.abs(r["T"][0] - k("T")).max() <= 1e-9 and np.abs(r["miss"][0] - k("miss")).max() <= 1e-9
    assert np.abs(r["theta"][0] - k("true_angles")).max() < 3 * rt.GOLD_TOL


@pytest.mark.parametrize("uc,m", [("1", 6), ("3", 1), ("4", 11)])
def test_shooting_vs_oracle_multi_candidate(engine, uc, m):
    """Two DELTA_S candidates batched in one call == the oracle run per candidate."""
    from oracle import shoot_oracle as S
    z, grd, gamma, box = _setup(uc)
    steps = np.array([rt.SIGMA / 3.0, rt.SIGMA / 2.0])
    msz = rt.candidate_rows(uc, steps)
    lo_a = 2 * np.pi / 60 if uc == "1" else 0.0
    angles = np.linspace(lo_a, np.pi / 2, 40)
    true = np.array([0.35, 0.6, 1.15, 1.45]) + 0.0021   # away from the critical angle pi/4 (exit point discontinuous there)
    engine.set_field(z, grd)
    tx, ty, _ = rt.targets_from_angles(engine, m, uc, true, steps[0])
    r = rt.shooting_search(m, z, grd, uc, (tx, ty), angles, steps=steps, engine=engine)
    assert r["theta"].shape == (2, 4) and r["iterations"] >= 25
    tol_th = 1e-10 if m < 10 else 1e-7   # golden-in-golden (op11) resolves theta per step to 1.5e-8 only
    for ci in range(2):
        tr = S.Tracer(m, z, grd, steps[ci], int(msz[ci]), gamma, box, (-2.0, -2.0))
        o = S.shooting_search(tr, (tx, ty), angles)
        assert np.array_equal(r["index"][ci], o["index"])
        assert np.abs(r["theta"][ci] - o["theta"]).max() <= tol_th
        if m < 10:
            assert np.array_equal(r["nsteps"][ci], o["nsteps"])
            assert np.abs(r["T"][ci] - o["T"]).max() <= 1e-9 and np.abs(r["miss"][ci] - o["miss"]).max() <= 1e-9
    assert np.abs(r["theta"][0] - true).max() < 3 * rt.GOLD_TOL if m < 10 else True
    idx, dev = rt.sweep_two_point(r, tol_T=1e-3)
    assert dev[0] == 0.0 and np.isfinite(dev).all()


def test_shooting_4096_angles_properties(engine):
    """configs[4] scale on one axis: 4096 launch angles x 512 targets x 3 candidates.  Properties: the
    targets were generated from known launch angles -> recovered to golden's resolution with ~0 miss;
    device-pointer and host-pointer paths agree bit for bit; run-to-run deterministic."""
    import torch
    uc, m = "1", 6
    z, grd, gamma, box = _setup(uc)
    engine.set_field(z, grd)
    steps = rt.SIGMA / np.array([3.0, 2.0, 1.01])
    msz = rt.candidate_rows(uc, steps)
    A, Tn = 4096, 512
    angles = np.linspace(2 * np.pi / 60, np.pi / 2, A)
    true = np.linspace(0.2, 1.5, Tn) + 1.3e-4
    tx, ty, T_true = rt.targets_from_angles(engine, m, uc, true, steps[0])
    r = rt.shooting_search(m, z, grd, uc, (tx, ty), angles, steps=steps, engine=engine)
    # near the critical angle (pi/4: total reflection <-> refraction) the exit point jumps between the bottom
    # and the top edge, so a bracket there holds a discontinuity; everywhere else the map is smooth
    ok = np.abs(true - np.pi / 4) > 0.12
    assert np.abs(r["theta"][0] - true)[ok].max() < 3 * rt.GOLD_TOL
    assert r["miss"][0][ok].max() < 2e-6 and np.abs(r["T"][0] - T_true)[ok].max() < 2e-6
    assert (np.abs(angles[r["index"][0]] - true) <= (angles[1] - angles[0]))[ok].all()
    # coarser steps: still hit the target (miss << step), travel time near the fine answer
    assert (r["miss"][:2, ok] < 1e-5).all() and np.abs(r["T"][1] - r["T"][0])[ok].max() < 0.05
    assert r["coarse_miss"].shape == (3, Tn) and (r["miss"][0][ok] <= r["coarse_miss"][0][ok] + 1e-12).all()
    # device pointers
    d = lambda a: torch.as_tensor(a).cuda()
    co = engine.shoot_coarse(m, d(angles), (-2.0, -2.0), d(steps), d(msz), (d(tx), d(ty)), gamma, box)
    lo, hi = rt.brackets_from_index(co.index, d(angles))
    rf = engine.shoot_refine(m, lo, hi, (-2.0, -2.0), d(steps), d(msz), (d(tx), d(ty)), gamma, box)
    assert np.array_equal(rf.theta.cpu().numpy(), r["theta"]) and np.array_equal(co.index.cpu().numpy(), r["index"])
    assert np.array_equal(rf.result[0].cpu().numpy(), r["miss"]) and rf.iterations == r["iterations"]
    # one initialisation per group of pairs, two launches per golden pass of a group, three to finish
    assert rf.stats["ray_steps"] > 0 and rf.stats["launches"] >= 2 * rf.iterations + 4


def test_shoot_edge_cases(engine):
    from raytracing_b200 import _cabi
    z, grd, gamma, box = _setup("3")
    engine.set_field(z, grd)
    step = rt.SIGMA / 2
    msz = rt.candidate_rows("3", [step])
    ang = np.linspace(0.1, 1.5, 8)
    e = np.zeros(0)
    co = engine.shoot_coarse(6, ang, (-2.0, -2.0), [step], msz, (e, e), gamma, box)      # no targets
    assert co.miss2.shape == (1, 0)
    rf = engine.shoot_refine(6, np.zeros((1, 0)), np.zeros((1, 0)), (-2.0, -2.0), [step], msz, (e, e), gamma, box)
    assert rf.theta.shape == (1, 0) and rf.iterations == 0
    # a bracket already narrower than GOLD_TOL: zero iterations, theta = midpoint
    lo = np.array([[0.7]]); hi = lo + 1e-9
    rf = engine.shoot_refine(6, lo, hi, (-2.0, -2.0), [step], msz, (np.array([0.0]), np.array([1.0])), gamma, box)
    assert rf.iterations == 0 and rf.theta[0, 0] == (hi[0, 0] + lo[0, 0]) / 2 and rf.result[4, 0, 0] > 10
    with pytest.raises(_cabi.RtbError):
        engine.shoot_coarse(6, ang, (-2.0, -2.0), [step], [0], (np.array([0.0]), np.array([1.0])), gamma, box)


@pytest.mark.parametrize("uc,m,n_t", [("1", 6, 700), ("3", 8, 37)])
def test_speculated_passes_give_the_same_search(engine, uc, m, n_t):
    """rtb_shoot_refine with two golden passes per launch (both outcomes of a pass traced before it is decided,
    RTB_OPT_SHOOT_SPECULATE 2) == one pass per launch, bit for bit: theta, results, pass count; and fewer launches.
    The counted ray-steps are those of the probes the passes consumed, the same either way."""
    z, grd, gamma, box = _setup(uc)
    engine.set_field(z, grd)
    steps = np.array([rt.SIGMA / 3.0, rt.SIGMA / 1.3, rt.SIGMA / 2.0])
    msz = rt.candidate_rows(uc, steps)
    lo_a = 2 * np.pi / 60 if uc == "1" else 0.0
    angles = np.linspace(lo_a, np.pi / 2, 257)
    true = np.linspace(0.25, 1.5, n_t) + 1.7e-4
    tx, ty, _ = rt.targets_from_angles(engine, m, uc, true, steps[0])
    co = engine.shoot_coarse(m, angles, (-2.0, -2.0), steps, msz, (tx, ty), gamma, box)
    lo, hi = rt.brackets_from_index(co.index, angles)
    lo[1, ::5] = hi[1, ::5] - 1e-9                      # some pairs converged from the start, some brackets narrower
    hi[2, 1::3] = lo[2, 1::3] + 0.37 * (hi[2, 1::3] - lo[2, 1::3])
    out = []
    try:
        for mode in (1, 2):
            engine.set_option("shoot_speculate", mode)
            out.append(engine.shoot_refine(m, lo, hi, (-2.0, -2.0), steps, msz, (tx, ty), gamma, box))
    finally:
        engine.set_option("shoot_speculate", 0)
    a, b = out
    assert np.array_equal(a.theta, b.theta) and np.array_equal(a.result, b.result, equal_nan=True)
    assert a.iterations == b.iterations and a.stats["ray_steps"] == b.stats["ray_steps"]
    assert b.stats["launches"] < a.stats["launches"]
