"""GPU parity tests added in round 2: dense fans on the paths the benchmarks time (warp-staged records, ray queue)
against the oracle and the reference-generated fixtures; ray queue / TMA trajectory stores against the static grid /
per-lane stores bit for bit; rtb_sweep; the edge-of-box fixture; the strided Snell metric; the non-finite flag."""
import os

import numpy as np
import pytest

import raytracing_b200 as rt
from _util import SCEN_METHODS, assert_close, check_steps, scenario_setup, tolerances
from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _dev(*arrays):
    import torch
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrays]


def _oracle_final(s, uc, m, th, x0, y0):
    from oracle import rt_oracle as O
    return O.trace(m, s["z"], s["grd"], s["step"], s["M"], s["gamma"], s["box"], False, th, x0, y0, full=False, nthreads=16)


def _measured_tolerance(s, uc, m, th, x0, y0, base, tol):
    """tests/_util.py's policy applied to THESE rays: how far the reference algorithm's own outputs move when every
    sin/cos result is replaced by a neighbouring double (oracle noise probe).  Needed where the reference is
    ill-conditioned and the dense fans reach places the 31-ray default fans of sensitivity.json do not: curvature_t
    divides a cancelling difference by a curvature just above GOLD_TOL (RT:361-363) for rays ever closer to the
    vertical, finite_diff amplifies noise.  Returns (tolerance dict, rays whose step count the noise itself moves)."""
    from oracle import rt_oracle as O
    from _util import SAFETY
    dev_all = dev_d = 0.0
    unstable = np.zeros(th.size, dtype=bool)
    try:
        for seed in (1, 2, 3, 1001, 1002, 1003):     # results of sin/cos +-1 ulp; from 1000 on their arguments too
            O.set_trig_noise(seed)
            r = _oracle_final(s, uc, m, th, x0, y0)
            same = base["d_ray"][2] == r["d_ray"][2]
            unstable |= ~same
            dev_all = max(dev_all, float(np.abs(base["final"] - r["final"])[:, same].max()))
            dev_d = max(dev_d, float(np.abs(base["d_ray"][:2] - r["d_ray"][:2])[:, same].max()))
    finally:
        O.set_trig_noise(0)
    return dict(tol, all=max(tol["all"], SAFETY * dev_all), d_ray=max(tol["d_ray"], SAFETY * dev_d)), unstable


# ------------------------------------------------------------------------------------------------------------------
# Dense fans: the WHOLE fan is launched (>= 2^20 rays: warp-staged cell records, 128-thread blocks, ray queue), sampled
# rays are compared with the oracle and with the rows the unmodified reference produced for rays of these very fans
# (tests/golden/ref_traj.npz synth_*; gen_golden.py synthetic_fans).
# ------------------------------------------------------------------------------------------------------------------
DENSE = {"2": (1 << 20, 20), "3": (1 << 26, 20), "4": (1 << 24, 20)}   # fan size, log2 of the launched slice


def _slices(uc, idx):
    """Slices of 2^20 consecutive rays of the benchmark fan that hold the fixture rays."""
    R, lg = DENSE[uc]
    n = 1 << lg
    starts = []
    for i in idx:
        lo = min(max(int(i) - n // 2, 0), R - n)
        if not any(a <= i < a + n for a in starts):
            starts.append(lo)
    return R, n, starts


@pytest.mark.parametrize("uc,m", [(uc, m) for uc in "234" for m in SCEN_METHODS[uc]])
def test_dense_fan_vs_oracle_and_reference(engine, golden_traj, sensitivity, uc, m):
    import torch
    s = scenario_setup(uc)
    engine.set_field(s["z"], s["grd"])
    tol = tolerances(sensitivity, uc, m)
    gidx = golden_traj["synth_s%s/idx" % uc]
    key = "synth_s%s_op%d/" % (uc, m)
    R, n, starts = _slices(uc, gidx)
    rng = np.random.default_rng(100 * int(uc) + m)
    for lo in starts:
        th, x0, y0 = rt.make_fan(uc, R, lo, lo + n)
        r = engine.trace(m, *_dev(th, x0, y0), s["step"], s["M"], s["gamma"], s["box"], full=False)   # static grid
        d = r.d_ray.cpu().numpy()
        last = r.last_row.cpu().numpy()
        assert r.stats["ray_steps"] == int(d[2].sum()) and r.stats["nonfinite_rays"] == 0
        # (a) oracle on sampled rays: a run of 40 consecutive rays (one warp and its neighbour: exit steps straddle
        #     inside a staged warp) + 24 scattered ones + the fixture rays of this slice
        a0 = int(rng.integers(0, n - 40))
        pick = np.unique(np.concatenate([np.arange(a0, a0 + 40), rng.integers(0, n, 24),
                                         [i - lo for i in gidx if lo <= i < lo + n]])).astype(int)
        o = _oracle_final(s, uc, m, th[pick], x0[pick], y0[pick])
        what = "dense s%s op%d slice %d" % (uc, m, lo)
        if m in (3, 4, 5, 7, 10):
            tol_m, unstable = _measured_tolerance(s, uc, m, th[pick], x0[pick], y0[pick], o, tol)
            assert unstable.mean() < 0.1 and np.abs(d[2, pick] - o["d_ray"][2])[unstable].max(initial=0) <= 1
            pick, o = pick[~unstable], {k: (v[..., ~unstable] if isinstance(v, np.ndarray) else v) for k, v in o.items()}
            tol = dict(tol, all=tol_m["all"], d_ray=tol_m["d_ray"])
        same = check_steps(d[2, pick], o["d_ray"][2], tol, what)
        what += " rays %s..%s" % (lo + pick[0], lo + pick[-1])
        assert_close(last[:, pick][:, same], o["final"][:, same], tol["all"], what + " last_row")
        assert_close(d[:2, pick][:, same], o["d_ray"][:2, same], tol["d_ray"], what + " d_ray")
        # (b) the reference's own rows for the fixture rays
        for j, i in enumerate(gidx):
            if lo <= i < lo + n:
                k = int(i - lo)
                gd = golden_traj[key + "d_ray"][:, j]
                ok = check_steps(d[2, k:k + 1], gd[2:3], tol, what + " golden")
                if ok.all():
                    assert_close(last[:, k], golden_traj[key + "last_row"][:, j], tol["all"], what + " golden last_row")
                    assert_close(d[:2, k], gd[:2], tol["d_ray"], what + " golden d_ray")
        # (c) the ray queue (persistent warps, compaction) gives the same bits as the static grid
        if m in (6, 1, 11):
            st = engine.trace(m, *_dev(th, x0, y0), s["step"], s["M"], s["gamma"], s["box"], full=False, ray_queue=True)
            assert torch.equal(st.d_ray, r.d_ray) and torch.equal(st.last_row, r.last_row)


@pytest.mark.parametrize("gamma", [0.6, 1.7, 3.0, 4.5])
@pytest.mark.parametrize("m", [10, 11])
def test_anisotropic_golden_other_gammas(engine, sensitivity, m, gamma):
    """The anisotropic golden search (evaluating passes + the Taylor tail, step_fast.cuh golden_aniso_tail) on the aniso
    scenario's field with other anisotropy factors than the scenario's 3: the bound on the expansion's error depends on
    gamma.  A dense run of 4096 rays (staged warps) against the oracle on sampled rays; step counts exact."""
    s = scenario_setup("4")
    engine.set_field(s["z"], s["grd"])
    tol = tolerances(sensitivity, "4", m)
    n = 4096
    th, x0, y0 = rt.make_fan("4", 1 << 20, 300000, 300000 + n)
    s = dict(s, gamma=gamma)
    r = engine.trace(m, *_dev(th, x0, y0), s["step"], s["M"], gamma, s["box"], full=False, want_px_stats=True)
    d, last = r.d_ray.cpu().numpy(), r.last_row.cpu().numpy()
    assert r.stats["nonfinite_rays"] == 0
    pick = np.unique(np.concatenate([np.arange(1000, 1040), np.linspace(0, n - 1, 24).astype(int)]))
    o = _oracle_final(s, "4", m, th[pick], x0[pick], y0[pick])
    tol_m, unstable = _measured_tolerance(s, "4", m, th[pick], x0[pick], y0[pick], o, tol)
    assert unstable.mean() < 0.1
    ok = ~unstable
    what = "aniso gamma %.1f op%d" % (gamma, m)
    assert np.array_equal(d[2, pick][ok], o["d_ray"][2][ok]), what + " steps"
    assert_close(last[:, pick][:, ok], o["final"][:, ok], tol_m["all"], what + " last_row")
    assert_close(d[:2, pick][:, ok], o["d_ray"][:2, ok], tol_m["d_ray"], what + " d_ray")


def test_interface_dense_fan_vs_oracle(engine, sensitivity):
    """interface 2^18 rays op6 + Snell metric in FINAL mode (two-pass with checkpoints) on the queue kernels."""
    from oracle import rt_oracle as O
    s = scenario_setup("1")
    engine.set_field(s["z"], s["grd"])
    tol = tolerances(sensitivity, "1", 6)
    R = 1 << 18
    th, x0, y0 = rt.make_fan("1", R)
    r = engine.trace(6, *_dev(th, x0, y0), s["step"], s["M"], s["gamma"], s["box"], interface=True, full=False, ray_queue=True)
    pick = np.unique(np.concatenate([np.arange(7000, 7040), np.linspace(0, R - 1, 24).astype(int)]))
    o = O.trace(6, s["z"], s["grd"], s["step"], s["M"], s["gamma"], s["box"], True, th[pick], x0[pick], y0[pick], full=True,
                nthreads=16)
    d = r.d_ray.cpu().numpy()
    same = check_steps(d[2, pick], o["d_ray"][2], tol, "interface dense")
    assert_close(r.last_row.cpu().numpy()[:, pick][:, same], o["final"][:, same], tol["all"], "interface dense last_row")
    assert_close(r.errors.cpu().numpy()[pick][same], o["errors"][same], tol["err"], "interface dense errors")


# ------------------------------------------------------------------------------------------------------------------
# Ray queue + warp compaction == static grid, bit for bit
# ------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", ["sorted", "shuffled", "mixed_steps", "interface_errors", "aniso_stats", "op7"])
def test_ray_queue_equals_static_grid(engine, case):
    import torch
    uc, m = {"interface_errors": ("1", 4), "aniso_stats": ("4", 10), "op7": ("3", 7)}.get(case, ("3", 6))
    s = scenario_setup(uc)
    engine.set_field(s["z"], s["grd"])
    R = (1 << 17) + 36
    th, x0, y0 = rt.make_fan(uc, R)
    kw = dict(full=False, want_px_stats=True, interface=(uc == "1"))
    M = s["M"]
    if case in ("shuffled", "op7"):
        th = np.random.default_rng(5).permutation(th)           # neighbours in a warp end thousands of steps apart
    if case == "mixed_steps":                                    # DELTA_S candidates of different length interleaved
        div = np.random.default_rng(6).choice([0.5, 1.0, 2.0, 20.0], R)
        steps = rt.SIGMA / div
        msz = np.ceil(80 / steps).astype(np.int64) + 1
        msz[::97] = 1                                            # rays without a single step
        kw.update(steps=_dev(steps)[0], max_sizes=_dev(msz)[0])
        M = int(msz.max())
    if case == "aniso_stats":
        M = 1200
    a = engine.trace(m, *_dev(th, x0, y0), None if case == "mixed_steps" else s["step"], M, s["gamma"], s["box"],
                     ray_queue=True, **kw)
    b = engine.trace(m, *_dev(th, x0, y0), None if case == "mixed_steps" else s["step"], M, s["gamma"], s["box"],
                     static_grid=True, **kw)
    for name in ("d_ray", "last_row", "px_stats") + (("errors",) if uc == "1" else ()):
        ta, tb = a[name], b[name]
        assert torch.equal(torch.nan_to_num(ta, nan=-1.0), torch.nan_to_num(tb, nan=-1.0)), (case, name)
    assert a.stats["ray_steps"] == b.stats["ray_steps"] == int(a.d_ray[2].sum().item())
    if case == "mixed_steps":
        assert bool((a.d_ray[2][::97] == 0).all())


def test_sort_fan_gives_the_same_results(engine):
    """RTB_F_SORT_FAN (Engine.trace(sort_fan=True)): an unsorted fan traced in the order of (x0, y0, theta0) inside the
    library, results handed back at the caller's ray index: the same bits as the unsorted call -- device and host
    pointers, shuffled angles, scattered start points, per-ray DELTA_S candidates, the Snell metric."""
    import torch
    s = scenario_setup("3")
    engine.set_field(s["z"], s["grd"])
    R = (1 << 16) + 5
    rng = np.random.default_rng(11)
    th, x0, y0 = rt.make_fan("3", R)
    th = rng.permutation(th)
    kw = dict(full=False, want_px_stats=True)
    a = engine.trace(6, *_dev(th, x0, y0), s["step"], s["M"], s["gamma"], s["box"], **kw)
    b = engine.trace(6, *_dev(th, x0, y0), s["step"], s["M"], s["gamma"], s["box"], sort_fan=True, **kw)
    for name in ("d_ray", "last_row", "px_stats"):
        assert torch.equal(torch.nan_to_num(a[name], nan=-1.0), torch.nan_to_num(b[name], nan=-1.0)), name
    assert a.stats["ray_steps"] == b.stats["ray_steps"]
    # host pointers; start points on three lines, negative and repeated angles, two DELTA_S candidates mixed per ray
    R = 5000
    th = rng.uniform(-0.3, 1.5, R); th[::7] = th[3]
    x0 = rng.choice([-2.0, -1.5, -1.0], R); y0 = rng.choice([-2.0, -1.75], R)
    steps = rng.choice([s["step"], 2.5 * s["step"]], R)
    msz = np.where(steps == s["step"], s["M"], s["M"] // 2).astype(np.int64)
    kw = dict(full=False, want_px_stats=True, steps=steps, max_sizes=msz)
    a = engine.trace(6, th, x0, y0, None, s["M"], s["gamma"], s["box"], **kw)
    b = engine.trace(6, th, x0, y0, None, s["M"], s["gamma"], s["box"], sort_fan=True, **kw)
    for name in ("d_ray", "last_row", "px_stats"):
        assert np.array_equal(a[name], b[name], equal_nan=True), name
    s1 = scenario_setup("1")
    engine.set_field(s1["z"], s1["grd"])
    th, x0, y0 = rt.make_fan("1", 3000)
    pm = rng.permutation(3000)
    a = engine.trace(6, th[pm], x0[pm], y0[pm], s1["step"], s1["M"], s1["gamma"], s1["box"], interface=True, full=False)
    b = engine.trace(6, th[pm], x0[pm], y0[pm], s1["step"], s1["M"], s1["gamma"], s1["box"], interface=True, full=False,
                     sort_fan=True)
    assert np.array_equal(a.errors, b.errors, equal_nan=True) and np.array_equal(a.d_ray, b.d_ray)
    with pytest.raises(Exception, match="RTB_F_SORT_FAN"):
        engine.trace(6, th, x0, y0, s1["step"], 50, s1["gamma"], s1["box"], full=True, sort_fan=True)


# ------------------------------------------------------------------------------------------------------------------
# Trajectory rows staged in shared memory + 2-D TMA stores == per-lane stores, bit for bit
# ------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("uc,m,R,M,stride,tail_zero", [("2", 6, 1 << 16, 300, 1, True), ("3", 6, (1 << 16) + 10, 1500, 3, True),
                                                       ("3", 1, (1 << 16) + 10, 1400, 1, False), ("4", 11, 1 << 16, 150, 2, True),
                                                       ("3", 7, (1 << 16) + 7, 90, 1, True)])
def test_tma_trajectory_store_equals_per_lane(engine, uc, m, R, M, stride, tail_zero):
    import torch
    s = scenario_setup(uc)
    engine.set_field(s["z"], s["grd"])
    th, x0, y0 = rt.make_fan(uc, R)
    rows = (M + stride - 1) // stride
    out = []
    for mode in (0, 1):          # per-lane stores first, then the staged path (odd R: both runs take the per-lane path)
        engine.set_option("traj_store", mode)
        buf = torch.zeros(rows, 6, R, dtype=torch.float64, device="cuda") if not tail_zero else \
            torch.full((rows, 6, R), -3.0, dtype=torch.float64, device="cuda")
        r = engine.trace(m, *_dev(th, x0, y0), s["step"], M, s["gamma"], s["box"], full=True, out_stride=stride,
                         out=dict(s_ray=buf), tail_zero=tail_zero, want_px_stats=(uc == "4"))
        out.append(r)
    engine.set_option("traj_store", 0)
    assert torch.equal(out[0].s_ray, out[1].s_ray)
    assert torch.equal(out[0].d_ray, out[1].d_ray) and torch.equal(out[0].last_row, out[1].last_row)
    # and both equal the FINAL-mode run and the rows of the trajectory itself
    fin = engine.trace(m, *_dev(th, x0, y0), s["step"], M, s["gamma"], s["box"], full=False)
    assert torch.equal(fin.d_ray, out[1].d_ray) and torch.equal(fin.last_row, out[1].last_row)
    if stride == 1:
        k = torch.arange(0, R, 997, device="cuda")
        li = out[1].d_ray[2, k].long()
        assert torch.equal(out[1].s_ray[li, :, k].T, out[1].last_row[:, k])
        if tail_zero:                                           # rows after a ray's last step are zero (np.zeros, RT:802)
            for kk, l in zip(k.tolist()[:20], li.tolist()[:20]):
                assert not bool(out[1].s_ray[l + 1:, :, kk].any())


# ------------------------------------------------------------------------------------------------------------------
# Small rows
# ------------------------------------------------------------------------------------------------------------------
def test_strided_trajectory_snell_errors(engine):
    """a strided trajectory with the Snell metric -- the second pass must not read checkpoints the
    trajectory pass never wrote."""
    s = scenario_setup("1")
    engine.set_field(s["z"], s["grd"])
    a = (s["theta0"], s["x0"], s["y0"], s["step"], s["M"], s["gamma"], s["box"])
    engine.trace(6, *a, interface=True, full=False)                       # leaves FINAL-mode checkpoints in the pool
    full = engine.trace(4, *a, interface=True, full=True)
    st = engine.trace(4, *a, interface=True, full=True, out_stride=7)
    fin = engine.trace(4, *a, interface=True, full=False)
    assert np.array_equal(st.errors, full.errors) and np.array_equal(fin.errors, full.errors)
    assert np.array_equal(st.s_ray, full.s_ray[::7])


@pytest.mark.parametrize("exact", [False, True])
def test_rays_leaving_on_their_first_steps_vs_reference(engine, exact):
    """ref_edge.npz: op7's two bootstrap steps carry no box test (RT:838-866); op6 / op1 break at once."""
    g = np.load(os.path.join(GOLDEN, "ref_edge.npz"))
    s = scenario_setup("3")
    engine.set_field(s["z"], s["grd"])
    theta, pos_x = g["theta"], g["pos_x"]
    for m in (7, 6, 1):
        r = engine.trace(m, theta, pos_x, -2.0, s["step"], s["M"], s["gamma"], s["box"], full=True, exact=exact)
        d = g["s3_op%d/d_ray" % m]
        assert np.array_equal(r.d_ray[2], d[2]), (m, r.d_ray[2])
        assert_close(r.d_ray[:2], d[:2], 1e-10, "edge d_ray")
        assert_close(r.s_ray[:8], g["s3_op%d/rows" % m], 1e-10, "edge rows")
        fin = engine.trace(m, theta, pos_x, -2.0, s["step"], s["M"], s["gamma"], s["box"], full=False, exact=exact)
        assert np.array_equal(fin.d_ray, r.d_ray)


def test_nonfinite_guard_flag(engine):
    s = scenario_setup("3")
    engine.set_field(s["z"], s["grd"])
    th = np.array([0.3, np.nan, 0.9, 1.2])
    r = engine.trace(6, th, -2.0, -2.0, s["step"], 400, s["gamma"], s["box"], full=False)
    assert r.stats["nonfinite_rays"] == 1 and r.d_ray[2, 1] == 399      # stepped on to max_size like the reference
    assert np.isfinite(r.last_row[:, [0, 2, 3]]).all()
    r = engine.trace(6, th[[0, 2, 3]], -2.0, -2.0, s["step"], 400, s["gamma"], s["box"], full=False)
    assert r.stats["nonfinite_rays"] == 0


@pytest.mark.parametrize("uc,m", [("1", 6), ("2", 6), ("3", 4), ("4", 11)])
def test_rtb_sweep_equals_per_ray_reduction(engine, uc, m):
    """rtb_sweep (device-side per-candidate sum / max / count) == the same sweep traced ray by ray and reduced in numpy
    the way RT:954-957, 1354-1360 do."""
    s = scenario_setup(uc)
    engine.set_field(s["z"], s["grd"])
    divisors, options, _ = rt.sweep_candidates(uc)
    sel = np.linspace(0, len(options) - 1, 7).astype(int)
    options, divisors = options[sel], divisors[sel]
    msz = np.array([rt.max_size_for(uc, st, d + 1) for st, d in zip(options, divisors)], dtype=np.int64)
    th, x0, y0 = s["theta0"], s["x0"], s["y0"]
    R, Cn = len(th), len(options)
    metric = {"1": "snell", "2": "closure"}.get(uc, "px_cv")
    skip = 1 if metric == "px_cv" else 0
    sm, mx, cnt, stats = engine.sweep(m, th, x0, y0, options, msz, s["gamma"], s["box"], metric, skip_lo=skip, skip_hi=skip)
    r = engine.trace(m, np.tile(th, Cn), np.tile(x0, Cn), np.tile(y0, Cn), None, int(msz.max()), s["gamma"], s["box"],
                     interface=(uc == "1"), full=False, want_px_stats=True, steps=np.repeat(options, R),
                     max_sizes=np.repeat(msz, R))
    assert stats["ray_steps"] == r.stats["ray_steps"]
    if uc == "1":
        e = r.errors.reshape(Cn, R)
        assert (cnt == R).all() and np.array_equal(mx, e.max(axis=1))
        assert_close(sm / cnt, e.mean(axis=1), 1e-13, "sweep mean")
    elif uc == "2":
        last = r.last_row.reshape(6, Cn, R)
        done = r.d_ray[2].reshape(Cn, R)[:, 0] == msz - 1
        want = [100 * np.linalg.norm(np.array([1, 0]) - (last[0:2, j, 0] if done[j] else 0)) / (2 * np.pi) for j in range(Cn)]
        assert (cnt == 1).all()
        assert_close(sm, want, 1e-13, "closure")
    else:
        st = r.px_stats.reshape(3, Cn, R)
        cv = 100 * st[2] / st[1]
        assert (cnt == R - 2).all()
        assert_close(sm / cnt, cv[:, 1:R - 1].mean(axis=1), 1e-12, "cv mean")
        assert np.array_equal(mx, cv[:, 1:R - 1].max(axis=1))


def test_fused_gather_block_cyclic_two_contexts():
    """rtb_set_gather_cyclic: block-cyclic shards (dist.cyclic_indices) land at their global indices."""
    import torch
    from raytracing_b200.dist import _DevArray, cyclic_indices
    s = scenario_setup("3")
    R, block = 5 * 1024 + 300, 1024
    dev = torch.device("cuda", 0)
    th, x0, y0 = rt.make_fan("3", R)
    eng = [rt.Engine(0), rt.Engine(0)]
    bufs = [e.gather_alloc(3 * 8 * R)[0] for e in eng]
    views = [torch.as_tensor(_DevArray(b, R), device=dev) for b in bufs]
    for v in views:
        v.fill_(-7.0)
    whole = torch.full((3, R), -7.0, dtype=torch.float64, device=dev)
    for rank, e in enumerate(eng):
        e.set_field(s["z"], s["grd"])
        idx = cyclic_indices(R, 2, rank, block)
        e.set_gather_cyclic(bufs, R, block, 2, rank)
        p = e.trace(6, *(torch.from_numpy(a[idx]).to(dev) for a in (th, x0, y0)), s["step"], s["M"], s["gamma"], s["box"],
                    full=False)
        whole[:, torch.from_numpy(idx).to(dev)] = p.d_ray
    torch.cuda.synchronize()
    assert torch.equal(views[0], whole) and torch.equal(views[1], whole) and not bool((whole == -7.0).any())
    for e, b in zip(eng, bufs):
        e.set_gather(None, 0, 0)
        e.gather_close(b, own=True)
        e.close()


def test_cli_benchmark_mode(capsys):
    """A15 on the GPU: the benchmark statistic (RT:1516-1539) over kernel-time samples of the traced scenario."""
    from raytracing_b200 import cli
    info = cli.run("3", 6, "default", benchmark=True, trials=5, show=False)
    out = capsys.readouterr().out
    assert "Completion time per scenario:" in out and 0 < info["completion_time_s"] < 1.0


def test_build_id_matches_tree(engine):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(GOLDEN), "..", "tools"))
    import build_id
    assert engine.build_id() == build_id.digest()
