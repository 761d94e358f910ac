"""GPU parity tests: the CUDA path through the C ABI vs the oracle and the golden fixtures."""
import os
import numpy as np
import pytest

from _util import SCEN_METHODS, assert_close, assert_close_rays, check_steps, scenario_setup, tolerances

pytestmark = pytest.mark.gpu

CASES = [(uc, m) for uc, ms in SCEN_METHODS.items() for m in ms]


def _oracle(s, uc, m, theta0=None, x0=None, y0=None, **kw):
    from oracle import rt_oracle as O
    theta0 = s["theta0"] if theta0 is None else theta0
    x0 = s["x0"] if x0 is None else x0
    y0 = s["y0"] if y0 is None else y0
    return O.trace(m, s["z"], s["grd"], kw.get("step", s["step"]), kw.get("M", s["M"]), s["gamma"], s["box"],
                   uc == "1", theta0, x0, y0, nthreads=8)


@pytest.mark.parametrize("uc", ["1", "2", "3"])
@pytest.mark.parametrize("exact", [True, False])
def test_field_vs_reference(engine, golden_field, uc, exact):
    """A1: n_gradient vs the reference's own values (golden), incl. knots, clamped points."""
    s = scenario_setup(uc)
    engine.set_field(s["z"], s["grd"])
    x, y = golden_field["s%s/x" % uc], golden_field["s%s/y" % uc]
    n, gx, gy = engine.n_gradient(x, y, exact=exact)
    scale = max(1.0, np.abs(golden_field["s%s/g" % uc]).max())
    tol = 1e-14 if exact else 1e-13
    assert np.abs(n - golden_field["s%s/n" % uc]).max() <= tol
    assert np.abs(gx - golden_field["s%s/g" % uc][:, 0]).max() <= tol * scale
    assert np.abs(gy - golden_field["s%s/g" % uc][:, 1]).max() <= tol * scale


@pytest.mark.parametrize("uc,m", CASES)
def test_default_fan_vs_oracle_and_golden(engine, golden_traj, sensitivity, uc, m):
    """A2-A13: full trajectories of the reference's default fans, every (scenario, method)."""
    s = scenario_setup(uc)
    engine.set_field(s["z"], s["grd"])
    tol = tolerances(sensitivity, uc, m)
    r = engine.trace(m, s["theta0"], s["x0"], s["y0"], s["step"], s["M"], s["gamma"], s["box"], interface=(uc == "1"),
                     full=True, want_px_stats=True)
    assert r.s_ray.shape == (s["M"], 6, len(s["theta0"]))
    o = _oracle(s, uc, m)
    what = "s%s op%d" % (uc, m)
    same = check_steps(r.d_ray[2], o["d_ray"][2], tol, what)
    ch = tol["chaotic"]
    assert_close_rays(r.s_ray[:, :, same], o["s_ray"][:, :, same], tol["all"], what + " s_ray", ch)
    assert_close_rays(r.d_ray[:2, same], o["d_ray"][:2, same], tol["d_ray"], what + " d_ray", ch)
    assert_close_rays(r.last_row[:, same], o["final"][:, same], tol["all"], what + " last_row", ch)
    if uc == "1":
        assert_close_rays(r.errors[same], o["errors"][same], tol["err"], what + " errors", ch)
        # the physics metric the reference reports (RT:1395) agrees regardless
        assert abs(np.mean(r.errors) - np.mean(o["errors"])) <= max(1e-9, 10 * tol["err"] / len(r.errors) ** 0.5)
    # golden rows straight from the reference
    key = "default_s%s_op%d/" % (uc, m)
    same_g = check_steps(r.d_ray[2], golden_traj[key + "d_ray"][2], tol, what + " golden")
    stride = int(golden_traj[key + "stride"])
    assert_close_rays(r.s_ray[::stride][:, :, same_g], golden_traj[key + "rows"][:, :, same_g], tol["all"],
                      what + " golden rows", ch)
    # rows after a ray's last step stay zero (np.zeros, RT:802)
    for k in range(0, r.s_ray.shape[2], 7):
        assert not r.s_ray[int(r.d_ray[2, k]) + 1:, :, k].any()
    # ray-step count = sum d_ray[2]
    assert r.stats["ray_steps"] == int(r.d_ray[2].sum())
    # online p_x statistics == numpy on the full trajectory (RT:1358-1359)
    for k in range(1, r.s_ray.shape[2] - 1, 5):
        px = np.ma.masked_equal(r.s_ray[:, 2, k], 0).compressed()
        assert r.px_stats[0, k] == px.size
        assert abs(r.px_stats[1, k] - np.mean(px)) <= 1e-13 * abs(np.mean(px))
        assert abs(r.px_stats[2, k] - np.std(px)) <= 1e-9 * abs(np.mean(px)) * 1e-3 + 1e-7 * np.std(px)


@pytest.mark.parametrize("uc,m", [("1", 6), ("1", 4), ("1", 7), ("3", 1), ("3", 9), ("4", 10), ("2", 8)])
def test_output_modes_consistent(engine, uc, m):
    """FINAL and STRIDED modes are bit-identical to the FULL run; two-pass Snell == FULL-mode Snell;
    host-pointer and device-pointer paths agree; caller-zeroed buffers with NO_TAIL_ZERO agree."""
    import torch
    s = scenario_setup(uc)
    engine.set_field(s["z"], s["grd"])
    a = (s["theta0"], s["x0"], s["y0"], s["step"], s["M"], s["gamma"], s["box"])
    full = engine.trace(m, *a, interface=(uc == "1"), full=True)
    fin = engine.trace(m, *a, interface=(uc == "1"), full=False)
    assert np.array_equal(full.d_ray, fin.d_ray) and np.array_equal(full.last_row, fin.last_row)
    if uc == "1":
        assert np.array_equal(full.errors, fin.errors)
    st = engine.trace(m, *a, full=True, out_stride=7)
    assert np.array_equal(st.s_ray, full.s_ray[::7])
    dev = [torch.from_numpy(np.ascontiguousarray(v)).cuda() for v in (s["theta0"], s["x0"], s["y0"])]
    pre = torch.zeros(s["M"], 6, len(s["theta0"]), dtype=torch.float64, device="cuda")
    d = engine.trace(m, *dev, s["step"], s["M"], s["gamma"], s["box"], interface=(uc == "1"), full=True,
                     out=dict(s_ray=pre), tail_zero=False)
    assert np.array_equal(d.s_ray.cpu().numpy(), full.s_ray) and np.array_equal(d.d_ray.cpu().numpy(), full.d_ray)


def test_edge_cases(engine):
    """Empty and ragged fans, max_size 1 and 2, rays that leave the box on their first step."""
    s = scenario_setup("3")
    engine.set_field(s["z"], s["grd"])
    g, box, h = s["gamma"], s["box"], s["step"]
    r = engine.trace(6, np.zeros(0), np.zeros(0), np.zeros(0), h, 100, g, box, full=True)
    assert r.d_ray.shape == (3, 0) and r.stats["ray_steps"] == 0
    for R in (1, 31, 33, 129):
        th = np.linspace(0.1, 1.5, R)
        r = engine.trace(6, th, -2.0, -2.0, h, 300, g, box, full=True)
        o = _oracle(s, "3", 6, th, np.full(R, -2.0), np.full(R, -2.0), M=300)
        assert np.array_equal(r.d_ray[2], o["d_ray"][2])
        assert_close(r.s_ray, o["s_ray"], 1e-10, "ragged R=%d" % R)
    th = np.array([0.3, 0.9])
    r = engine.trace(6, th, -2.0, -2.0, h, 1, g, box, full=True)          # max_size 1: only row 0
    assert r.s_ray.shape[0] == 1 and not r.d_ray.any() and r.s_ray[0, 0, 0] == -2.0
    r = engine.trace(7, th, -2.0, -2.0, h, 2, g, box, full=True)          # op7 inside its bootstrap
    o = _oracle(s, "3", 7, th, np.full(2, -2.0), np.full(2, -2.0), M=2)
    assert np.array_equal(r.d_ray[2], [1, 1]) and np.abs(r.s_ray - o["s_ray"]).max() < 1e-12
    th = np.array([np.pi, 2.5, 4.0])                                         # leave the box on step 1
    r = engine.trace(6, th, -2.0, -2.0, h, 50, g, box, full=True)
    o = _oracle(s, "3", 6, th, np.full(3, -2.0), np.full(3, -2.0), M=50)
    assert np.array_equal(r.d_ray[2], o["d_ray"][2]) and np.array_equal(r.d_ray[2], [1, 1, 1])
    assert not r.s_ray[2:].any()


def test_errors_reported_not_fatal(engine):
    from raytracing_b200 import _cabi
    s = scenario_setup("3")
    engine.set_field(s["z"], s["grd"])
    with pytest.raises(ValueError):
        engine.trace(12, s["theta0"], -2.0, -2.0, s["step"], 10, 1, s["box"])
    with pytest.raises(_cabi.RtbError):
        engine.trace(6, s["theta0"], -2.0, -2.0, -1.0, 10, 1, s["box"])


def test_sharded_equals_unsharded(engine):
    """8(e): tracing contiguous blocks of a fan and concatenating == tracing the whole fan, bit for bit."""
    import raytracing_b200 as rt
    from raytracing_b200.dist import shard_bounds
    s = scenario_setup("3")
    engine.set_field(s["z"], s["grd"])
    R = 1000
    th, x0, y0 = rt.make_fan("3", R)
    whole = engine.trace(6, th, x0, y0, s["step"], s["M"], s["gamma"], s["box"], full=False)
    for world in (2, 8):
        parts = []
        for rank in range(world):
            lo, hi = shard_bounds(R, world, rank)
            a = rt.make_fan("3", R, lo, hi)
            parts.append(engine.trace(6, *a, s["step"], s["M"], s["gamma"], s["box"], full=False))
        assert np.array_equal(np.concatenate([p.d_ray for p in parts], 1), whole.d_ray)
        assert np.array_equal(np.concatenate([p.last_row for p in parts], 1), whole.last_row)


@pytest.mark.parametrize("uc,m", [("2", 6), ("1", 4), ("4", 11)])
def test_warp_staged_records_equal_per_lane_loads(engine, uc, m):
    """Fans of >= 2^16 rays read the cell record once per warp through shared memory (field.cuh, warp staging);
    smaller launches load it per lane.  Same table, same arithmetic: tracing a dense fan whole (staged) and in two
    halves (per lane) must agree bit for bit, trajectories included."""
    import raytracing_b200 as rt
    s = scenario_setup(uc)
    engine.set_field(s["z"], s["grd"])
    R = 1 << 16
    M = min(s["M"], 400)
    th, x0, y0 = rt.make_fan(uc, R)
    whole = engine.trace(m, th, x0, y0, s["step"], M, s["gamma"], s["box"], full=True, out_stride=57)
    halves = [engine.trace(m, th[a:b], x0[a:b], y0[a:b], s["step"], M, s["gamma"], s["box"], full=True, out_stride=57)
              for a, b in ((0, R // 2), (R // 2, R))]
    for name in ("d_ray", "last_row"):
        assert np.array_equal(np.concatenate([h[name] for h in halves], 1), whole[name], equal_nan=True), name
    assert np.array_equal(np.concatenate([h.s_ray for h in halves], 2), whole.s_ray, equal_nan=True)


def test_fused_gather_two_contexts_one_gpu():
    """rtb_set_gather: the kernel epilogue of every 'rank' stores its d_ray columns into every gather array.  Two
    contexts on one GPU stand in for two ranks (the IPC mapping itself is covered by tests/mgpu_fused_gather.py)."""
    import torch
    import raytracing_b200 as rt
    from raytracing_b200.dist import _DevArray, shard_bounds
    from raytracing_b200 import _cabi
    s = scenario_setup("3")
    R = 5000
    dev = torch.device("cuda", 0)
    th, x0, y0 = rt.make_fan("3", R)
    eng = [rt.Engine(0), rt.Engine(0)]
    bufs = [e.gather_alloc(3 * 8 * R)[0] for e in eng]
    views = [torch.as_tensor(_DevArray(b, R), device=dev) for b in bufs]
    for v in views:
        v.fill_(-7.0)
    parts = []
    for rank, e in enumerate(eng):
        e.set_field(s["z"], s["grd"])
        lo, hi = shard_bounds(R, 2, rank)
        e.set_gather(bufs, R, lo)
        parts.append(e.trace(6, *(torch.from_numpy(a[lo:hi]).to(dev) for a in (th, x0, y0)), s["step"], s["M"],
                             s["gamma"], s["box"], full=False))
    torch.cuda.synchronize()
    whole = torch.cat([p.d_ray for p in parts], 1)
    assert torch.equal(views[0], whole) and torch.equal(views[1], whole)
    with pytest.raises(_cabi.RtbError):            # a fan that does not fit the window is refused
        eng[1].trace(6, *(torch.from_numpy(a).to(dev) for a in (th, x0, y0)), s["step"], s["M"], s["gamma"], s["box"],
                     full=False)
    eng[0].set_gather(None, 0, 0)                  # off again: the arrays are no longer touched
    views[0].fill_(-7.0)
    eng[0].trace(6, *(torch.from_numpy(a[:100]).to(dev) for a in (th, x0, y0)), s["step"], s["M"], s["gamma"], s["box"],
                 full=False)
    torch.cuda.synchronize()
    assert bool((views[0] == -7.0).all())
    for e, b in zip(eng, bufs):
        e.set_gather(None, 0, 0)
        e.gather_close(b, own=True)
        e.close()


def test_fused_gather_across_processes():
    """Two processes, two GPUs, CUDA IPC peer mappings (skipped on a one-GPU box)."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29577", os.path.join(here, "mgpu_fused_gather.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "fused gather ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("uc,m,div", [("3", 9, 2.74), ("3", 5, 2.53), ("2", 9, 325), ("3", 6, 2.55), ("3", 3, 2.34),
                                      ("3", 9, 1.0), ("3", 6, 0.6)])
def test_coarse_steps_vs_oracle(engine, uc, m, div):
    """The reference's calibrated DELTA_S (RT:1413-1455) are 7-8x the default step, and the coarse end of its sweeps
    larger still: per-step turns leave the small-angle forms of step_fast.cuh (the midpoint golden search hands over to
    the evaluating one beyond 2^-6, the rotations and the step-length series to their general branches).  Same parity
    bar against the CPU restatement as for the default step."""
    import raytracing_b200 as rt
    s = scenario_setup(uc)
    engine.set_field(s["z"], s["grd"])
    if uc == "2":
        step, M = 2 * np.pi / div, rt.max_size_for(uc, 2 * np.pi / div, int(div) + 1)
    else:
        step = rt.SIGMA / div
        M = rt.max_size_for(uc, step, rt.DELTA_S_DIVISOR_FISHEYE + 1)
    r = engine.trace(m, s["theta0"], s["x0"], s["y0"], step, M, s["gamma"], s["box"], full=True)
    o = _oracle(s, uc, m, step=step, M=M)
    assert np.array_equal(r.d_ray[2], o["d_ray"][2])
    assert_close(r.s_ray, o["s_ray"], 1e-10, "s_ray")
    assert_close(r.d_ray[:2], o["d_ray"][:2], 1e-10, "d_ray")
