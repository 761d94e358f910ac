"""Generate the golden fixtures in tests/golden/ from the UNMODIFIED reference.

Runs only in the build container (needs /root/reference); the fixtures are committed so the
tests run anywhere.  Usage:  python tests/golden/gen_golden.py [traj|field|sweep|sens|shoot|edge|bench|all]

  ref_traj.npz    trazar outputs of every (scenario, method) on the default fans (d_ray, errors,
                  strided s_ray rows, last row per ray) and on slices of the synthetic fans
  ref_field.npz   n_gradient at random points and next to knots, per scenario
  ref_sweep.npz   search_delta metrics over the DELTA_S sweep + find_index result (RT:1303-1385)
  ref_shoot.npz   two-point shooting problems solved with RT.trazar rays + RT.golden refinement
  ref_edge.npz    rays leaving the box on their first / second step (op7's bootstrap steps carry no box test)
  ref_bench.npz   remove_outliers_iqr + the median-of-last-30 % / 0.5 % loop of the benchmark mode on recorded samples
  sensitivity.json  the reference algorithm's response to +-1 ulp sin/cos noise (oracle probe,
                  cross-checked against the same probe on the real reference for a few cases)
"""
import json
import os
import sys
import warnings
from multiprocessing import Pool

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
warnings.filterwarnings("ignore")

from oracle import ref_loader as L  # noqa: E402

SCEN = {"1": list(range(1, 10)), "2": list(range(1, 10)), "3": list(range(1, 10)), "4": [10, 11]}
ROWS_KEPT = 48


def synthetic_fans():
    """Slices of the benchmark fans (raytracing_b200.fans.make_fan) small enough for the reference."""
    from raytracing_b200.fans import make_fan
    out = {}
    for uc, R, idx in (("2", 1 << 20, [0, 1, 349525, 524288, 1048574, 1048575]),
                       ("3", 1 << 26, [0, 7, 22369621, 44739242, 67108863]),
                       ("4", 1 << 24, [3, 8388608, 16777215])):
        th = np.concatenate([make_fan(uc, R, i, i + 1)[0] for i in idx])
        out[uc] = (R, np.array(idx), th)
    return out


def _trace_task(args):
    uc, m, fan_theta, tag = args
    RT = L.load_reference()
    fan = None if fan_theta is None else dict(theta_v=fan_theta, ray_count=len(fan_theta))
    z, grd = L.setup_scenario(RT, uc, fan)
    s, d, _ct, e = RT.trazar(L.method(RT, m), z, grd, False, RT.DELTA_S, RT.DELTA_S_DIVISOR_FISHEYE + 1, uc)
    M, _, R = s.shape
    stride = max(1, M // ROWS_KEPT)
    last = np.stack([s[int(d[2, k]), :, k] for k in range(R)], axis=1)
    key = "%s_s%s_op%d" % (tag, uc, m)
    return key, dict(d_ray=d, errors=e, rows=s[::stride].copy(), stride=np.int64(stride), M=np.int64(M),
                     last_row=last, theta0=np.asarray(RT.theta_v[:R], dtype=np.float64),
                     px_sum=s[:, 2, :].sum(axis=0), T_last=last[4])


def gen_traj():
    tasks = [(uc, m, None, "default") for uc, ms in SCEN.items() for m in ms]
    for uc, (_R, _idx, th) in synthetic_fans().items():
        tasks += [(uc, m, th, "synth") for m in SCEN[uc]]
    tasks.sort(key=lambda t: -(100 if t[1] >= 10 else 10 if t[1] in (5, 9) else 1) * (42 if t[0] == "1" else 31))
    flat = {}
    with Pool(8) as pool:
        for key, d in pool.imap_unordered(_trace_task, tasks):
            print("traj", key, d["d_ray"][2].sum(), flush=True)
            for k, v in d.items():
                flat[key + "/" + k] = v
    for uc, (R, idx, _th) in synthetic_fans().items():
        flat["synth_s%s/R" % uc] = np.int64(R)
        flat["synth_s%s/idx" % uc] = idx
    np.savez_compressed(os.path.join(HERE, "ref_traj.npz"), **flat)


def gen_field():
    flat = {}
    for uc in "123":
        RT = L.load_reference()
        z, grd = L.setup_scenario(RT, uc)
        rng = np.random.default_rng(int(uc))
        n = 600
        x = rng.uniform(RT.limx_i - 0.2, RT.limx_s + 0.2, n)
        y = rng.uniform(RT.limy_i - 0.2, RT.limy_s + 0.2, n)
        tu, tv, _ = grd[0].tck
        # on and right next to knots, and clamped points outside the grid
        ku = rng.choice(tu[4:-4], 60); kv = rng.choice(tv[4:-4], 60)
        x = np.concatenate([x, kv, np.nextafter(kv, -np.inf), np.nextafter(kv, np.inf), [tv[0] - 1, tv[-1] + 1, tv[0], tv[-1]]])
        y = np.concatenate([y, ku, np.nextafter(ku, np.inf), np.nextafter(ku, -np.inf), [tu[0] - 1, tu[-1] + 1, tu[-1], tu[0]]])
        res = [RT.n_gradient(np.array([a, b]), grd, z) for a, b in zip(x, y)]
        flat["s%s/x" % uc] = x; flat["s%s/y" % uc] = y
        flat["s%s/n" % uc] = np.array([r[0] for r in res])
        flat["s%s/g" % uc] = np.array([r[1] for r in res])
        print("field", uc, len(x), flush=True)
    np.savez_compressed(os.path.join(HERE, "ref_field.npz"), **flat)


def _sweep_task(args):
    uc, m, step, div = args
    RT = L.load_reference()
    z, grd = L.setup_scenario(RT, uc)
    r = RT.search_delta(L.method(RT, m), z, grd, step, div + 1, uc)
    if uc in "34":  # reduce the p_x matrix exactly as RT:1354-1360 does
        cvs = np.zeros(RT.ray_count - 2)
        for i in range(1, RT.ray_count - 1):
            masked = np.ma.masked_equal(r[:, i], 0).compressed()
            cvs[i - 1] = 100 * np.std(masked) / np.mean(masked)
        r = np.mean(cvs)
    return r


def gen_sweep():
    from raytracing_b200 import api
    flat = {}
    with Pool(8) as pool:
        for uc, m in (("2", 6), ("2", 1), ("3", 6), ("3", 4), ("1", 6), ("4", 11)):
            divisors, options, max_dev = api.sweep_candidates(uc)
            if uc == "4":  # 2 ms/step in the reference: keep every 8th candidate
                divisors, options = divisors[::8], options[::8]
            res = pool.map(_sweep_task, [(uc, m, s, d) for s, d in zip(options, divisors)])
            key = "s%s_op%d" % (uc, m)
            if uc == "1":
                prom, mx = [r[0] for r in res], [r[1] for r in res]
                idx = api.find_index_interface(prom, mx, max_dev)
                flat[key + "/mean"] = np.array(prom); flat[key + "/max"] = np.array(mx)
            else:
                idx = (api.find_index_fisheye if uc == "2" else api.find_index_vert)(res, max_dev)
                flat[key + "/metric"] = np.array(res)
            flat[key + "/index"] = np.int64(-1 if idx is None else idx)
            flat[key + "/divisors"] = divisors
            print("sweep", key, "index", idx, flush=True)
    np.savez_compressed(os.path.join(HERE, "ref_sweep.npz"), **flat)


def gen_sens():
    """Oracle noise probe on the default fans: per (scenario, method) the max deviation of positions,
    of all s_ray columns, of the Snell errors, and which rays change their step count."""
    from oracle import rt_oracle as O
    import raytracing_b200 as rt
    out = {}
    for uc, ms in SCEN.items():
        z, grd = rt.build_field(uc)
        c = rt.scenario_constants(uc)
        theta0, x0, y0 = rt.default_fan(uc)
        M = rt.max_size_for(uc, rt.DELTA_S, 91)
        for m in ms:
            O.set_trig_noise(0)
            base = O.trace(m, z, grd, rt.DELTA_S, M, c[0], c[5:9], uc == "1", theta0, x0, y0, nthreads=8)
            dev = dict(pos=0.0, all=0.0, d_ray=0.0, err=0.0, unstable_rays=[])
            for seed in (1, 2, 3):
                O.set_trig_noise(seed)
                r = O.trace(m, z, grd, rt.DELTA_S, M, c[0], c[5:9], uc == "1", theta0, x0, y0, nthreads=8)
                same = base["d_ray"][2] == r["d_ray"][2]
                dev["unstable_rays"] = sorted(set(dev["unstable_rays"]) | set(np.nonzero(~same)[0].tolist()))
                ds = np.abs(base["s_ray"] - r["s_ray"])[:, :, same]
                dev["pos"] = max(dev["pos"], float(ds[:, 0:2].max()))
                dev["all"] = max(dev["all"], float(ds.max()))
                dev["d_ray"] = max(dev["d_ray"], float(np.abs(base["d_ray"][:2] - r["d_ray"][:2])[:, same].max()))
                dev["err"] = max(dev["err"], float(np.abs(base["errors"] - r["errors"])[same].max()))
            O.set_trig_noise(0)
            out["s%s_op%d" % (uc, m)] = dev
            print("sens", uc, m, dev, flush=True)
    json.dump(out, open(os.path.join(HERE, "sensitivity.json"), "w"), indent=1)


def _shoot_task(args):
    """Two-point search (oracle/shoot_oracle.py procedure) with every ray traced by the REAL RT.trazar and
    the refinement done by the REAL RT.golden."""
    uc, m, div, true_angles, A = args
    from oracle import shoot_oracle as S
    RT = L.load_reference()
    fan = dict(theta_v=np.zeros(1))
    z, grd = L.setup_scenario(RT, uc, fan)
    step = RT.SIGMA / div
    box = (RT.limx_i, RT.limx_s, RT.limy_i, RT.limy_s)

    def trace_fn(th):
        fan["theta_v"] = np.asarray(th, dtype=np.float64)
        s, d, _ct, _e = RT.trazar(L.method(RT, m), z, grd, False, step, RT.DELTA_S_DIVISOR_FISHEYE + 1, uc)
        return s, d
    tr = S.Tracer(m, z, grd, step, 0, RT.gamma, box, (-2.0, -2.0), trace_fn=trace_fn)
    lo_a = 2 * np.pi / 60 if uc == "1" else 0.0
    angles = np.linspace(lo_a, np.pi / 2, A)
    e_true, _ = tr(true_angles)
    tx, ty = e_true[0], e_true[1]
    ex, ns = tr(angles)
    out = dict(angles=angles, true_angles=np.asarray(true_angles), tx=tx, ty=ty, T_true=e_true[2], step=np.float64(step),
               coarse_exit=ex, coarse_nsteps=ns)
    res = {k: [] for k in ("theta", "miss", "T", "exit_x", "exit_y", "nsteps", "index")}
    for t in range(len(tx)):
        m2 = (ex[0] - tx[t]) ** 2 + (ex[1] - ty[t]) ** 2
        j = int(np.argmin(m2))
        lo, hi = angles[max(j - 1, 0)], angles[min(j + 1, A - 1)]
        f = lambda th: float(((tr([th])[0][:2, 0] - np.array([tx[t], ty[t]])) ** 2).sum())
        theta = RT.golden(f, lo, hi)
        e, n1 = tr([theta])
        for k, v in zip(res, (theta, np.hypot(e[0, 0] - tx[t], e[1, 0] - ty[t]), e[2, 0], e[0, 0], e[1, 0], n1[0], j)):
            res[k].append(v)
    out.update({k: np.array(v) for k, v in res.items()})
    return "s%s_op%d" % (uc, m), out


def gen_shoot():
    """ref_shoot.npz: small two-point problems solved through the real reference (parity anchor of the
    parts; the procedure itself is ours, see oracle/shoot_oracle.py)."""
    tasks = [("1", 6, 3.0, np.array([0.3, 0.6, 0.9, 1.0, 1.2, 1.4]) + 0.00371, 64),
             ("1", 8, 2.0, np.array([0.5, 1.1, 1.3]) + 0.00113, 48),
             ("3", 8, 2.0, np.array([0.4, 0.9, 1.3]) + 0.00229, 48)]
    flat = {}
    with Pool(3) as pool:
        for key, d in pool.imap_unordered(_shoot_task, tasks):
            print("shoot", key, d["theta"] - d["true_angles"], d["miss"], flush=True)
            for k, v in d.items():
                flat[key + "/" + k] = v
    np.savez_compressed(os.path.join(HERE, "ref_shoot.npz"), **flat)


def gen_edge():
    """ref_edge.npz: rays that leave the box on their first or second step.  op7 takes its two bootstrap steps
    before the loop that holds the box test (RT:838-866), so such a ray still reaches row 3; every other method
    breaks at once (RT:878)."""
    flat = {}
    theta = np.array([np.pi, np.pi, 2.8, -2.5, 0.3])
    pos_x = np.array([-2.0, -1.996, -2.0, -1.996, -2.0])
    for m in (7, 6, 1):
        RT = L.load_reference()
        z, grd = L.setup_scenario(RT, "3", dict(theta_v=theta, pos_x=pos_x, ray_count=len(theta)))
        s, d, _ct, _e = RT.trazar(L.method(RT, m), z, grd, False, RT.DELTA_S, RT.DELTA_S_DIVISOR_FISHEYE + 1, "3")
        flat["s3_op%d/d_ray" % m] = d
        flat["s3_op%d/rows" % m] = s[:8].copy()
        print("edge op%d" % m, d[2], flush=True)
    flat["theta"] = theta
    flat["pos_x"] = pos_x
    np.savez_compressed(os.path.join(HERE, "ref_edge.npz"), **flat)


def gen_bench():
    """ref_bench.npz: the benchmark statistic of RT:1516-1538 on recorded sample streams.  remove_outliers_iqr is the
    reference's own function (RT:123-138); the loop around it lives inline in main() (it spawns worker processes and
    reads input()), so its four statistic statements are replayed here verbatim on the recorded samples."""
    RT = L.load_reference()
    flat = {}
    for case, (seed, n, drift) in enumerate(((1, 300, 0.0), (2, 200, 0.02), (3, 400, 0.004))):
        rng = np.random.default_rng(seed)
        rounds, benchmarks = [], []
        while True:
            base = 1.0 + drift / (1 + len(rounds))                 # rounds settle like a warming-up machine
            arr = base + 0.01 * rng.standard_normal(n)
            arr[rng.integers(0, n, 7)] += rng.uniform(0.2, 3.0, 7)  # outliers for the IQR filter
            rounds.append(arr.copy())
            cleaned_arr = RT.remove_outliers_iqr(arr)                                   # RT:1530
            new_size = len(cleaned_arr)                                                 # RT:1531
            benchmarks.append(np.median(cleaned_arr[int(-0.3 * new_size):]))            # RT:1532
            if len(benchmarks) >= 2:                                                    # RT:1534-1537
                diff_percent = 100 * abs(benchmarks[-1] - benchmarks[-2]) / max(benchmarks[-1], benchmarks[-2])
                if diff_percent < 0.5:
                    break
        flat["case%d/samples" % case] = np.stack(rounds)
        flat["case%d/benchmarks" % case] = np.array(benchmarks)
        flat["case%d/final" % case] = np.float64(np.mean(benchmarks[-2:]))              # RT:1539
        flat["case%d/cleaned0" % case] = RT.remove_outliers_iqr(rounds[0])
        print("bench case", case, len(rounds), "rounds ->", flat["case%d/final" % case], flush=True)
    np.savez_compressed(os.path.join(HERE, "ref_bench.npz"), **flat)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("sens", "all"):
        gen_sens()
    if what in ("field", "all"):
        gen_field()
    if what in ("traj", "all"):
        gen_traj()
    if what in ("sweep", "all"):
        gen_sweep()
    if what in ("shoot", "all"):
        gen_shoot()
    if what in ("edge", "all"):
        gen_edge()
    if what in ("bench", "all"):
        gen_bench()
