"""Drop-in surface of RT_bench.py's hot path: trazar, search_delta, n_gradient, the DELTA_S
finder and the benchmark statistic, with the reference's names, argument meaning and return
values, executed by librtb.so on the GPU.
"""
import numpy as np

from . import constants as K
from .engine import Engine, default_fan, max_size_for, method_id
from .field import build_field

_ENGINES = {}


def get_engine(device=0):
    """Process-wide Engine per device (created on first use; raises without a GPU)."""
    if device not in _ENGINES:
        _ENGINES[device] = Engine(device)
    return _ENGINES[device]


def interpolacion_gpu(x, y, Z, delta=None, engine=None):
    """interpolacion(x, y, Z) (RT_bench.py:435-464) with np.gradient and the three RectBivariateSpline fits done on
    the GPU (rtb_fit_field): returns (z, grd) handles every function here accepts in place of the scipy objects.
    The gradient is bit-exact, the spline coefficients equal scipy's to a few ulp of max|gradient|."""
    eng = engine or get_engine()
    return eng.fit_field(x, y, Z, K.DELTA if delta is None else delta)


_GPU_FIELDS = {}


def build_field_gpu(user_choice, engine=None):
    """build_field with the fit on the GPU: genZ on the host (the scenario functions are numpy expressions of the
    reference, RT:201-261), everything after it on the device.  Cached per (engine, scenario function)."""
    from .field import genZ
    eng = engine or get_engine()
    uc = K.scenario_id(user_choice)
    f = K.SCENARIO_FUNCS[uc]
    lim = K.constants(uc)[5:9]
    key = (id(eng), f.__name__, lim)
    if key not in _GPU_FIELDS:
        x, y, _X, _Y, Z = genZ(lim[0], lim[1], lim[2], lim[3], f)
        _GPU_FIELDS[key] = interpolacion_gpu(x, y, Z, engine=eng)
    return _GPU_FIELDS[key]


def n_gradient(vector, grd, z, engine=None):
    """RT_bench.py:141-156: returns (n, array([dn/dx, dn/dy])) at `vector` = (x, y)."""
    eng = engine or get_engine()
    eng.set_field(z, grd)
    n, gx, gy = eng.n_gradient([vector[0]], [vector[1]])
    return n[0], np.array([gx[0], gy[0]])


def _scenario(user_choice):
    uc = K.scenario_id(user_choice)
    c = K.constants(uc)
    return uc, c[0], c[5:9]


def trazar(selected_func, z, grd, show, step, divisor, user_choice, fan=None, engine=None, full=True,
           out_stride=1, exact=False):
    """RT_bench.py:766-948.  Returns (s_ray, d_ray, compute_times, errors) as numpy arrays.

    selected_func: op1..op11 as id, name or the reference's function object.  `fan` =
    (theta0, x0, y0) overrides the scenario's default fan of constants() (RT_bench.py:809-813).
    compute_times (RT_bench.py:882) is kernel time / R for every ray.  With full=False s_ray is
    None (large fans; see trace()).
    """
    eng = engine or get_engine()
    uc, gamma, box = _scenario(user_choice)
    eng.set_field(z, grd)
    theta0, x0, y0 = fan if fan is not None else default_fan(uc)
    M = max_size_for(uc, step, divisor)
    r = eng.trace(selected_func, theta0, x0, y0, step, M, gamma, box, interface=(uc == "1"), full=full,
                  out_stride=out_stride, want_errors=True, exact=exact)
    R = r.d_ray.shape[1]
    compute_times = np.full(R, r.stats["kernel_ms"] * 1e-3 / max(R, 1))
    if show and uc == "1":
        _print_snell_table(r, theta0)
    return r.s_ray, r.d_ray, compute_times, r.errors


def _simulated_angles(r):
    """angsim of RT_bench.py:912-916 for every ray, from the stored trajectory (None without the stride-1 s_ray)."""
    if r.s_ray is None or r.s_ray.shape[0] <= int(r.d_ray[2].max()):
        return None
    out = np.empty(r.d_ray.shape[1])
    for k in range(out.size):
        i = r.d_ray[2, k]
        ia, ib = int(9 * i / 10), int(9.5 * i / 10)
        with np.errstate(invalid="ignore", divide="ignore"):
            out[k] = 180 * np.arctan(np.abs((r.s_ray[ib, 0, k] - r.s_ray[ia, 0, k]) /
                                            (r.s_ray[ib, 1, k] - r.s_ray[ia, 1, k]))) / np.pi
    return out


def _format_num(num):  # RT_bench.py:929-943
    if num < 0:
        return "{: >10.8f}".format(num) if abs(num) < 10 else "{: >10.7f}".format(num)
    return "{: >10.9f}".format(num) if num < 10 else "{: >10.8f}".format(num)


def _print_snell_table(r, theta0):
    """RT_bench.py:921-945.  SimAng needs the two sample rows of the trajectory; without the stride-1 s_ray (FINAL or
    strided output) the column shows the two values error and expected angle allow, SnellAng -+ Err."""
    sim = _simulated_angles(r)
    for k in range(r.d_ray.shape[1]):
        a, b = r.last_row[0, k], r.last_row[1, k]
        e = r.errors[k]
        th = theta0[k]
        if th < np.pi / 4:
            d = 90 - 180 * th / np.pi
        elif th == np.pi / 4:
            d = 0
        else:
            d = 180 * np.arcsin(np.sqrt(2) * np.sin(np.pi / 2 - th)) / np.pi
        c = _format_num(sim[k]) if sim is not None else "%s|%s" % (_format_num(d - e).strip(), _format_num(d + e).strip())
        print(f"Coords: [ {_format_num(a)} , {_format_num(b)} ] | SimAng: {c} | SnellAng: {_format_num(d)} | "
              f"Err: {_format_num(e)} | InitAng: {_format_num(th * 180 / np.pi)}")


def px_cv(s_ray_px):
    """Coefficient of variation (%) of one ray's non-zero p_x entries, RT_bench.py:1358-1359."""
    masked = np.ma.masked_equal(s_ray_px, 0).compressed()
    return 100 * np.std(masked) / np.mean(masked)


def search_delta(option, z, grd, step, divisor, user_choice, fan=None, engine=None):
    """RT_bench.py:950-958: (mean, max) Snell error | fisheye closure % | p_x matrix."""
    uc = K.scenario_id(user_choice)
    rays, _, _, errors = trazar(option, z, grd, False, step, divisor, uc, fan=fan, engine=engine)
    if uc == "1":
        return np.mean(errors), np.max(errors)
    if uc == "2":
        return 100 * np.linalg.norm(np.array([1, 0]) - np.array((rays[-1, 0][0], rays[-1, 1][0]))) / (2 * np.pi)
    return rays[:, 2, :]


# ------------------------------------------------------------------------------------------
# DELTA_S finder RT_bench.py:1299-1385
# ------------------------------------------------------------------------------------------
def sweep_candidates(user_choice):
    """(divisors, delta_s_options, MAX_DEVIATION) of the sweep, RT_bench.py:1302-1312."""
    uc = K.scenario_id(user_choice)
    if uc == "1":
        divisors = np.arange(K.DELTA_S_DIVISOR_UPPER_LIMIT, K.DELTA_S_DIVISOR_LOWER_LIMIT - K.DELTA_STEP, -K.DELTA_STEP)
        return divisors, K.SIGMA / divisors, K.MAX_DEVIATION
    if uc == "2":
        divisors = np.arange(K.DELTA_S_DIVISOR_FISHEYE_UPPER_LIMIT,
                             K.DELTA_S_DIVISOR_FISHEYE_LOWER_LIMIT - K.DELTA_STEP_FISHEYE, -K.DELTA_STEP_FISHEYE)
        return divisors, 2 * np.pi / divisors, 5
    divisors = np.arange(K.DELTA_S_DIVISOR_VERT_UPPER_LIMIT, K.DELTA_S_DIVISOR_VERT_LOWER_LIMIT - 2 * K.DELTA_STEP,
                         -K.DELTA_STEP)
    return divisors, K.SIGMA / divisors, 0.05


def find_index_interface(errors, max_errors, max_deviation):
    """RT_bench.py:1323-1332."""
    if not any(e > max_deviation for e in errors) or not any(e < max_deviation for e in errors):
        return None
    for i in reversed(range(len(errors))):
        if errors[i] < max_deviation and max_errors[i] < 0.8:
            if all(e < max_deviation for e in errors[:i]) and all(e < 0.8 for e in max_errors[:i]):
                return i
    return None


def find_index_fisheye(errors, max_deviation):
    """RT_bench.py:1339-1347."""
    if not any(e > max_deviation for e in errors) or not any(e < max_deviation for e in errors):
        return None
    for i in range(len(errors)):
        if errors[i] > max_deviation:
            return i - 1
    return None


def find_index_vert(errors, max_deviation):
    """RT_bench.py:1363-1373."""
    if not any(e > max_deviation for e in errors) or not any(e < max_deviation for e in errors):
        return None
    for i in range(len(errors)):
        if i > 1:
            if errors[i] > max_deviation:
                if all(e < max_deviation for e in errors[:i - 1]):
                    return i - 1
    return None


def sweep_metrics(option, z, grd, user_choice, delta_s_options, divisors, fan=None, engine=None, group=None):
    """The executor.map(search_delta, ...) of RT_bench.py:1317-1318 as batched GPU traces over candidates x rays with
    the per-candidate reductions done on the device (rtb_sweep), reduced to what the selection rules consume:
      interface -> (error_prom[C], max_error[C]); fisheye -> closure %[C];
      vert/aniso -> scenario_average_CV[C] (RT_bench.py:1354-1360).
    With `group` (a torch.distributed process group) the fan is sharded over the ranks in contiguous blocks and the
    per-candidate partial sums / maxima / counts are all-reduced (dist.reduce_sum_max).
    """
    eng = engine or get_engine()
    uc, gamma, box = _scenario(user_choice)
    eng.set_field(z, grd)
    theta0, x0, y0 = fan if fan is not None else default_fan(uc)
    theta0 = np.asarray(theta0, dtype=np.float64)
    R = len(theta0)
    x0, y0 = (np.broadcast_to(np.asarray(a, dtype=np.float64), (R,)) for a in (x0, y0))
    msz = np.array([max_size_for(uc, s, d + 1) for s, d in zip(delta_s_options, divisors)], dtype=np.int64)
    metric = {"1": "snell", "2": "closure"}.get(uc, "px_cv")
    lo, hi = 0, R
    if group is not None:
        import torch.distributed as dist
        from .dist import shard_bounds
        lo, hi = shard_bounds(R, dist.get_world_size(group), dist.get_rank(group))
    # the reference leaves its first and last ray out of the CV average (range(1, ray_count - 1), RT:1356);
    # the closure error is that of the fan's first ray (rays[-1, 0:2, 0], RT:956)
    skip_lo = 1 if (metric == "px_cv" and lo == 0) else 0
    skip_hi = 1 if (metric == "px_cv" and hi == R) else 0
    if metric == "closure":
        hi = min(hi, 1) if lo == 0 else lo
    sm, mx, cnt, _stats = eng.sweep(option, theta0[lo:hi], x0[lo:hi], y0[lo:hi], delta_s_options, msz, gamma, box, metric,
                                    skip_lo=skip_lo, skip_hi=skip_hi)
    if group is not None:
        import torch
        from .dist import reduce_sum_max
        dev = torch.device("cuda", eng.device) if torch.distributed.get_backend(group) == "nccl" else torch.device("cpu")
        st = torch.as_tensor(np.stack([sm, cnt.astype(np.float64)]), device=dev)
        mt = torch.as_tensor(mx, device=dev)
        reduce_sum_max(st, mt, group)
        sm, cnt, mx = st[0].cpu().numpy(), st[1].cpu().numpy(), mt.cpu().numpy()
    with np.errstate(invalid="ignore", divide="ignore"):
        mean = sm / cnt
    if uc == "1":
        return list(mean), list(mx)
    return list(mean)


def find_delta_s(option, z, grd, user_choice, fan=None, engine=None):
    """The DELTA_S finder of RT_bench.py:1299-1385.  Returns dict(index, divisor, DELTA_S, metrics);
    index None mirrors 'No suitable divisor was found' (the reference then sys.exit()s, RT:1406)."""
    uc = K.scenario_id(user_choice)
    divisors, options, max_dev = sweep_candidates(uc)
    metrics = sweep_metrics(option, z, grd, uc, options, divisors, fan=fan, engine=engine)
    if uc == "1":
        index = find_index_interface(metrics[0], metrics[1], max_dev)
    elif uc == "2":
        index = find_index_fisheye(metrics, max_dev)
    else:
        index = find_index_vert(metrics, max_dev)
    out = dict(index=index, divisors=divisors, metrics=metrics, divisor=None, DELTA_S=None)
    if index is not None:
        if uc == "2":
            out["divisor"] = round(divisors[index])          # RT:1379
            out["DELTA_S"] = 2 * np.pi / out["divisor"]
        else:
            out["divisor"] = round(divisors[index], 2)       # RT:1383
            out["DELTA_S"] = K.SIGMA / out["divisor"]
    return out


# ------------------------------------------------------------------------------------------
# Benchmark statistic RT_bench.py:1486-1542
# ------------------------------------------------------------------------------------------
def remove_outliers_iqr(data):
    """RT_bench.py:123-138."""
    q1, q3 = np.percentile(data, 25), np.percentile(data, 75)
    iqr = q3 - q1
    return data[(data >= q1 - 1.5 * iqr) & (data <= q3 + 1.5 * iqr)]


def benchmark_statistic(sample_fn, trial=100, replicas=1, max_rounds=None):
    """RT_bench.py:1516-1538 with `sample_fn()` -> seconds replacing one worker's np.sum(result[2]):
    rounds of trial*replicas samples, IQR filter, median of the last 30 %, repeat until two
    consecutive medians agree to 0.5 %, return their mean.  Like the reference the loop is unbounded by default
    (`while True`, RT:1517); `max_rounds` is an optional safety cap for callers with noisy samplers (a deviation:
    the reference has none)."""
    import itertools
    benchmarks = []
    arr = np.zeros(trial * replicas)
    for _ in (itertools.count() if max_rounds is None else range(max_rounds)):
        for idx in range(arr.size):
            arr[idx] = sample_fn()
        cleaned = remove_outliers_iqr(arr)
        benchmarks.append(np.median(cleaned[int(-0.3 * len(cleaned)):]))
        if len(benchmarks) >= 2:
            diff = 100 * abs(benchmarks[-1] - benchmarks[-2]) / max(benchmarks[-1], benchmarks[-2])
            if diff < 0.5:
                break
    return float(np.mean(benchmarks[-2:]))
