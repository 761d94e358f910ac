"""Non-interactive mirror of RT_bench.py's prompts: every `input()` of the reference
(scenario RT:1555-1583, algorithm RT:1206-1294, DELTA_S search / calibrated RT:1296-1460, benchmark RT:1486-1542) is a
flag, the RESULTS block is printed with the reference's own wording (RT:1390-1402, 1465-1478, 1486, 1541), and the
work runs on the GPU through librtb.so.  Plotting (graficar) is out of scope.

  python -m raytracing_b200 --scenario interface --method 6 --delta-s calibrated
  python -m raytracing_b200 --scenario 4 --method 2 --delta-s search --benchmark
"""
import argparse
import json
import sys
import time

import numpy as np

from . import constants as K
from .api import benchmark_statistic, build_field_gpu, find_delta_s, get_engine, px_cv, trazar
from .field import build_field

SCENARIOS = {"interface": "1", "fisheye": "2", "vert_heterogeneous": "3", "anisotropy": "4"}
MESSAGES = [  # RT:1211-1221, 1273-1276
    " 1st order Taylor  + analytical 2-point momentum-impulse",
    " 1st order Taylor  + d_theta/d_s Runge-Kutta (AnDF)",
    " 2-point curvature + d_theta/d_s Runge-Kutta",
    " 2-point curvature + analytical 2-point momentum-impulse",
    " 2-point curvature + optimized  2-point momentum-impulse",
    " 2nd order Taylor  + d_theta/d_s Runge-Kutta (HySA)",
    " 2nd order Taylor  + 4-point difference method (MxSA)",
    " 2nd order Taylor  + analytical 2-point momentum-impulse",
    " 2nd order Taylor  + optimized  2-point momentum-impulse",
    " 2-point curvature + optimized 2-point momentum-impulse",
    " 2nd order Taylor  + optimized 2-point momentum-impulse",
]


def menu_to_method(uc, choice):
    """Menu answer -> op number: isotropic scenarios 1..9 -> op1..op9 (RT:1238-1263); anisotropy 1, 2 -> op10, op11
    (RT:1286-1290)."""
    choice = int(choice)
    if uc == "4":
        if choice not in (1, 2):
            raise ValueError("anisotropy offers algorithms 1 and 2 (op10, op11), got %d" % choice)
        return 9 + choice
    if not 1 <= choice <= 9:
        raise ValueError("isotropic scenarios offer algorithms 1..9, got %d" % choice)
    return choice


def results_metric(uc, s_ray, errors):
    """The number the RESULTS block prints (RT:1391-1402)."""
    if uc == "2":
        return 100 * np.linalg.norm(np.array([1, 0]) - s_ray[-1, 0:2, 0]) / (2 * np.pi)
    if uc == "1":
        return np.mean(errors)
    R = s_ray.shape[2]
    return np.mean([px_cv(s_ray[:, 2, i]) for i in range(1, R - 1)])


def print_results(uc, metric, out=None):
    out = out or sys.stdout
    print("\nRESULTS", file=out)
    if uc == "2":
        print("Closure error ", metric, "%", file=out)
    elif uc == "1":
        print("Average ray error: ", metric, "degrees", file=out)
    else:
        print("Average ray Coefficient of Variation: ", metric, file=out)


def run(scenario, method_choice, delta_s="default", benchmark=False, trials=100, show=True, out=None, device=0,
        fit="device"):
    """One pass through the reference's main(): returns a dict with everything printed."""
    out = out or sys.stdout
    uc = SCENARIOS.get(str(scenario), str(scenario))
    uc = K.scenario_id(uc)
    m = menu_to_method(uc, method_choice)
    eng = get_engine(device)
    z, grd = build_field_gpu(uc, eng) if fit == "device" else build_field(uc)    # interpolacion, RT:1588
    step, divisor = K.DELTA_S, K.DELTA_S_DIVISOR_FISHEYE
    info = dict(scenario=uc, method=m, algorithm=MESSAGES[m - 1].strip())
    if delta_s == "search":                                   # RT:1299-1406
        res = find_delta_s(m, z, grd, uc, engine=eng)
        if res["index"] is None:
            print("\nNo suitable divisor was found. Try using another search interval (*_UPPER_LIMIT, *_LOWER_LIMIT). "
                  "Exiting...", file=out)
            info["error"] = "no suitable divisor"
            return info
        step = res["DELTA_S"]
        if uc == "2":
            divisor = res["divisor"]
        print("\nSelected DELTA_S divisor: ", res["divisor"], file=out)
        info["divisor"] = res["divisor"]
    elif delta_s == "calibrated":                             # RT:1408-1456
        step, divisor = K.calibrated_step(uc, m)
    elif delta_s != "default":
        raise ValueError("--delta-s must be default, calibrated or search")
    t1 = time.perf_counter()
    s_ray, d_ray, compute_times, errors = trazar(m, z, grd, show, step, divisor + 1, uc, engine=eng)
    t2 = time.perf_counter()
    metric = results_metric(uc, s_ray, errors)
    print_results(uc, metric, out)
    print("Total travelled distance: ", np.sum(d_ray[1, :]), file=out)                     # RT:1486
    info.update(DELTA_S=float(step), metric=float(metric), total_distance=float(np.sum(d_ray[1, :])),
                ray_steps=int(d_ray[2].sum()), wall_s=t2 - t1, kernel_s=float(np.sum(compute_times)))
    if benchmark:                                             # RT:1502-1541 with GPU kernel time as the sample
        def sample():
            return float(np.sum(trazar(m, z, grd, False, step, divisor + 1, uc, engine=eng, full=False)[2]))
        final = benchmark_statistic(sample, trial=trials)
        print("\nCompletion time per scenario:", final, "seconds.", file=out)
        info["completion_time_s"] = final
    return info


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m raytracing_b200", description=__doc__.split("\n\n")[0])
    ap.add_argument("--scenario", required=True, help="1-4 or interface | fisheye | vert_heterogeneous | anisotropy")
    ap.add_argument("--method", required=True, type=int, help="the algorithm menu answer: 1-9 (1-2 for anisotropy)")
    ap.add_argument("--delta-s", default="default", choices=["default", "calibrated", "search"])
    ap.add_argument("--benchmark", action="store_true")
    ap.add_argument("--trials", type=int, default=100)
    ap.add_argument("--quiet", action="store_true", help="do not print the per-ray Snell table the reference shows for "
                                                         "the interface scenario (RT:1387, 1470)")
    ap.add_argument("--json", action="store_true", help="also print the results as one JSON line")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--fit", default="device", choices=["device", "scipy"],
                    help="where interpolacion() runs: rtb_fit_field on the GPU, or scipy on the host as the reference")
    a = ap.parse_args(argv)
    info = run(a.scenario, a.method, a.delta_s, a.benchmark, a.trials, not a.quiet, device=a.device, fit=a.fit)
    if a.json:
        print(json.dumps(info))
    return 0 if "error" not in info else 1


if __name__ == "__main__":
    sys.exit(main())
