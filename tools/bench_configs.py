#!/usr/bin/env python
"""BASELINE.json configs[2..4] as bench lines (bench.py --workload vert|aniso|shoot delegates here).

  vert  : 'vert_heterogeneous', 2^26 rays op6, sharded (block-cyclic) over the N GPUs, FINAL output +
          online p_x statistics (the CV metric of RT_bench.py:1354-1360); strong scaling
  aniso : 'anisotropy' gamma=3, 2^24 rays, op10 and op11, same sharding; strong scaling
  shoot : two-point shooting search + DELTA_S sweep on 'interface': 4096 targets x 4096 launch angles x
          the 200 sweep candidates of RT_bench.py:1303-1304; coarse fan sharded by angle, MIN/argmin
          all-reduce (the path's only real exchange step), refinement sharded by target; strong scaling

Same JSON line shape as bench.py (one line per workload/method); launched by torchrun for N > 1.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

METRIC, UNIT = "fp64_ray_steps_per_s", "ray-steps/s"


def _dist_setup():
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    return world, rank, local, dist


def _max_sum(vals, world, dist, dev):
    import torch
    t = torch.tensor(vals, dtype=torch.float64, device=dev)
    if world == 1:
        return t.tolist(), t.tolist()
    mx, sm = t.clone(), t.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    return mx.tolist(), sm.tolist()


def run_trace_workload(name, args, rays_log2=None):
    import torch
    import raytracing_b200 as rt
    from raytracing_b200 import constants as K
    from raytracing_b200.dist import cyclic_indices
    from raytracing_b200.fans import fan_at
    from bench import ClockSampler
    world, rank, local, dist = _dist_setup()
    dev = torch.device("cuda", local)
    uc, methods, lg = {"vert": ("3", [6], 26), "aniso": ("4", [11, 10], 24)}[name]
    lg = rays_log2 or lg
    R = 1 << lg
    eng = rt.Engine(local)
    eng.use_torch_stream()
    z, grd = rt.build_field(uc)
    eng.set_field(z, grd)
    c = rt.scenario_constants(uc)
    gamma, box = c[0], c[5:9]
    step = rt.DELTA_S
    M = rt.max_size_for(uc, step, 91)
    idx = cyclic_indices(R, world, rank)
    th, x0, y0 = (torch.from_numpy(a).to(dev) for a in fan_at(uc, R, idx))
    Rl = idx.size
    out = dict(d_ray=torch.empty(3, Rl, dtype=torch.float64, device=dev),
               last_row=torch.empty(6, Rl, dtype=torch.float64, device=dev),
               px_stats=torch.empty(3, Rl, dtype=torch.float64, device=dev))
    peak = eng.fp64_peak()
    for m in methods:
        def one_step():
            return eng.trace(m, th, x0, y0, step, M, gamma, box, full=False, want_px_stats=True, out=out)
        for _ in range(args.warmup):
            one_step()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ksum, steps_done, launches = 0.0, 0, 0
        for _ in range(args.steps):
            r = one_step()
            ksum += r.stats["kernel_ms"]; steps_done += r.stats["ray_steps"]; launches += r.stats["launches"]
        e1.record()
        torch.cuda.synchronize()
        t_ms = e0.elapsed_time(e1)
        if world > 1:
            dist.barrier()
        clocks = sampler.stop() if rank == 0 else None
        st = out["px_stats"]
        # scenario_average_CV RT:1354-1360 (the reference skips its first and last ray, theta = 0 and pi/2, where
        # p_x ~ 0 makes the ratio meaningless; here every ray with |mean p_x| > 1e-6 counts)
        okr = st[1].abs() > 1e-6
        cv = torch.where(okr, 100 * st[2] / st[1], torch.zeros_like(st[1]))
        cv_max = float((cv.sum() / okr.sum().clamp(min=1)).item())
        (t_max, k_max, _, cvm, _), (_, _, total_steps, _, launches) = _max_sum(
            [t_ms, ksum, float(steps_done), cv_max, float(launches)], world, dist, dev)
        if rank == 0:
            value = total_steps / (t_max * 1e-3)
            ach = (total_steps / world) * K.F_ALG[m] / (k_max * 1e-3) / 1e12
            print(json.dumps({
                "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": t_max / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "%s, 2^%d rays total, op%d, DELTA_S=SIGMA/20, max_size=%d, FINAL output + online "
                                       "p_x statistics" % ({"3": "vert_heterogeneous", "4": "anisotropy (gamma=3)"}[uc], lg, m, M),
                           "rays_total": R, "rays_per_gpu": Rl, "ray_steps_per_step": int(total_steps // args.steps),
                           "l2": "inputs+outputs %.1f GB per GPU > L2" % (15 * 8 * Rl / 1e9),
                           "parallelism": "block-cyclic ray shards x%d, no data-path collective" % world,
                           "px_cv_mean_percent_max_over_ranks": cvm},
                "roofline": {"bound": "fp64", "achieved": ach, "peak": peak / 1e12, "unit": "TFLOP/s",
                             "frac": ach / (peak / 1e12), "traffic": None, "flop_per_ray_step": K.F_ALG[m],
                             "kernel": "trace_kernel<%d,%s,final,extra>" % (m, "iso" if gamma == 1 else "aniso")},
                "gpu_launches": int(launches), "clocks": clocks}), flush=True)


def run_shoot_workload(args, n_cand=200, A=4096, Tn=4096, method=6):
    import torch
    import raytracing_b200 as rt
    from raytracing_b200 import constants as K
    from bench import ClockSampler
    world, rank, local, dist = _dist_setup()
    dev = torch.device("cuda", local)
    uc = "1"
    eng = rt.Engine(local)
    eng.use_torch_stream()
    z, grd = rt.build_field(uc)
    eng.set_field(z, grd)
    divisors, options, _ = rt.sweep_candidates(uc)                 # RT:1303-1304: 200 candidates
    sel = np.linspace(0, len(options) - 1, n_cand).astype(int)
    steps = options[sel]
    angles = np.linspace(2 * np.pi / 60, np.pi / 2, A)              # the scenario's fan range RT:259-260
    true = np.linspace(0.2, 1.5, Tn) + 1.3e-4
    tx, ty, T_true = rt.targets_from_angles(eng, method, uc, true, steps[0])
    grp = True if world > 1 else None

    def one_step():
        return rt.shooting_search(method, z, grd, uc, (tx, ty), angles, steps=steps, engine=eng, group=grp,
                                  device_arrays=True)
    for _ in range(args.warmup):
        r = one_step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ksum, steps_done, launches = 0.0, 0, 0
    for _ in range(args.steps):
        r = one_step()
        ksum += r["stats"]["kernel_ms"]; steps_done += r["stats"]["ray_steps"]; launches += r["stats"]["launches"]
    e1.record()
    torch.cuda.synchronize()
    t_ms = e0.elapsed_time(e1)
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    (t_max, k_max, _, _), (_, _, total_steps, total_launches) = _max_sum([t_ms, ksum, float(steps_done), float(launches)],
                                                                        world, dist, dev)
    if rank == 0:
        th = r["theta"].cpu().numpy() if hasattr(r["theta"], "cpu") else np.asarray(r["theta"])
        miss = r["miss"].cpu().numpy() if hasattr(r["miss"], "cpu") else np.asarray(r["miss"])
        ok = np.abs(true - np.pi / 4) > 0.12
        peak = eng.fp64_peak()
        value = total_steps / (t_max * 1e-3)
        ach = (total_steps / world) * K.F_ALG[method] / (k_max * 1e-3) / 1e12
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": t_max / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "two-point shooting search + DELTA_S sweep, interface, %d targets x %d launch angles x %d "
                                   "DELTA_S candidates, op%d, golden refinement (%d iterations)" % (Tn, A, n_cand, method, r["iterations"]),
                       "ray_steps_per_step": int(total_steps // args.steps),
                       "parallelism": ("coarse fan sharded by angle + NCCL MIN/argmin all-reduce, refinement sharded by "
                                       "target, result all_gather, x%d" % world) if world > 1 else "1 GPU",
                       "l2": "per-step buffers (%.0f MB) rewritten every golden iteration" % (2 * n_cand * Tn * 12 * 8 / 1e6),
                       "check": {"max_abs_theta_minus_truth_finest": float(np.abs(th[0] - true)[ok].max()),
                                 "max_miss_finest": float(miss[0][ok].max())}},
            "roofline": {"bound": "fp64", "achieved": ach, "peak": peak / 1e12, "unit": "TFLOP/s", "frac": ach / (peak / 1e12),
                         "traffic": None, "flop_per_ray_step": K.F_ALG[method], "kernel": "trace_kernel<%d,iso,final,extra>" % method},
            "gpu_launches": int(total_launches), "clocks": clocks}), flush=True)


def main(argv=None):
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", required=True, choices=["vert", "aniso", "shoot"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--rays-log2", type=int, default=None, help="override the fan size (vert/aniso)")
    ap.add_argument("--candidates", type=int, default=200)
    args = ap.parse_args(argv)
    if args.workload == "shoot":
        run_shoot_workload(args, n_cand=args.candidates)
    else:
        run_trace_workload(args.workload, args, args.rays_log2)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
