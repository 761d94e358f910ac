#!/usr/bin/env python
"""bench.py -- FP64 ray-steps/s of the RT_bench.py hot path on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --steps K --warmup W    # CPU arm: the oracle port on host cores

ONE JSON line.  The headline (`value`, `e2e`, `roofline`, ...) is BASELINE.json configs[1]: 'fisheye', 2^20 rays per
GPU, FP64, method op6 (HySA), default DELTA_S = SIGMA/20, max_size = N*(90+1) = 910 (909 ray-steps per ray); at N>1
every rank traces its own 2^20-ray shard of an N*2^20 fan (weak scaling, no data-path collective; the d_ray result
gather is inside the timed step).  A "step" is one pass of the hot path (one trazar) over the whole synthetic fan.

The same line carries `configs`: the other four BASELINE configs, each with its own value / ms_per_step / roofline /
e2e (and, at N>1, its result gather inside the timed step):
  interface_default     configs[0]  'interface', default fan (42 rays), op6, default DELTA_S, FULL s_ray + Snell errors
  vert_64M              configs[2]  'vert_heterogeneous', 2^26 rays op6, block-cyclic shards over the N GPUs (STRONG
                                    scaling), FINAL output + online p_x statistics (the CV metric of RT:1354-1360)
  aniso_16M_op11/op10   configs[3]  'anisotropy' gamma=3, 2^24 rays, both methods of the menu (RT:1268-1294); strong
  shoot_4096x4096x200   configs[4]  two-point shooting search + DELTA_S sweep on 'interface': 4096 targets x 4096 launch
                                    angles x the 200 sweep candidates of RT:1303-1304; coarse fan sharded by angle,
                                    NCCL MIN/argmin all-reduce (the path's only real exchange step), refinement
                                    sharded by target; strong
The secondary configs run a bounded number of steps (1 warm-up + at most 2 timed) so the default run stays within
minutes; `--configs headline` skips them.  `--impl reference` gives every config its CPU leg (oracle port, all host
threads, bounded samples).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCENARIO = "2"
METHOD = 6
RAYS_PER_GPU = 1 << 20
METRIC = "fp64_ray_steps_per_s"
UNIT = "ray-steps/s"
# FP64 pipe of one B200: 148 SMs x 4 sub-partitions, one warp-wide DFMA (64 flop) per 2 cycles each
SMS, SMSP, FP64_CYCLES_PER_WARP_INST = 148, 4, 2.0


def workload_name(n_gpus):
    return ("fisheye (Maxwell fisheye), %d x 2^20 rays, op%d (HySA), DELTA_S=SIGMA/20, max_size=910, "
            "FP64, FINAL output (d_ray + last row)" % (n_gpus, METHOD))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(len(r) >= 6 and r[2 + j] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def host_cores():
    """All host threads this process may use (torchrun exports OMP_NUM_THREADS=1, which the oracle's
    explicit omp_set_num_threads overrides)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ----------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (C port of the reference algorithm, OpenMP over rays) on bounded samples of the same workloads.
# The one place besides tests/ and smoke() that executes oracle/ -- as the baseline, never as the product.
# ----------------------------------------------------------------------------------------------------------------
def _cpu_trace(uc, method, th, x0, y0, nthreads, full=False, interface=False):
    import raytracing_b200 as rt
    from oracle import rt_oracle as O
    z, grd = rt.build_field(uc)
    c = rt.scenario_constants(uc)
    M = rt.max_size_for(uc, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1)
    zs, g0, g1 = (O.HostSpline.from_scipy(s) for s in (z, grd[0], grd[1]))
    O.trace(method, zs, (g0, g1), rt.DELTA_S, M, c[0], c[5:9], interface, th[:min(64, len(th))], x0[:64], y0[:64],
            full=False, nthreads=nthreads)                                  # warm the threads and the tables
    t0 = time.perf_counter()
    r = O.trace(method, zs, (g0, g1), rt.DELTA_S, M, c[0], c[5:9], interface, th, x0, y0, full=full, nthreads=nthreads)
    return r["ray_steps"], time.perf_counter() - t0


def cpu_port_rate(sample_rays, nthreads):
    """Headline workload: a strided slice of the 2^20-ray fisheye fan."""
    import raytracing_b200 as rt
    stride = RAYS_PER_GPU // sample_rays
    th, x0, y0 = rt.make_fan(SCENARIO, RAYS_PER_GPU)
    return _cpu_trace(SCENARIO, METHOD, th[::stride][:sample_rays], x0[:sample_rays], y0[:sample_rays], nthreads)


def cpu_config_legs(cores):
    """CPU legs of the secondary configs: {name: {value, unit, cores, kind, sample}}."""
    import raytracing_b200 as rt
    out = {}

    def entry(steps, dt, sample):
        return {"value": steps / dt, "unit": UNIT, "cores": cores, "kind": "port", "seconds": dt, "sample": sample}
    th, x0, y0 = rt.default_fan("1")
    s, dt = _cpu_trace("1", 6, th, x0, y0, cores, full=True, interface=True)
    out["interface_default"] = entry(s, dt, "the reference's own case as is: default fan (42 rays), op6, full s_ray + Snell "
                                            "errors (%d ray-steps), oracle C port, OpenMP over rays" % s)
    for name, uc, lg, m, n in (("vert_64M", "3", 26, 6, 4096), ("aniso_16M_op11", "4", 24, 11, 512),
                               ("aniso_16M_op10", "4", 24, 10, 512)):
        th, x0, y0 = rt.make_fan(uc, 1 << lg)
        st = (1 << lg) // n
        s, dt = _cpu_trace(uc, m, th[::st][:n], x0[:n], y0[:n], cores)
        out[name] = entry(s, dt, "%d-ray strided slice of the 2^%d-ray fan (%d ray-steps), op%d" % (n, lg, s, m))
    out["shoot_4096x4096x200"] = entry(*_cpu_shoot_sample(cores))
    return out


def _cpu_shoot_sample(cores):
    """Two-point search (oracle/shoot_oracle.py: rays by the C port, golden literally) on a bounded sub-problem."""
    import raytracing_b200 as rt
    from oracle import shoot_oracle as S
    uc, m = "1", 6
    z, grd = rt.build_field(uc)
    c = rt.scenario_constants(uc)
    _div, options, _ = rt.sweep_candidates(uc)
    steps = [options[0], options[-1]]
    angles = np.linspace(2 * np.pi / 60, np.pi / 2, 64)
    true = np.linspace(0.3, 1.4, 8) + 1.3e-4
    total, t0 = 0, time.perf_counter()
    for st in steps:
        M = int(rt.candidate_rows(uc, [st])[0])
        tr = S.Tracer(m, z, grd, st, M, c[0], c[5:9], (-2.0, -2.0))
        counted = {"n": 0}
        def traced(thetas, _tr=tr, _c=counted):
            e, ns = _tr(thetas)
            _c["n"] += int(ns.sum())
            return e, ns
        e_true, _ = traced(true)
        S.shooting_search(traced, (e_true[0], e_true[1]), angles)
        total += counted["n"]
    dt = time.perf_counter() - t0
    return total, dt, ("2 DELTA_S candidates (finest, coarsest) x 64 launch angles x 8 targets (%d ray-steps): "
                       "oracle/shoot_oracle.py, rays by the C port (8 threads per call), golden in Python" % total)


def run_reference(args):
    """--impl reference: the reference's algorithm on the host cores (oracle port; the reference itself
    is a Python script that cannot travel to the GPU box)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    sample = 4096
    for _ in range(args.warmup):
        cpu_port_rate(256, cores)
    steps_total, t_total = 0, 0.0
    for _ in range(args.steps):
        s, dt = cpu_port_rate(sample, cores)
        steps_total += s
        t_total += dt
    v = steps_total / t_total
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.gpus), "note": "CPU arm runs a bounded sample per step"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "%d-ray strided slice of the 2^20-ray fan per step (%d ray-steps), "
                                       "oracle/rt_oracle.c with OpenMP over rays" % (sample, steps_total // args.steps)},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if args.configs == "all":
        line["configs"] = cpu_config_legs(cores)
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------------------
# CUDA arm
# ----------------------------------------------------------------------------------------------------------------
class Dist:
    def __init__(self):
        import torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def max_sum(self, vals):
        import torch
        t = torch.tensor(vals, dtype=torch.float64, device=self.dev)
        if self.world == 1:
            return t.tolist(), t.tolist()
        mx, sm = t.clone(), t.clone()
        self.dist.all_reduce(mx, op=self.dist.ReduceOp.MAX)
        self.dist.all_reduce(sm, op=self.dist.ReduceOp.SUM)
        return mx.tolist(), sm.tolist()


def load_ncu_latest(kernel_id):
    """Executed-instruction counts of the kernels, from the ncu captures tools/ncu_summary.py json condensed into
    profiles/ncu_latest.json.  They describe one version of the trace kernels: used only when its kernel id (digest of
    the kernel sources, rtb_kernel_id) is the loaded library's, so the numbers are traceable to a capture of THIS code."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "ncu_latest.json")))
    except (OSError, ValueError):
        return None, "profiles/ncu_latest.json missing"
    if d.get("kernel_id") != kernel_id:
        return None, "profiles/ncu_latest.json is of kernels %s, the library's are %s" % (d.get("kernel_id"), kernel_id)
    return d, None


def roofline_block(method, kernel_name, per_gpu_steps, kernel_ms, fp64_peak, clocks_mhz, ncu, ncu_note, ncu_key, F_ALG):
    """`frac` is the contract's figure: algorithmic flop (SURVEY.md 8(d) F_alg: B-spline-form field evaluation, 25 flop
    per sin/cos) x ray-steps / time / measured DFMA-stream peak -- it exceeds 1 when the kernel executes fewer flop than
    the algorithmic count.  `pipe_frac` is the hardware's view: executed FP64 warp-instructions x 2 issue cycles /
    sub-partition cycles of the launch."""
    ach = per_gpu_steps * F_ALG[method] / (kernel_ms * 1e-3) / 1e12
    nominal = SMS * SMSP * 64 / FP64_CYCLES_PER_WARP_INST * (clocks_mhz or 1965.0) * 1e6 / 1e12
    rb = {"bound": "fp64", "achieved": ach, "peak": fp64_peak / 1e12, "unit": "TFLOP/s", "frac": ach / (fp64_peak / 1e12),
          "frac_is": "algorithmic flop (F_alg = %d per ray-step, SURVEY.md 8(d)) / measured DFMA-stream peak; > 1 means the "
                     "kernel executes fewer flop than the algorithmic count -- see pipe_frac" % F_ALG[method],
          "peak_source": "measured live: rtb_fp64_peak (DFMA stream on all SMs; MEASURED_PEAKS.json has no FP64 figure); "
                         "nominal pipe %.1f TFLOP/s at %.0f MHz" % (nominal, clocks_mhz or 1965.0),
          "peak_nominal": nominal, "kernel": kernel_name, "kernel_ms": kernel_ms, "flop_per_ray_step": F_ALG[method],
          "pipe_frac": None, "traffic": None, "executed": None}
    k = (ncu or {}).get("kernels", {}).get(ncu_key)
    if k is None:
        rb["executed_note"] = ncu_note or ("no ncu capture of %s in profiles/ncu_latest.json" % ncu_key)
        return rb
    inst = k["dfma"] + k["dmul"] + k["dadd"] + k.get("dsetp", 0.0)
    flop = 2 * k["dfma"] + k["dmul"] + k["dadd"]
    cycles = kernel_ms * 1e-3 * (clocks_mhz or 1965.0) * 1e6
    rb["pipe_frac"] = (per_gpu_steps / 32.0) * inst * FP64_CYCLES_PER_WARP_INST / (SMS * SMSP * cycles)
    rb["traffic"] = k.get("dram_bytes_per_launch")
    rb["traffic_unit"] = "DRAM bytes per launch of the ncu capture (dram__bytes_read.sum + dram__bytes_write.sum)"
    rb["executed"] = {"fp64_inst_per_ray_step": inst, "dfma": k["dfma"], "dmul": k["dmul"], "dadd": k["dadd"],
                      "flop_per_ray_step": flop, "tflops": per_gpu_steps * flop / (kernel_ms * 1e-3) / 1e12,
                      "inst_per_warp_step": k.get("inst_per_warp_step"),
                      "ncu_fp64_pipe_active_frac": k.get("fp64_pipe_active_frac"),
                      "ncu_l1_data_pipe_frac": k.get("lsu_pipe_frac"), "source": k.get("source"),
                      "captured_on": k.get("workload")}
    return rb


def run_cuda(args):
    import torch
    import raytracing_b200 as rt
    from raytracing_b200 import constants as K

    D = Dist()
    world, rank, local, dist, dev = D.world, D.rank, D.local, D.dist, D.dev
    eng = rt.Engine(local)
    eng.use_torch_stream()
    z, grd = rt.build_field(SCENARIO)
    eng.set_field(z, grd)
    c = rt.scenario_constants(SCENARIO)
    gamma, box = c[0], c[5:9]
    step = rt.DELTA_S
    M = rt.max_size_for(SCENARIO, step, rt.DELTA_S_DIVISOR_FISHEYE + 1)
    R = RAYS_PER_GPU
    th_h, x0_h, y0_h = rt.make_fan(SCENARIO, R * world, rank * R, (rank + 1) * R)  # this rank's shard

    # device-resident inputs / preallocated outputs for `value`
    th, x0, y0 = (torch.from_numpy(a).to(dev) for a in (th_h, x0_h, y0_h))
    out = dict(d_ray=torch.empty(3, R, dtype=torch.float64, device=dev),
               last_row=torch.empty(6, R, dtype=torch.float64, device=dev))
    gathered = torch.empty(world * 3 * R, dtype=torch.float64, device=dev) if world > 1 else None
    fused = None
    if world > 1 and args.gather == "fused":   # d_ray written into every GPU's array by the kernel itself (dist.FusedGather)
        from raytracing_b200.dist import FusedGather
        try:
            fused = FusedGather(eng, R * world, rank * R)
        except RuntimeError as e:   # raised on every rank alike: fall back to the NCCL all-gather
            if rank == 0:
                print("bench: fused gather unavailable (%s), using NCCL all_gather" % e, file=sys.stderr)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    last_view = [None]

    def one_step():
        r = eng.trace(METHOD, th, x0, y0, step, M, gamma, box, full=False, out=out)
        if fused is not None:  # peers were written by the kernel; ranks only agree that they are done
            last_view[0] = fused.finish()
        elif world > 1:  # result gather as a separate collective (NCCL over NVLink)
            dist.all_gather_into_tensor(gathered, out["d_ray"].view(-1))
        return r

    launches = 0
    for _ in range(max(args.warmup, 3)):
        one_step()
    fp64_peak = eng.fp64_peak()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_ms, ray_steps = [], 0
    D.barrier()
    torch.cuda.synchronize()
    for k in range(args.steps):
        flush.zero_()  # L2 flush between timed iterations (outside the event pair)
        ev[k][0].record()
        r = one_step()
        ev[k][1].record()
        kernel_ms.append(r.stats["kernel_ms"])
        ray_steps += r.stats["ray_steps"]
        launches += r.stats["launches"]
    torch.cuda.synchronize()
    D.barrier()
    t_ms = sum(a.elapsed_time(b) for a, b in ev)
    # the event pairs must bracket the kernels (same stream): a step cannot be shorter than its kernel
    if t_ms < 0.99 * sum(kernel_ms):
        raise RuntimeError("timing events (%.3f ms) do not bracket the kernels (%.3f ms): stream mismatch"
                           % (t_ms, sum(kernel_ms)))
    clocks = sampler.stop() if rank == 0 else None
    gather_mode, gather_check = None, None
    if world > 1:
        gather_mode = ("peer stores from the kernel epilogue over NVLink (CUDA IPC, double-buffered) + 1-element all-reduce"
                       if fused is not None else "NCCL all_gather after the kernel")
        # the gathered array must be what an all-gather of every rank's d_ray gives, bit for bit
        ref = torch.empty(world * 3 * R, dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(ref, out["d_ray"].contiguous().view(-1))
        ref = ref.view(world, 3, R).permute(1, 0, 2).reshape(3, world * R)
        got = last_view[0] if fused is not None else gathered.view(world, 3, R).permute(1, 0, 2).reshape(3, world * R)
        okf = torch.tensor([1.0 if torch.equal(got, ref) else 0.0], dtype=torch.float64, device=dev)
        dist.all_reduce(okf, op=dist.ReduceOp.MIN)
        gather_check = "ok: every rank's (3, %d) array == all_gather(d_ray), bit for bit" % (world * R) if okf.item() == 1.0 \
            else "MISMATCH between the gathered array and all_gather(d_ray)"
    if fused is not None:   # the remaining legs use the same engine without peers
        fused.close()
        fused = None

    # end to end through the public host-buffer API: pinned host inputs, H2D + D2H inside the call
    pin = lambda a: torch.from_numpy(a).pin_memory().numpy()
    th_p, x0_p, y0_p = pin(th_h), pin(x0_h), pin(y0_h)
    out_h = dict(d_ray=torch.empty(3, R, dtype=torch.float64).pin_memory().numpy(),
                 last_row=torch.empty(6, R, dtype=torch.float64).pin_memory().numpy())
    for _ in range(2):
        eng.trace(METHOD, th_p, x0_p, y0_p, step, M, gamma, box, full=False, out=out_h)
    D.barrier()
    torch.cuda.synchronize()
    e2e_t0 = time.perf_counter()
    e2e_steps, h2d, d2h = 0, 0, 0
    for _ in range(args.steps):
        r2 = eng.trace(METHOD, th_p, x0_p, y0_p, step, M, gamma, box, full=False, out=out_h)
        e2e_steps += r2.stats["ray_steps"]
        h2d, d2h = r2.stats["h2d_bytes"], r2.stats["d2h_bytes"]
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - e2e_t0) * 1e3

    # FULL-trajectory mode (the reference's (max_size,6,R) s_ray, 45.8 GB): HBM-write roofline
    full_info, trazar_full = None, None
    if rank == 0 and not args.no_full:
        s_ray = torch.empty(M, 6, R, dtype=torch.float64, device=dev)
        best = None
        for _ in range(3):
            rf = eng.trace(METHOD, th, x0, y0, step, M, gamma, box, full=True, out=dict(s_ray=s_ray, **out))
            if best is None or rf.stats["kernel_ms"] < best.stats["kernel_ms"]:
                best = rf
        gbs = best.stats["ray_steps"] * 48 / (best.stats["kernel_ms"] * 1e-3) / 1e9
        full_info = {"ms": best.stats["kernel_ms"], "ray_steps_per_s": best.stats["ray_steps"] / (best.stats["kernel_ms"] * 1e-3),
                     "hbm_write_gbs": gbs, "kernel": "trace_kernel<6,iso,traj=per-lane st.global.cs>"}
        del s_ray
        torch.cuda.empty_cache()
        trazar_full = e2e_trazar_full(rt, eng, z, grd, step)

    # max over ranks of the device time; sum of the work
    (t_ms, e2e_ms, _, _, _), (_, _, ray_steps, e2e_steps, launches) = D.max_sum(
        [t_ms, e2e_ms, float(ray_steps), float(e2e_steps), float(launches)])
    ray_steps, e2e_steps, launches = int(ray_steps), int(e2e_steps), int(launches)

    ncu, ncu_note = load_ncu_latest(eng.kernel_id())
    ctx = dict(D=D, eng=eng, rt=rt, K=K, args=args, fp64_peak=fp64_peak, ncu=ncu, ncu_note=ncu_note)
    configs = run_secondary_configs(ctx) if args.configs == "all" else None

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        value = ray_steps / (t_ms * 1e-3)
        k_ms = float(np.mean(kernel_ms))
        per_gpu_steps = ray_steps / world / args.steps
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": t_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(world), "scenario": "fisheye", "method": "op6",
                           "rays_per_gpu": R, "max_size": M, "ray_steps_per_step": ray_steps // args.steps,
                           "l2": "flushed (256 MiB write) between timed steps",
                           "parallelism": ("rays sharded x%d, d_ray gathered on every GPU: %s" % (world, gather_mode))
                           if world > 1 else "1 GPU", "build_id": eng.build_id(), "kernel_id": eng.kernel_id()},
                "roofline": roofline_block(METHOD, "trace_kernel<6,iso,final>", per_gpu_steps, k_ms, fp64_peak,
                                           (clocks or {}).get("sm_mhz"), ncu, ncu_note, "op6_final_fisheye", K.F_ALG),
                "e2e": {"value": e2e_steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / args.steps,
                        "api": "Engine.trace with pinned host numpy buffers (RTB_PTR_HOST): rtb_trace pipelines the fan in 8 chunks "
                               "(H2D copy, kernel, D2H copy on three streams); FINAL output (d_ray + last row)"},
                "gpu_launches": launches, "clocks": clocks}
        if gather_check:
            line["gather_check"] = gather_check
        if full_info:
            full_info.update({"hbm_peak_gbs": hbm_peak, "hbm_frac": full_info["hbm_write_gbs"] / hbm_peak,
                              "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"})
            line["full_trajectory_mode"] = full_info
        if trazar_full:
            line["e2e_trazar_full"] = trazar_full
        if configs is not None:
            line["configs"] = configs
        if world == 1 and not args.no_cpu:
            cores = host_cores()
            s, dt = cpu_port_rate(4096, cores)
            line["cpu_baseline"] = {"value": s / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "4096-ray strided slice of the same fan (%d ray-steps), oracle C port, "
                                              "OpenMP over rays" % s}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def e2e_trazar_full(rt, eng, z, grd, step):
    """The drop-in rt.trazar() exactly as the reference is called (RT:766-948): pageable numpy in, the FULL
    (max_size, 6, R) s_ray out.  45.8 GB for the 2^20-ray fan -- PCIe-bound by construction (48 B per ray-step must
    cross the bus); FINAL / STRIDED outputs are the answer to that, this is the number for the reference's own
    return value.  The fan is the largest power of two whose s_ray fits a quarter of the free host memory (<= 2^20)."""
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 64 << 30
    M = rt.max_size_for(SCENARIO, step, rt.DELTA_S_DIVISOR_FISHEYE + 1)
    lg = 20
    while lg > 14 and M * 6 * 8 * (1 << lg) > min(avail // 4, 48 << 30):
        lg -= 1
    R = 1 << lg
    fan = rt.make_fan(SCENARIO, RAYS_PER_GPU)
    st = RAYS_PER_GPU // R
    fan = tuple(np.ascontiguousarray(a[::st][:R]) for a in fan)
    rt.trazar(METHOD, z, grd, False, step, rt.DELTA_S_DIVISOR_FISHEYE + 1, SCENARIO, fan=fan, engine=eng, full=False)
    t0 = time.perf_counter()
    s_ray, d_ray, _ct, _err = rt.trazar(METHOD, z, grd, False, step, rt.DELTA_S_DIVISOR_FISHEYE + 1, SCENARIO, fan=fan,
                                        engine=eng)
    dt = time.perf_counter() - t0
    steps = int(d_ray[2].sum())
    return {"value": steps / dt, "unit": UNIT, "seconds": dt, "rays": R, "s_ray_bytes": int(s_ray.nbytes),
            "d2h_gbs": s_ray.nbytes / dt / 1e9,
            "api": "raytracing_b200.trazar(op6, z, grd, False, DELTA_S, 91, '2', fan=...): the reference's call and return "
                   "values, pageable numpy, full s_ray (the device array is allocated, written and copied back inside "
                   "the call)"}


# ----------------------------------------------------------------------------------------------------------------
# BASELINE.json configs[0], [2], [3], [4]
# ----------------------------------------------------------------------------------------------------------------
def _timed(D, fn, steps, warmup):
    """warmup + `steps` timed calls of fn() between CUDA events on the current stream (barrier + synchronize on both
    sides, max over ranks taken by the caller).  Returns (ms total, list of fn results)."""
    import torch
    for _ in range(warmup):
        fn()
    D.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    rs = [fn() for _ in range(steps)]
    e1.record()
    torch.cuda.synchronize()
    D.barrier()
    return e0.elapsed_time(e1), rs


def _wall(D, fn, steps, warmup):
    import torch
    for _ in range(warmup):
        fn()
    D.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    rs = [fn() for _ in range(steps)]
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3, rs


def run_trace_config(ctx, name, uc, lg, m, ncu_key):
    """vert / aniso: a 2^lg-ray fan dealt block-cyclically over the ranks (strong scaling), FINAL output + online p_x
    statistics; at N>1 every GPU also receives every ray's d_ray (fused peer stores, block-cyclic index map) inside
    the timed step."""
    import torch
    D, eng, rt, K, args = ctx["D"], ctx["eng"], ctx["rt"], ctx["K"], ctx["args"]
    from raytracing_b200.dist import FusedGather, cyclic_indices
    from raytracing_b200.fans import fan_at
    world, rank, dev = D.world, D.rank, D.dev
    BLOCK = 4096
    R = 1 << lg
    eng.set_field(*rt.build_field(uc))
    c = rt.scenario_constants(uc)
    gamma, box = c[0], c[5:9]
    step = rt.DELTA_S
    M = rt.max_size_for(uc, step, 91)
    idx = cyclic_indices(R, world, rank, BLOCK)
    fan_h = fan_at(uc, R, idx)
    Rl = idx.size
    th, x0, y0 = (torch.from_numpy(a).to(dev) for a in fan_h)
    out = dict(d_ray=torch.empty(3, Rl, dtype=torch.float64, device=dev),
               last_row=torch.empty(6, Rl, dtype=torch.float64, device=dev),
               px_stats=torch.empty(3, Rl, dtype=torch.float64, device=dev))
    fused, gather_mode = None, None
    if world > 1:
        try:
            fused = FusedGather(eng, R, cyclic_block=BLOCK)
            gather_mode = "d_ray of all 2^%d rays on every GPU by peer stores from the kernel epilogue (block-cyclic index map)" % lg
        except RuntimeError:
            gather_mode = "none (peer mapping unavailable)"
    view = [None]

    def one_step():
        r = eng.trace(m, th, x0, y0, step, M, gamma, box, full=False, want_px_stats=True, out=out)
        if fused is not None:
            view[0] = fused.finish()
        return r
    steps = max(1, min(args.steps, 2))
    sampler = ClockSampler(D.local)
    if rank == 0:
        sampler.start()
    t_ms, rs = _timed(D, one_step, steps, 1)
    clocks = sampler.stop() if rank == 0 else None
    ksum = sum(r.stats["kernel_ms"] for r in rs)
    steps_done = sum(r.stats["ray_steps"] for r in rs)
    launches = sum(r.stats["launches"] for r in rs)
    gather_check = None
    if fused is not None:    # the fused array == scatter of an all-gather of the shards, bit for bit (sampled blocks)
        mine = torch.from_numpy(idx).to(dev)
        okf = torch.tensor([1.0 if torch.equal(view[0][:, mine], out["d_ray"]) else 0.0], dtype=torch.float64, device=dev)
        chk = view[0][2].sum()                        # every rank holds the same whole array
        chk_all = [torch.zeros_like(chk) for _ in range(world)]
        D.dist.all_gather(chk_all, chk)
        if any(float(a) != float(chk) for a in chk_all):
            okf.zero_()
        D.dist.all_reduce(okf, op=D.dist.ReduceOp.MIN)
        gather_check = "ok: own shard found at its global indices on this GPU, the same whole array on every GPU" \
            if okf.item() == 1.0 else "MISMATCH"
        total_from_gather = float(view[0][2].sum().item())
        fused.close()
    # scenario_average_CV RT:1354-1360 (rays with |mean p_x| > 1e-6; the reference skips its first and last ray)
    st = out["px_stats"]
    okr = st[1].abs() > 1e-6
    cv = torch.where(okr, 100 * st[2] / st[1], torch.zeros_like(st[1]))
    cv_mean = float((cv.sum() / okr.sum().clamp(min=1)).item())

    # end to end: host buffers (pinned numpy) through Engine.trace, H2D + D2H inside the call
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    fan_p = [pin(a) for a in fan_h]
    out_h = {k: torch.empty(v.shape, dtype=torch.float64).pin_memory().numpy() for k, v in out.items()}
    e2e_ms, rs2 = _wall(D, lambda: eng.trace(m, *fan_p, step, M, gamma, box, full=False, want_px_stats=True, out=out_h), 1, 1)
    e2e_steps = sum(r.stats["ray_steps"] for r in rs2)
    h2d, d2h = rs2[-1].stats["h2d_bytes"], rs2[-1].stats["d2h_bytes"]
    del fan_p, out_h

    (t_max, k_max, e2e_max, cvm, _, _, _), (_, _, _, _, total_steps, e2e_total, launches) = D.max_sum(
        [t_ms, ksum, e2e_ms, cv_mean, float(steps_done), float(e2e_steps), float(launches)])
    entry = None
    if rank == 0:
        scen = {"3": "vert_heterogeneous", "4": "anisotropy (gamma=3)"}[uc]
        entry = {"value": total_steps / (t_max * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": 1,
                 "ms_per_step": t_max / steps, "scaling": "strong",
                 "workload": "%s, 2^%d rays total, op%d, DELTA_S=SIGMA/20, max_size=%d, FINAL output + online p_x statistics"
                             % (scen, lg, m, M),
                 "rays_total": R, "rays_per_gpu": Rl, "ray_steps_per_step": int(total_steps // steps),
                 "l2": "inputs + outputs %.1f GB per GPU > L2" % (15 * 8 * Rl / 1e9),
                 "parallelism": ("block-cyclic ray shards (4096-ray blocks) x%d, no data-path collective; " % world + gather_mode)
                 if world > 1 else "1 GPU",
                 "px_cv_mean_percent_max_over_ranks": cvm,
                 "roofline": roofline_block(m, "trace_kernel<%d,%s,final,stats>" % (m, "iso" if gamma == 1 else "aniso"),
                                            total_steps / world / steps, k_max / steps, ctx["fp64_peak"],
                                            (clocks or {}).get("sm_mhz"), ctx["ncu"], ctx["ncu_note"], ncu_key, K.F_ALG),
                 "e2e": {"value": e2e_total / (e2e_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                         "d2h_bytes_per_step": d2h, "ms_per_step": e2e_max,
                         "api": "Engine.trace with pinned host numpy buffers, chunked H2D / kernel / D2H pipeline"},
                 "gpu_launches": int(launches), "clocks": clocks}
        if gather_check:
            entry["gather_check"] = gather_check
            entry["ray_steps_from_gathered_d_ray"] = total_from_gather
    del th, x0, y0, out
    torch.cuda.empty_cache()
    return entry


def run_interface_default(ctx):
    """configs[0]: the reference's own CPU-runnable case -- 'interface', default fan of 42 rays, op6, default DELTA_S,
    the full (30228, 6, 42) s_ray and the Snell errors.  42 rays are one GPU's work: rank 0 runs it."""
    import torch
    D, eng, rt, K = ctx["D"], ctx["eng"], ctx["rt"], ctx["K"]
    if D.rank != 0:
        return None
    uc, m = "1", 6
    z, grd = rt.build_field(uc)
    eng.set_field(z, grd)
    c = rt.scenario_constants(uc)
    th, x0, y0 = rt.default_fan(uc)
    M = rt.max_size_for(uc, rt.DELTA_S, 91)
    dev = D.dev
    a = [torch.from_numpy(np.ascontiguousarray(v)).to(dev) for v in (th, x0, y0)]
    out = dict(s_ray=torch.empty(M, 6, len(th), dtype=torch.float64, device=dev))
    for _ in range(2):
        r = eng.trace(m, *a, rt.DELTA_S, M, c[0], c[5:9], interface=True, full=True, out=out)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    n = 5
    for _ in range(n):
        r = eng.trace(m, *a, rt.DELTA_S, M, c[0], c[5:9], interface=True, full=True, out=out)
    e1.record()
    torch.cuda.synchronize()
    t_ms = e0.elapsed_time(e1) / n
    # end to end: the drop-in call with the reference's signature and return values (pageable numpy, 61 MB s_ray)
    rt.trazar(m, z, grd, False, rt.DELTA_S, 91, uc, engine=eng)
    t0 = time.perf_counter()
    s_ray, d_ray, _ct, errors = rt.trazar(m, z, grd, False, rt.DELTA_S, 91, uc, engine=eng)
    dt = time.perf_counter() - t0
    steps = int(d_ray[2].sum())
    return {"value": r.stats["ray_steps"] / (t_ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": n, "warmup": 2,
            "ms_per_step": t_ms, "scaling": "none (42 rays: one GPU)",
            "workload": "interface, default fan (42 rays), op6, DELTA_S=SIGMA/20, max_size=%d, FULL s_ray + Snell errors" % M,
            "ray_steps_per_step": steps, "mean_snell_error_deg": float(np.mean(errors)),
            "note": "a latency chain, not a throughput case: 42 rays of up to 8763 dependent steps, spread one ray per warp",
            "roofline": {"bound": "latency", "achieved": None, "frac": None,
                         "us_per_step_of_longest_ray": 1e3 * r.stats["kernel_ms"] / float(d_ray[2].max())},
            "e2e": {"value": steps / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "h2d_bytes_per_step": 3 * 8 * len(th),
                    "d2h_bytes_per_step": int(s_ray.nbytes + d_ray.nbytes + errors.nbytes),
                    "api": "raytracing_b200.trazar(op6, z, grd, False, DELTA_S, 91, '1'): the reference's signature and "
                           "return values, pageable numpy"},
            "gpu_launches": int(r.stats["launches"]) * n}


def shoot_roofline(ctx, method, per_gpu, kernel_ms, clocks, K):
    """The FP64 roofline block of the trace kernel (executed counts from the dense interface capture) plus what ncu
    measured on a launch of the refinement itself: its probe rays sit in different field cells, so the launch is bound
    by the L1 data pipe / L1 misses, not by FP64."""
    rb = roofline_block(method, "trace_kernel<%d,iso,final,exit>" % method, per_gpu, kernel_ms, ctx["fp64_peak"],
                        (clocks or {}).get("sm_mhz"), ctx["ncu"], ctx["ncu_note"], "op6_exit_interface", K.F_ALG)
    k = (ctx["ncu"] or {}).get("kernels", {}).get("shoot_refine")
    if k:
        rb["refine_launch"] = {"bound_measured": "l1 data pipe", "ncu_l1_data_pipe_frac": k.get("lsu_pipe_frac"),
                               "ncu_fp64_pipe_active_frac": k.get("fp64_pipe_active_frac"), "ncu_l1_hit_rate": k.get("l1_hit_rate"),
                               "ncu_issue_active_frac": k.get("issue_active_frac"), "source": k.get("source")}
    return rb


def run_shoot_config(ctx, n_cand=200, A=4096, Tn=4096, method=6):
    import torch
    D, eng, rt, K, args = ctx["D"], ctx["eng"], ctx["rt"], ctx["K"], ctx["args"]
    world, rank, dev = D.world, D.rank, D.dev
    uc = "1"
    z, grd = rt.build_field(uc)
    eng.set_field(z, grd)
    _divisors, options, _ = rt.sweep_candidates(uc)                 # RT:1303-1304: 200 candidates
    sel = np.linspace(0, len(options) - 1, n_cand).astype(int)
    steps_c = options[sel]
    angles = np.linspace(2 * np.pi / 60, np.pi / 2, A)              # the scenario's fan range RT:259-260
    true = np.linspace(0.2, 1.5, Tn) + 1.3e-4
    tx, ty, _T = rt.targets_from_angles(eng, method, uc, true, steps_c[0])
    grp = True if world > 1 else None

    def one_step():
        return rt.shooting_search(method, z, grd, uc, (tx, ty), angles, steps=steps_c, engine=eng, group=grp,
                                  device_arrays=(world > 1))
    steps = max(1, min(args.steps, 2))
    sampler = ClockSampler(D.local)
    if rank == 0:
        sampler.start()
    # the search takes host arrays in and gives host arrays out (device_arrays only inside the multi-GPU exchange):
    # the timed call IS the end-to-end call; its kernel time is reported beside it
    t_ms, rs = _timed(D, one_step, steps, 1)
    clocks = sampler.stop() if rank == 0 else None
    ksum = sum(r["stats"]["kernel_ms"] for r in rs)
    steps_done = sum(r["stats"]["ray_steps"] for r in rs)
    launches = sum(r["stats"]["launches"] for r in rs)
    (t_max, k_max, _, _), (_, _, total_steps, total_launches) = D.max_sum([t_ms, ksum, float(steps_done), float(launches)])
    if rank != 0:
        return None
    r = rs[-1]
    th = r["theta"].cpu().numpy() if hasattr(r["theta"], "cpu") else np.asarray(r["theta"])
    miss = r["miss"].cpu().numpy() if hasattr(r["miss"], "cpu") else np.asarray(r["miss"])
    ok = np.abs(true - np.pi / 4) > 0.12
    per_gpu = total_steps / world / steps
    bytes_io = (A + 2 * Tn + 2 * n_cand) * 8, 9 * n_cand * Tn * 8
    return {"value": total_steps / (t_max * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": 1,
            "ms_per_step": t_max / steps, "scaling": "strong",
            "workload": "two-point shooting search + DELTA_S sweep, interface, %d targets x %d launch angles x %d DELTA_S "
                        "candidates, op%d, golden refinement (%d passes)" % (Tn, A, n_cand, method, r["iterations"]),
            "ray_steps_per_step": int(total_steps // steps),
            "parallelism": ("coarse fan sharded by angle + NCCL MIN/argmin all-reduce, refinement sharded by target, result "
                            "all_gather, x%d" % world) if world > 1 else "1 GPU",
            "check": {"max_abs_theta_minus_truth_finest": float(np.abs(th[0] - true)[ok].max()),
                      "max_miss_finest": float(miss[0][ok].max())},
            "roofline": shoot_roofline(ctx, method, per_gpu, k_max / steps, clocks, K),
            "e2e": {"value": total_steps / (t_max * 1e-3), "unit": UNIT, "ms_per_step": t_max / steps,
                    "h2d_bytes_per_step": bytes_io[0], "d2h_bytes_per_step": bytes_io[1] if world == 1 else 0,
                    "api": "raytracing_b200.shooting_search: numpy arrays in (angles, targets, candidates), %s"
                           % ("numpy results out -- the timed call is the end-to-end call" if world == 1 else
                              "results gathered on every GPU (device tensors)")},
            "gpu_launches": int(total_launches), "clocks": clocks}


def run_secondary_configs(ctx):
    out = {}
    for name, fn in (("interface_default", lambda: run_interface_default(ctx)),
                     ("vert_64M", lambda: run_trace_config(ctx, "vert_64M", "3", 26, 6, "op6_stats_vert")),
                     ("aniso_16M_op11", lambda: run_trace_config(ctx, "aniso_16M_op11", "4", 24, 11, "op11_stats_aniso")),
                     ("aniso_16M_op10", lambda: run_trace_config(ctx, "aniso_16M_op10", "4", 24, 10, "op10_stats_aniso")),
                     ("shoot_4096x4096x200", lambda: run_shoot_config(ctx))):
        if ctx["args"].only and name not in ctx["args"].only.split(","):
            continue
        t0 = time.perf_counter()
        e = fn()
        if e is not None:
            e["bench_seconds"] = time.perf_counter() - t0
            out[name] = e
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--configs", default="all", choices=["all", "headline"],
                    help="all = the headline + BASELINE configs[0], [2], [3], [4] in the `configs` object (default)")
    ap.add_argument("--only", default="", help="comma-separated names of the `configs` entries to run (default: all of them)")
    ap.add_argument("--no-full", action="store_true", help="skip the FULL-trajectory (45.8 GB s_ray) legs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--gather", default="fused", choices=["fused", "nccl"],
                    help="N>1: d_ray reaches every GPU by peer stores from the kernel (fused) or by an NCCL all-gather")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
