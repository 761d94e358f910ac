#!/usr/bin/env python
"""bench.py -- FP64 ray-steps/s of the RT_bench.py hot path on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --steps K --warmup W    # CPU arm: the oracle port on host cores

A "step" is one pass of the hot path (one trazar) over the whole synthetic fan.  Workload at
N=1 is BASELINE.json configs[1]: 'fisheye', 2^20 rays, FP64, method op6 (HySA), default
DELTA_S = SIGMA/20, max_size = N*(90+1) = 910 (909 ray-steps per ray).  At N>1 every rank traces
its own 2^20-ray shard of an N*2^20 fan (weak scaling, no data-path collective; the d_ray
result gather over NCCL is inside the timed step).
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
# Executed (not algorithmic) FP64 work of the op6 FINAL kernel per ray-step, counted by ncu on this workload
# (profiles/r1f_ncu_full_op6_fisheye1M.csv + the per-instruction page): 64.1 DFMA + 24.3 DMUL + 17.3 DADD warp
# instructions per warp-step.  The algorithmic figure of SURVEY.md 8(d) (376 flop: B-spline-form field evaluation,
# 25 flop per sin or cos) is larger than what the kernel executes, so `roofline.frac` can exceed 1; the executed
# figures below are the hardware-side view of the same launch.
NCU_OP6 = {"dfma": 59.5, "dmul": 19.4, "dadd": 13.3, "l1_data_pipe_frac": 0.59, "fp64_pipe_active_frac": 0.637,
           "dram_bytes_per_launch": 26.645e6 + 19.594e6, "source": "profiles/r1j_ncu_full_op6_fisheye1M.csv"}
METRIC = "fp64_ray_steps_per_s"
UNIT = "ray-steps/s"


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


def cpu_port_rate(sample_rays, nthreads):
    """Time the oracle (C port of the reference algorithm) on a bounded slice of the same fan."""
    import raytracing_b200 as rt
    from oracle import rt_oracle as O
    z, grd = rt.build_field(SCENARIO)
    c = rt.scenario_constants(SCENARIO)
    M = rt.max_size_for(SCENARIO, rt.DELTA_S, rt.DELTA_S_DIVISOR_FISHEYE + 1)
    stride = RAYS_PER_GPU // sample_rays
    th, x0, y0 = rt.make_fan(SCENARIO, RAYS_PER_GPU)
    th, x0, y0 = th[::stride][:sample_rays], x0[:sample_rays], y0[:sample_rays]
    zs, g0, g1 = (O.HostSpline.from_scipy(s) for s in (z, grd[0], grd[1]))
    O.trace(METHOD, zs, (g0, g1), rt.DELTA_S, M, c[0], c[5:9], False, th[:64], x0[:64], y0[:64], full=False,
            nthreads=nthreads)
    t0 = time.perf_counter()
    r = O.trace(METHOD, zs, (g0, g1), rt.DELTA_S, M, c[0], c[5:9], False, th, x0, y0, full=False, nthreads=nthreads)
    dt = time.perf_counter() - t0
    return r["ray_steps"], dt


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
    print(json.dumps(line))


def run_cuda(args):
    import torch
    import raytracing_b200 as rt
    from raytracing_b200 import constants as K

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)

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

    def one_step():
        r = eng.trace(METHOD, th, x0, y0, step, M, gamma, box, full=False, out=out)
        if fused is not None:  # peers were written by the kernel; ranks only agree that they are done
            fused.finish()
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
    if world > 1:
        dist.barrier()
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
    if world > 1:
        dist.barrier()
    t_ms = sum(a.elapsed_time(b) for a, b in ev)
    # the event pairs must bracket the kernels (same stream): a step cannot be shorter than its kernel
    if t_ms < 0.99 * sum(kernel_ms):
        raise RuntimeError("timing events (%.3f ms) do not bracket the kernels (%.3f ms): stream mismatch"
                           % (t_ms, sum(kernel_ms)))
    clocks = sampler.stop() if rank == 0 else None
    gather_mode = None
    if world > 1:
        gather_mode = ("peer stores from the kernel epilogue over NVLink (CUDA IPC) + 1-element all-reduce"
                       if fused is not None else "NCCL all_gather after the kernel")
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
    if world > 1:
        dist.barrier()
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
    full_info = None
    if rank == 0 and not args.no_full:
        s_ray = torch.empty(M, 6, R, dtype=torch.float64, device=dev)
        for _ in range(2):
            rf = eng.trace(METHOD, th, x0, y0, step, M, gamma, box, full=True, out=dict(s_ray=s_ray, **out))
        gbs = rf.stats["ray_steps"] * 48 / (rf.stats["kernel_ms"] * 1e-3) / 1e9
        full_info = {"ms": rf.stats["kernel_ms"], "ray_steps_per_s": rf.stats["ray_steps"] / (rf.stats["kernel_ms"] * 1e-3),
                     "hbm_write_gbs": gbs}
        del s_ray

    # max over ranks of the device time; sum of the work
    agg = torch.tensor([t_ms, e2e_ms, float(ray_steps), float(e2e_steps), float(launches)], dtype=torch.float64, device=dev)
    if world > 1:
        mx = agg.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = agg.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        t_ms, e2e_ms = mx[0].item(), mx[1].item()
        ray_steps, e2e_steps, launches = int(sm[2].item()), int(sm[3].item()), int(sm[4].item())

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
        achieved = per_gpu_steps * K.F_ALG[METHOD] / (k_ms * 1e-3) / 1e12
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": t_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(world), "scenario": "fisheye", "method": "op6",
                           "rays_per_gpu": R, "max_size": M, "ray_steps_per_step": ray_steps // args.steps,
                           "l2": "flushed (256 MiB write) between timed steps",
                           "parallelism": ("rays sharded x%d, d_ray gathered on every GPU: %s" % (world, gather_mode))
                           if world > 1 else "1 GPU"},
                "roofline": {"bound": "fp64", "achieved": achieved, "peak": fp64_peak / 1e12, "unit": "TFLOP/s",
                             "frac": achieved / (fp64_peak / 1e12),
                             "traffic": NCU_OP6["dram_bytes_per_launch"], "traffic_unit": "DRAM bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum; the call's inputs + outputs are 101 MB, mostly absorbed by the 126 MB L2)",
                             "kernel": "trace_kernel<6,iso,final>", "kernel_ms": k_ms,
                             "flop_per_ray_step": K.F_ALG[METHOD],
                             "executed": {"fp64_inst_per_ray_step": NCU_OP6["dfma"] + NCU_OP6["dmul"] + NCU_OP6["dadd"],
                                          "flop_per_ray_step": 2 * NCU_OP6["dfma"] + NCU_OP6["dmul"] + NCU_OP6["dadd"],
                                          "tflops": per_gpu_steps * (2 * NCU_OP6["dfma"] + NCU_OP6["dmul"] + NCU_OP6["dadd"])
                                          / (k_ms * 1e-3) / 1e12,
                                          "ncu_fp64_pipe_active_frac": NCU_OP6["fp64_pipe_active_frac"],
                                          "ncu_l1_data_pipe_frac": NCU_OP6["l1_data_pipe_frac"],
                                          "source": NCU_OP6["source"]},
                             "peak_source": "measured live: rtb_fp64_peak (DFMA stream on all SMs); "
                                            "MEASURED_PEAKS.json has no FP64 figure"},
                "e2e": {"value": e2e_steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / args.steps,
                        "api": "Engine.trace with pinned host numpy buffers (RTB_PTR_HOST): rtb_trace pipelines the fan in 8 chunks "
                               "(H2D copy, kernel, D2H copy on three streams)"},
                "gpu_launches": launches, "clocks": clocks}
        if full_info:
            full_info.update({"hbm_peak_gbs": hbm_peak, "hbm_frac": full_info["hbm_write_gbs"] / hbm_peak,
                              "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"})
            line["full_trajectory_mode"] = full_info
        if world == 1 and not args.no_cpu:
            cores = host_cores()
            s, dt = cpu_port_rate(4096, cores)
            line["cpu_baseline"] = {"value": s / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "4096-ray strided slice of the same fan (%d ray-steps), oracle C port, "
                                              "OpenMP over rays" % s}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-full", action="store_true", help="skip the FULL-trajectory (45.8 GB s_ray) leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--gather", default="fused", choices=["fused", "nccl"],
                    help="N>1: d_ray reaches every GPU by peer stores from the kernel (fused) or by an NCCL all-gather")
    ap.add_argument("--workload", default="fisheye", choices=["fisheye", "vert", "aniso", "shoot"],
                    help="fisheye = the headline configs[1] line (default); vert / aniso / shoot = configs[2..4], "
                         "run by tools/bench_configs.py")
    args = ap.parse_args()
    if args.workload != "fisheye" and args.impl == "cuda":
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_configs
        return bench_configs.main(["--workload", args.workload, "--gpus", str(args.gpus), "--steps", str(args.steps),
                                   "--warmup", str(args.warmup)])
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
