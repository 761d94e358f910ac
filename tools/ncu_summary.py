"""Dev script: condense `ncu -i X.ncu-rep --page raw --csv` exports.

  ncu_summary.py table X.raw.csv OUT.csv                  compact metric,unit,value table for profiles/
  ncu_summary.py json OUT.json KEY=PREFIX [KEY=PREFIX..]  profiles/ncu_latest.json: per kernel the executed FP64
        instructions per ray-step, instructions per warp-step, pipe utilisations and DRAM bytes of the capture
        PREFIX.raw.csv, normalised by the ray-step count PREFIX.plain.log reports (tools/prof_trace.py prints
        `... ray_steps=N`).  Stamped with the digest of the trace-kernel sources (tools/build_id.py kernel_digest =
        rtb_kernel_id() of the library that was profiled), so bench.py only uses it for the kernels it describes.
"""
import csv
import json
import os
import re
import sys

KEEP = re.compile(r"Kernel Name|Block Size|Grid Size|gpu__time_duration\.sum|dram__bytes_(read|write)\.sum$|"
                  r"dram__bytes_(read|write)\.sum\.per_second|launch__registers_per_thread$|launch__occupancy_limit|"
                  r"sm__warps_active\.avg\.pct|sm__inst_executed_pipe_\w+\.avg\.pct_of_peak_sustained_active|"
                  r"sm__pipe_fp64_cycles_active\.avg|smsp__inst_executed\.sum$|smsp__issue_active\.avg\.pct|"
                  r"smsp__warps_eligible\.avg|smsp__average_warps_issue_stalled_\w+_per_issue_active|"
                  r"smsp__sass_thread_inst_executed_op_d(fma|mul|add)_pred_on\.sum|l1tex__t_sector_hit_rate|"
                  r"lts__t_sector_hit_rate|sm__throughput\.avg|smsp__cycles_active\.avg$|smsp__cycles_elapsed\.avg$|"
                  r"smsp__thread_inst_executed_per_inst_executed|sm__sass_inst_executed_op_global|"
                  r"l1tex__data_pipe_lsu_wavefronts\.avg\.pct|l1tex__t_bytes_pipe_lsu_mem_global_op_(ld|st)\.sum$|lts__t_bytes\.sum$")


def table(path, out):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        for k, vals in enumerate(rows[2:]):
            f.write("# launch %d\nmetric,unit,value\n" % k)
            for h, u, v in zip(hdr, units, vals):
                if KEEP.search(h) and v != "":
                    f.write('%s,%s,"%s"\n' % (h, u, v))


UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "us": 1e-3, "ms": 1.0, "s": 1e3,
              "usecond": 1e-3, "msecond": 1.0, "second": 1e3}


def kernel_entry(prefix, key):
    rows = list(csv.reader(open(prefix + ".raw.csv")))
    hdr, units, vals = rows[0], rows[1], rows[2]
    col = {h: (u, v) for h, u, v in zip(hdr, units, vals)}

    def num(name, scale=False):
        u, v = col[name]
        x = float(v.replace(",", ""))
        return x * UNIT_SCALE.get(u, 1.0) if scale else x
    m = re.search(r"ray_steps=(\d+)", open(prefix + ".plain.log").read())
    steps = int(m.group(1)) if m else None     # None: one launch out of many (shooting refinement): no per-step figures
    cyc = num("smsp__cycles_elapsed.avg")
    per = lambda op: num("smsp__sass_thread_inst_executed_op_%s_pred_on.sum.per_cycle_elapsed" % op) * cyc / steps if steps else None
    return {
        "kernel": col["Kernel Name"][1], "workload": key, "ray_steps": steps,
        "ncu_duration_ms": num("gpu__time_duration.sum", True),
        "dfma": per("dfma"), "dmul": per("dmul"), "dadd": per("dadd"),
        "inst_per_warp_step": num("smsp__inst_executed.sum") / (steps / 32.0) if steps else None,
        "fp64_pipe_active_frac": num("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active") / 100.0,
        "issue_active_frac": num("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100.0,
        "lsu_pipe_frac": num("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed") / 100.0,
        "warps_active_frac": num("sm__warps_active.avg.pct_of_peak_sustained_active") / 100.0,
        "registers": int(num("launch__registers_per_thread")),
        "l1_hit_rate": num("l1tex__t_sector_hit_rate.pct") / 100.0,
        "dram_bytes_per_launch": num("dram__bytes_read.sum", True) + num("dram__bytes_write.sum", True),
        "source": "profiles/" + os.path.basename(prefix) + ".csv",
    }


def to_json(out, pairs):
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import build_id
    d = {"kernel_id": build_id.kernel_digest(), "how": "ncu --set full --clock-control none, one launch per kernel after warm-up "
         "launches (tools/gpu_*.sh); per ray-step = thread instructions / ray-steps of the launch", "kernels": {}}
    for p in pairs:
        key, prefix = p.split("=", 1)
        d["kernels"][key] = kernel_entry(prefix, key)
    json.dump(d, open(out, "w"), indent=1)
    print("wrote", out, "kernel_id", d["kernel_id"], "kernels", list(d["kernels"]))


if __name__ == "__main__":
    if sys.argv[1] == "table":
        table(sys.argv[2], sys.argv[3])
    elif sys.argv[1] == "json":
        to_json(sys.argv[2], sys.argv[3:])
    else:   # old form: ncu_summary.py X.raw.csv OUT.csv
        table(sys.argv[1], sys.argv[2])
