"""Dev script: turn `ncu -i X.ncu-rep --page raw --csv` into a compact metric,unit,value table for profiles/."""
import csv
import re
import sys

KEEP = re.compile(r"Kernel Name|Block Size|Grid Size|gpu__time_duration\.sum|dram__bytes_(read|write)\.sum$|"
                  r"dram__bytes_(read|write)\.sum\.per_second|launch__registers_per_thread$|launch__occupancy_limit|"
                  r"sm__warps_active\.avg\.pct|sm__inst_executed_pipe_\w+\.avg\.pct_of_peak_sustained_active|"
                  r"sm__pipe_fp64_cycles_active\.avg|smsp__inst_executed\.sum$|smsp__issue_active\.avg\.pct|"
                  r"smsp__warps_eligible\.avg|smsp__average_warps_issue_stalled_\w+_per_issue_active|"
                  r"smsp__sass_thread_inst_executed_op_d(fma|mul|add)_pred_on\.sum|l1tex__t_sector_hit_rate|"
                  r"lts__t_sector_hit_rate|sm__throughput\.avg|smsp__cycles_active\.avg$|"
                  r"smsp__thread_inst_executed_per_inst_executed|sm__sass_inst_executed_op_global|"
                  r"l1tex__t_bytes_pipe_lsu_mem_global_op_(ld|st)\.sum$|lts__t_bytes\.sum$")


def main(path, out):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        for k, vals in enumerate(rows[2:]):
            f.write("# launch %d\nmetric,unit,value\n" % k)
            for h, u, v in zip(hdr, units, vals):
                if KEEP.search(h) and v != "":
                    f.write('%s,%s,"%s"\n' % (h, u, v))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
