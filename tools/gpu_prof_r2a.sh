#!/bin/bash
# Round 2, call A: write-bandwidth microbenchmark + ncu --set full of the kernels that had no profile in round 1
# (TRAJ flavour, p_x-statistics flavour on vert, op9, op11, bracket_kernel).  One GPU.
O=gpurun_out/r2a; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $O/smi.txt 2>&1
timeout 120 tools/bin/hbm_write_microbench > $O/hbm_write_microbench.txt 2>&1
cat $O/hbm_write_microbench.txt
prof() { # name, kernel regex, skip, command...
  local name=$1 re=$2 skip=$3; shift 3
  timeout 300 "$@" > $O/$name.plain.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$re -s $skip -c 1 -o $O/$name -f "$@" > $O/$name.ncu.log 2>&1
  tail -1 $O/$name.plain.log
  ncu -i $O/$name.ncu-rep --page raw --csv > $O/$name.raw.csv 2>/dev/null
}
prof op6_full_fisheye1M trace_kernel 2 python tools/prof_trace.py 2 6 20 full 3
prof op6_stats_vert4M trace_kernel 2 python tools/prof_trace.py 3 6 22 stats 3
prof op6_final_vert4M trace_kernel 2 python tools/prof_trace.py 3 6 22 final 3
prof op9_final_fisheye1M trace_kernel 2 python tools/prof_trace.py 2 9 20 final 3
prof op11_stats_aniso1M trace_kernel 2 python tools/prof_trace.py 4 11 20 stats 3
prof op6_final_fisheye1M trace_kernel 2 python tools/prof_trace.py 2 6 20 final 3
prof bracket_shoot bracket_kernel 0 python tools/shoot_probe.py
ls -la $O
