#!/bin/bash
# ncu captures of the shipped kernels -> profiles/ (tools/ncu_summary.py json builds profiles/ncu_latest.json from them),
# the launch list of the bench command, the parity report and the method timing table.  One GPU.
O=gpurun_out/ncu_final; mkdir -p $O
prof() { # name, kernel regex, skip, command...
  local name=$1 re=$2 skip=$3; shift 3
  timeout 300 "$@" > $O/$name.plain.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$re -s $skip -c 1 -o $O/$name -f "$@" > $O/$name.ncu.log 2>&1
  tail -1 $O/$name.plain.log
  ncu -i $O/$name.ncu-rep --page raw --csv > $O/$name.raw.csv 2>/dev/null
  # gpurun brings back at most 64 MiB: keep the per-instruction page of the headline kernel, drop the other reports
  if [ "$name" = "op6_final_fisheye" ] || [ "$name" = "op11_stats_aniso" ]; then ncu -i $O/$name.ncu-rep --page source --csv > $O/$name.source.csv 2>/dev/null; fi
  rm -f $O/$name.ncu-rep
}
prof op6_final_fisheye trace_kernel 2 python tools/prof_trace.py 2 6 20 final 3
prof op6_full_fisheye trace_kernel 2 python tools/prof_trace.py 2 6 20 full 3
prof op6_stats_vert trace_kernel 2 python tools/prof_trace.py 3 6 22 stats 3
prof op9_final_fisheye trace_kernel 2 python tools/prof_trace.py 2 9 20 final 3
prof op11_stats_aniso trace_kernel 2 python tools/prof_trace.py 4 11 20 stats 3
prof op10_stats_aniso trace_kernel 2 python tools/prof_trace.py 4 10 20 stats 3
prof op6_exit_interface trace_kernel 2 python tools/prof_trace.py 1 6 18 exit 3
prof bracket_shoot bracket_kernel 1 python tools/shoot_probe.py 8 4096
prof shoot_refine trace_kernel 8 python tools/shoot_probe.py 200 4096 1
timeout 600 python bench.py --steps 2 --warmup 3 --configs headline --no-cpu > $O/bench_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench.csv python bench.py --steps 2 --warmup 3 --configs headline --no-cpu > $O/ncu_launch.log 2>&1
timeout 900 python tools/parity_report.py > $O/parity_report.txt 2>&1; tail -5 $O/parity_report.txt
timeout 600 python tools/time_methods.py > $O/method_timings.txt 2>&1
rm -f $O/*.ncu-rep.tmp; ls -la $O | head -50
