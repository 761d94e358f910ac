#!/bin/bash
# One ncu --set full capture with the source page: gpu_ncu_one.sh <name> <prof_trace.py args...>
name=$1; shift
O=gpurun_out/ncu_one; mkdir -p $O
timeout 300 python tools/prof_trace.py "$@" > $O/$name.plain.log 2>&1 || { tail -3 $O/$name.plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_kernel -s 2 -c 1 -o $O/$name -f python tools/prof_trace.py "$@" > $O/$name.ncu.log 2>&1
tail -1 $O/$name.plain.log
ncu -i $O/$name.ncu-rep --page raw --csv > $O/$name.raw.csv 2>/dev/null
ncu -i $O/$name.ncu-rep --page source --csv > $O/$name.source.csv 2>/dev/null
rm -f $O/$name.ncu-rep; ls -la $O
