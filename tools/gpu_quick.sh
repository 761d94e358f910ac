#!/bin/bash
# Dev: parity tests + method timings (+ optional ncu source-level capture of op6 with NCU=1)
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/time_methods.py 2>&1 | grep -v "exact=1" | tee $O/time_methods.log
if [ "$NCU" = "1" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:trace_kernel -c 1 -s 3 -o $O/op6_full -f python bench.py --steps 2 --warmup 3 --no-full --no-cpu > $O/ncu_full.log 2>&1
  ncu -i $O/op6_full.ncu-rep --page raw --csv > $O/op6_full_raw.csv 2>/dev/null
  ncu -i $O/op6_full.ncu-rep --page source --csv > $O/op6_full_source.csv 2>/dev/null
fi
