#!/bin/bash
# Round 2, call C: GPU tests again + A/B timings incl. build variants + ncu of the staged TRAJ kernel.
O=gpurun_out/r2c; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -30 $O/pytest_gpu.log
timeout 600 python tools/time_r2.py > $O/time_r2.log 2>&1; cat $O/time_r2.log
for v in q4 t4 rows2; do
  echo "== variant $v" >> $O/time_r2_variants.log
  RTB_LIB_PATH=raytracing_b200/csrc/variants/$v/librtb.so timeout 600 python tools/time_r2.py 2>&1 | grep -E "fisheye 2\^20 op6|vert 2\^22 op6 final|MIXED|SHUFFLED  " >> $O/time_r2_variants.log
done
cat $O/time_r2_variants.log
timeout 300 python tools/shoot_probe.py 200 > $O/shoot_probe.log 2>&1; head -2 $O/shoot_probe.log
timeout 300 python tools/prof_trace.py 2 6 20 full 3 > $O/op6_full.plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_kernel -s 2 -c 1 -o $O/op6_full_tma -f python tools/prof_trace.py 2 6 20 full 3 > $O/op6_full.ncu.log 2>&1
ncu -i $O/op6_full_tma.ncu-rep --page raw --csv > $O/op6_full_tma.raw.csv 2>/dev/null
