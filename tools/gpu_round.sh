#!/bin/bash
# One GPU box session: tests, smoke, bench (both arms), method timing table, ncu launch list + full capture.
set -x
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $O/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
timeout 600 python tools/time_methods.py > $O/time_methods.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-full --no-cpu > $O/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_kernel -c 1 -s 3 -o $O/op6_full -f python bench.py --steps 2 --warmup 3 --no-full --no-cpu > $O/ncu_full.log 2>&1
ncu -i $O/op6_full.ncu-rep --page raw --csv > $O/op6_full_raw.csv 2>/dev/null
tail -3 $O/pytest_gpu.log; cat $O/smoke.log | tail -3; cat $O/bench.json; cat $O/bench_ref.json; tail -30 $O/time_methods.log
