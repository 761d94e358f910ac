#!/bin/bash
# Round 2, call J (1 GPU): tests, the full bench line (both arms), sort_fan timing.
O=gpurun_out/r2j; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
grep -E "passed|failed|FAILED|Error|rc=" $O/pytest_gpu.log | head -20
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
time timeout 1200 python bench.py --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 600 python tools/time_r2.py 2>&1 | grep -E "SHUFFLED|FULL" > $O/time_r2.log; cat $O/time_r2.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
