#!/bin/bash
# Round 2, call I: tests + trajectory debug + bench headline.
O=gpurun_out/r2i; mkdir -p $O
python tools/debug_traj.py 2>&1 | grep -E "^case|last row" 
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
grep -E "passed|failed|FAILED|Error|rc=" $O/pytest_gpu.log | head -20
timeout 600 python tools/time_r2.py 2>&1 | grep -E "FULL" > $O/time_r2.log; cat $O/time_r2.log
