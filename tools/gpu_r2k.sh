#!/bin/bash
# Round 2, call K (1 GPU): tests + method timings after trimming the anisotropic golden pass.
O=gpurun_out/r2k; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
grep -E "passed|failed|FAILED|Error|rc=" $O/pytest_gpu.log | head -20
for m in 11 10; do timeout 300 python tools/prof_trace.py 4 $m 20 stats 2; done
timeout 300 python tools/prof_trace.py 2 9 20 final 2; timeout 300 python tools/prof_trace.py 2 5 20 final 2; timeout 300 python tools/prof_trace.py 2 6 20 final 2
