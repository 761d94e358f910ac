#!/bin/bash
# Round 2, call D: GPU tests + method timings after the golden-section rewrite.
O=gpurun_out/r2d; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -40 $O/pytest_gpu.log
timeout 600 python tools/time_methods.py > $O/time_methods.log 2>&1; cat $O/time_methods.log
timeout 300 python tools/prof_trace.py 4 11 20 stats 2 ; timeout 300 python tools/prof_trace.py 4 10 20 stats 2; timeout 300 python tools/prof_trace.py 2 9 20 final 2; timeout 300 python tools/prof_trace.py 2 5 20 final 2
