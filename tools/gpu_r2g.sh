#!/bin/bash
# Round 2, call G (1 GPU): tests, FULL-mode A/B, shooting-search group sweep, compute-sanitizer memcheck.
O=gpurun_out/r2g; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
grep -E "passed|failed|FAILED|Error|rc=" $O/pytest_gpu.log | head -20
timeout 600 python tools/time_r2.py 2>&1 | grep -E "FULL|interface|vert 2\^22" > $O/time_r2.log; cat $O/time_r2.log
timeout 600 python tools/shoot_probe.py 200 4096 1 2 4 8 16 > $O/shoot_probe.log 2>&1; grep "^G=" $O/shoot_probe.log
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python __graft_entry__.py smoke > $O/sanitizer_memcheck_smoke.log 2>&1; echo "memcheck smoke rc=$?" | tee -a $O/sanitizer_memcheck_smoke.log
tail -5 $O/sanitizer_memcheck_smoke.log
