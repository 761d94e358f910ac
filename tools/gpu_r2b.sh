#!/bin/bash
# Round 2, call B: GPU tests of the new paths + A/B timings (ray queue, TMA trajectory stores, shooting search).
O=gpurun_out/r2b; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_round2.py::test_dense_fan_vs_oracle_and_reference > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -30 $O/pytest_gpu.log
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -q -k dense_fan > $O/pytest_dense.log 2>&1; echo "pytest rc=$?" >> $O/pytest_dense.log
tail -30 $O/pytest_dense.log
timeout 600 python tools/time_r2.py > $O/time_r2.log 2>&1; cat $O/time_r2.log
timeout 300 python tools/shoot_probe.py 200 > $O/shoot_probe.log 2>&1; cat $O/shoot_probe.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; tail -3 $O/smoke.log
