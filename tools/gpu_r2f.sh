#!/bin/bash
# Round 2, call F (2 GPUs): GPU tests incl. the 2-process fused gather, FULL-mode A/B, bench at N=2 and N=1.
O=gpurun_out/r2f; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/smi.txt
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
grep -E "passed|failed|FAILED|Error|rc=" $O/pytest_gpu.log | head -20
timeout 600 python tools/time_r2.py 2>&1 | grep -E "FULL|interface|vert 2\^22" > $O/time_r2.log; cat $O/time_r2.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n2 rc=$?" >> $O/bench_n2.err
tail -3 $O/bench_n2.err; cat $O/bench_n2.json | cut -c1-1500
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 --configs headline > $O/bench_ref_n2.json 2> $O/bench_ref_n2.err; echo "ref n2 rc=$?"; cut -c1-300 $O/bench_ref_n2.json
