#!/bin/bash
# bench.py at N=8 (and N=4) on one 8-GPU box, the way the driver launches it.
O=gpurun_out/scale8; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/smi.txt
for N in 8 4; do
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench n$N rc=$?" >> $O/bench_n$N.err
tail -2 $O/bench_n$N.err; grep "^{" $O/bench_n$N.json | cut -c1-400
done
