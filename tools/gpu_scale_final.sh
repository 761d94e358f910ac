#!/bin/bash
# Final scaling pass on an 8-GPU box: headline bench at N=2,4,8 (own arm), configs[2..4] at N=8.
O=gpurun_out; mkdir -p $O
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
    bench.py --gpus $n --steps 10 --warmup 3 --no-full --no-cpu > $O/scale_n$n.json 2> $O/scale_n$n.err
  echo "n=$n rc=$?"; tail -1 $O/scale_n$n.json | cut -c1-400
done
bash tools/gpu_configs.sh 8
