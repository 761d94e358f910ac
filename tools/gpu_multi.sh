#!/bin/bash
# Multi-GPU box session: bench at N = 1..NG (torchrun launch exactly like the driver), both arms.
NG=${1:-2}
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/smi_multi.txt 2>&1
nvidia-smi topo -m >> $O/smi_multi.txt 2>&1
for n in 1 2 4 8; do
  [ $n -gt $NG ] && break
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-full > $O/scale_n1.json 2> $O/scale_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $n --steps 10 --warmup 3 --no-full > $O/scale_n$n.json 2> $O/scale_n$n.err
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --impl reference --gpus $n --steps 2 --warmup 1 > $O/scale_ref_n$n.json 2> $O/scale_ref_n$n.err
  fi
  echo "n=$n rc=$?"; tail -1 $O/scale_n$n.json
done
