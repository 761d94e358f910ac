#!/bin/bash
# configs[2..4] bench lines on NG GPUs (default 1)
NG=${1:-1}
O=gpurun_out; mkdir -p $O
run() { # workload extra-args...
  w=$1; shift
  if [ $NG -eq 1 ]; then timeout 900 python tools/bench_configs.py --workload $w --gpus 1 "$@"
  else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29533 tools/bench_configs.py --workload $w --gpus $NG "$@"; fi
}
run shoot --steps 3 --warmup 1 > $O/cfg_shoot_n$NG.json 2> $O/cfg_shoot_n$NG.err; tail -1 $O/cfg_shoot_n$NG.json | cut -c1-900
run vert --steps 2 --warmup 1 > $O/cfg_vert_n$NG.json 2> $O/cfg_vert_n$NG.err; tail -1 $O/cfg_vert_n$NG.json | cut -c1-900
run aniso --steps 1 --warmup 1 > $O/cfg_aniso_n$NG.json 2> $O/cfg_aniso_n$NG.err; tail -2 $O/cfg_aniso_n$NG.json | cut -c1-900
tail -n 3 $O/cfg_*_n$NG.err; true
