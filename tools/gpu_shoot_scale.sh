#!/bin/bash
# Shooting-search config alone at N = $1 (and the mgpu tests when N == 2).
N=${1:-2}; O=gpurun_out/shoot_n$N; mkdir -p $O
if [ "$N" = 2 ]; then timeout 900 python -m pytest tests -m gpu -q -k "mgpu or shoot or dist or gather" > $O/pytest.log 2>&1; tail -3 $O/pytest.log; fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 3 --warmup 3 --no-full --only shoot_4096x4096x200 > $O/bench.json 2> $O/bench.err; echo "rc=$?"
tail -2 $O/bench.err
python - <<PY
import json
for l in open("$O/bench.json"):
    if l.startswith("{"):
        d = json.loads(l); s = d["configs"]["shoot_4096x4096x200"]
        print("N=%d shoot %.4e ms %.1f kernel_ms %.2f check %s" % (d["n_gpus"], s["value"], s["ms_per_step"], s["roofline"].get("kernel_ms", -1), s["check"]))
PY
