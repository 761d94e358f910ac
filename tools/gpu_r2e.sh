#!/bin/bash
# Round 2, call E: GPU tests + the full bench line at N=1 (all configs) + both arms.
O=gpurun_out/r2e; mkdir -p $O
true

time timeout 1200 python bench.py --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
tail -5 $O/bench.err | head -3; cat $O/bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; cat $O/bench_ref.json | cut -c1-600
for m in 11 10; do timeout 300 python tools/prof_trace.py 4 $m 20 stats 2; done
