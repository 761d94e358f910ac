#!/bin/bash
# Dev: time kernel variants built by tools/build_variant.sh (RTB_LIB_PATH override) + the in-tree default.
O=gpurun_out; mkdir -p $O
V=raytracing_b200/csrc/variants
echo "== default" > $O/variants.log
timeout 300 python tools/time_methods.py 2>&1 | grep -E "op6|op1 |op3|op9|op11" >> $O/variants.log
for d in $V/*/; do
  n=$(basename $d)
  echo "== $n" >> $O/variants.log
  RTB_LIB_PATH=$d/librtb.so timeout 300 python tools/time_methods.py 2>&1 | grep -E "op6|op1 |op3|op9|op11" >> $O/variants.log
done
cat $O/variants.log
