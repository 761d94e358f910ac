#!/bin/bash
# Dev: build librtb variant into raytracing_b200/csrc/variants/<name>/librtb.so with extra nvcc flags.
# usage: tools/build_variant.sh <name> [extra nvcc flags...]   (only methods 1 and 6 are rebuilt quickly: METHODS env)
set -e
name=$1; shift
src=$(cd $(dirname $0)/../raytracing_b200/csrc && pwd)
out=$src/variants/$name; mkdir -p $out
ARCH="-gencode arch=compute_100a,code=sm_100a"
FL="-O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC -Xptxas -v $@"
pids=()
( nvcc $FL -c $src/rtb.cu -o $out/rtb.o 2> $out/rtb.log ) & pids+=($!)
( nvcc $FL -c $src/shoot.cu -o $out/shoot.o 2> $out/shoot.log ) & pids+=($!)
( nvcc $FL -c $src/sweep.cu -o $out/sweep.o 2> $out/sweep.log ) & pids+=($!)
( nvcc $FL -c $src/wavefront.cu -o $out/wavefront.o 2> $out/wavefront.log ) & pids+=($!)
( nvcc $FL -c $src/fit.cu -o $out/fit.o 2> $out/fit.log ) & pids+=($!)
for m in 1 2 3 4 5 6 7 8 9 10 11; do
  ( nvcc $FL -DRTB_METHOD=$m -c $src/trace_inst.cu -o $out/trace_m$m.o 2> $out/trace_m$m.log ) & pids+=($!)
done
for p in ${pids[@]}; do wait $p; done
nvcc -shared $ARCH -o $out/librtb.so $out/*.o
grep -A2 "trace_kernelILi6ELb1ELi[012]ELb0ELi0E\|trace_queue_kernelILi6ELb1ELi0E" $out/trace_m6.log | grep -E "Compiling|spill|Used"
