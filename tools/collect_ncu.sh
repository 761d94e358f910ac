#!/bin/bash
# After tools/gpu_ncu_final.sh came back: condense gpurun_out/ncu_final into profiles/ (tables, ncu_latest.json, source page,
# SASS of the op6 loop, launch list, parity report, method timings).
set -e
cd "$(dirname "$0")/.."
O=gpurun_out/ncu_final
P="op6_final_fisheye op6_full_fisheye op6_stats_vert op9_final_fisheye op11_stats_aniso op10_stats_aniso op6_exit_interface"
rm -rf /tmp/ncuj; mkdir -p /tmp/ncuj
for k in $P; do
  python tools/ncu_summary.py table $O/$k.raw.csv profiles/r2_ncu_$k.csv
  cp $O/$k.raw.csv /tmp/ncuj/r2_ncu_$k.raw.csv; cp $O/$k.plain.log /tmp/ncuj/r2_ncu_$k.plain.log
done
python tools/ncu_summary.py table $O/bracket_shoot.raw.csv profiles/r2_ncu_bracket_kernel.csv
python tools/ncu_summary.py table $O/shoot_refine.raw.csv profiles/r2_ncu_shoot_refine.csv
python tools/ncu_source_summary.py $O/op11_stats_aniso.source.csv $(grep -o "ray_steps=[0-9]*" $O/op11_stats_aniso.plain.log | cut -d= -f2) > profiles/r2_ncu_source_op11_stats_aniso.txt
cp $O/shoot_refine.raw.csv /tmp/ncuj/r2_ncu_shoot_refine.raw.csv; cp $O/shoot_refine.plain.log /tmp/ncuj/r2_ncu_shoot_refine.plain.log
python tools/ncu_summary.py json profiles/ncu_latest.json $(for k in $P shoot_refine; do echo -n "$k=/tmp/ncuj/r2_ncu_$k "; done)
python tools/ncu_source_summary.py $O/op6_final_fisheye.source.csv 953155584 > profiles/r2_ncu_source_op6_final_fisheye.txt
python tools/sass_loop.py raytracing_b200/csrc/trace_m6.o trace_kernelILi6ELb1ELi0ELb0ELi0E > profiles/r2_sass_loop_op6_final.txt
cuobjdump -sass raytracing_b200/csrc/trace_m6.o 2>/dev/null | awk '/Function : .*trace_kernelILi6ELb1ELi0ELb0ELi0E/{f=1} f&&/Function : /&&!/trace_kernelILi6ELb1ELi0ELb0ELi0E/{f=0} f' | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed 's/ *\/\* 0x[0-9a-f]* \*\///' >> profiles/r2_sass_loop_op6_final.txt
cp $O/launches_bench.csv profiles/r2_launches_bench_op6.csv
cp $O/parity_report.txt profiles/r2_parity_report.txt
cp $O/method_timings.txt profiles/r2_method_timings.txt
