#!/bin/bash
# ncu --set full of trace_kernel launches inside the shooting search's refinement (one group: 1.6M probe rays per launch).
O=gpurun_out/ncu_shoot; mkdir -p $O
timeout 300 python tools/shoot_probe.py 200 4096 1 > $O/plain.log 2>&1 || { tail -5 $O/plain.log; exit 1; }
grep "G=" $O/plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_kernel -s 8 -c 2 -o $O/shoot -f python tools/shoot_probe.py 200 4096 1 > $O/ncu.log 2>&1
ncu -i $O/shoot.ncu-rep --page raw --csv > $O/shoot.raw.csv 2>/dev/null
ncu -i $O/shoot.ncu-rep --page source --csv > $O/shoot.source.csv 2>/dev/null
rm -f $O/shoot.ncu-rep; ls -la $O
