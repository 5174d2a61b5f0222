#!/bin/bash
# ncu of the clustering kernels on cfg4
O=gpurun_out/r01ai; mkdir -p $O
CMD="python bench.py --workload cfg4 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > $O/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k67_union|k6_hook_min|k7_' -c 5 -o $O/prof_cfg4_k6 $CMD > $O/ncu.log 2>&1
ncu -i $O/prof_cfg4_k6.ncu-rep --page raw --csv > $O/cfg4_k6_raw.csv 2>/dev/null
ncu -i $O/prof_cfg4_k6.ncu-rep --page details > $O/cfg4_k6_details.txt 2>/dev/null
ls -la $O
