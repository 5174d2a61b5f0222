#!/bin/bash
# Round-1 closing pass (ncu): default bench (plain), then the launch list and one --set full capture of the same command
O=gpurun_out/r01ac; mkdir -p $O
timeout 600 python bench.py > $O/cfg2.json 2> $O/cfg2.err
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_cfg2.csv $CMD > $O/ncu_launches.log 2>&1
$CMD > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k1_small|k4_gram|kb_build|k5_emit|k5_count' -s 14 -c 12 -o $O/prof_cfg2 $CMD > $O/ncu_full.log 2>&1
ncu -i $O/prof_cfg2.ncu-rep --page raw --csv > $O/cfg2_ncu_raw.csv 2>/dev/null
ls -la $O
