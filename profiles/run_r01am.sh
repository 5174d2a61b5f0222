#!/bin/bash
# final state of round 1: default bench line, reference arm, launch list of cfg4 (plain run first)
O=gpurun_out/r01am; mkdir -p $O
timeout 600 python bench.py > $O/cfg2.json 2> $O/cfg2.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/ref.json 2> $O/ref.err
timeout 600 python bench.py --workload cfg4 --no-cpu-baseline > $O/cfg4.json 2> $O/cfg4.err
CMD="python bench.py --workload cfg4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_cfg4.csv $CMD > $O/ncu_launches.log 2>&1
ls $O
