#!/bin/bash
# Round-1 "t" GPU pass: full GPU suite (10-bit packed members are the default for N <= 1024), the
# parity tests again with 12-bit forced, smoke, default bench, reference arm, K4 routing sweep, ncu launch list.
O=gpurun_out/r01t; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > $O/gpu.txt
( time timeout 900 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
( time KSPIDER_MEMBER_BITS=12 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_gram.py -m gpu -x -q ) > $O/pytest_gpu_12bit.log 2>&1; echo "rc=$?" >> $O/pytest_gpu_12bit.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1
timeout 600 python bench.py > $O/cfg2.json 2> $O/cfg2.err
KSPIDER_MEMBER_BITS=12 timeout 600 python bench.py --no-cpu-baseline > $O/cfg2_12bit.json 2> $O/cfg2_12bit.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/ref.json 2> $O/ref.err
for g in 16 20 28; do KSPIDER_GRAM=$g timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_gram$g.json 2> $O/cfg2_gram$g.err; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_cfg2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > $O/ncu.log 2>&1
tail -3 $O/pytest_gpu.log $O/pytest_gpu_12bit.log $O/smoke.log; cat $O/cfg2.json
