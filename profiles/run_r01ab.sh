#!/bin/bash
# Round-1 closing pass (no ncu): GPU suite, smoke, default bench + reference arm, the other workloads with e2e
O=gpurun_out/r01ab; mkdir -p $O
( time timeout 900 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1
timeout 600 python bench.py > $O/cfg2.json 2> $O/cfg2.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/ref.json 2> $O/ref.err
for w in cfg3 cfg4 cfg5; do timeout 600 python bench.py --workload $w --no-cpu-baseline > $O/$w.json 2> $O/$w.err; done
tail -n 3 $O/pytest_gpu.log $O/smoke.log
