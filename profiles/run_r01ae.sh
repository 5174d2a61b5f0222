#!/bin/bash
# K6 with a per-CTA edge queue
O=gpurun_out/r01ae; mkdir -p $O
( timeout 900 python -m pytest tests -m gpu -x -q ) > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
for w in cfg4 cfg3 cfg2; do timeout 300 python bench.py --workload $w --no-cpu-baseline > $O/$w.json 2> $O/$w.err; done
tail -n 3 $O/pytest.log
