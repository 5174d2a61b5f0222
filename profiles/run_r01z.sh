#!/bin/bash
# K1 c32 with per-class cached head (40 registers, 6 CTAs/SM)
O=gpurun_out/r01z; mkdir -p $O
( timeout 900 python -m pytest tests -m gpu -x -q ) > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
for k in 3 2 4; do KSPIDER_K1_CTAS=$k timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_k$k.json 2> $O/cfg2_k$k.err; done
KSPIDER_GRAM=20 timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_g20.json 2> $O/cfg2_g20.err
KSPIDER_GRAM=26 timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_g26.json 2> $O/cfg2_g26.err
for w in cfg3 cfg4 cfg5; do timeout 300 python bench.py --workload $w --no-cpu-baseline --no-e2e > $O/${w}.json 2> $O/${w}.err; done
tail -n 3 $O/pytest.log
