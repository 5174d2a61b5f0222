#!/bin/bash
# K1 with 32-bit cell arithmetic (k1_small_c32) + carve-out beside K4
O=gpurun_out/r01x; mkdir -p $O
( timeout 900 python -m pytest tests -m gpu -x -q ) > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
( KSPIDER_K1_CELLS=64 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q ) > $O/pytest_k1_64.log 2>&1; echo "rc=$?" >> $O/pytest_k1_64.log
for k in 3 4; do KSPIDER_K1_CTAS=$k timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_k$k.json 2> $O/cfg2_k$k.err; done
KSPIDER_K1_CELLS=64 KSPIDER_K1_CTAS=3 timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_k3_cells64.json 2> $O/cfg2_k3_cells64.err
KSPIDER_GRAM=26 KSPIDER_K1_CTAS=3 timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_k3_g26.json 2> $O/cfg2_k3_g26.err
KSPIDER_GRAM=20 KSPIDER_K1_CTAS=3 timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_k3_g20.json 2> $O/cfg2_k3_g20.err
for w in cfg3 cfg4 cfg5; do timeout 300 python bench.py --workload $w --no-cpu-baseline --no-e2e > $O/${w}.json 2> $O/${w}.err; done
tail -n 3 $O/pytest.log $O/pytest_k1_64.log
