#!/bin/bash
# K1 with the max-shared carve-out beside K4 (so K4's CTAs can become resident while K1 runs)
O=gpurun_out/r01w; mkdir -p $O
( timeout 600 python -m pytest tests/test_gpu_gram.py tests/test_gpu_parity.py -m gpu -x -q ) > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
for k in 4 3 5 8; do KSPIDER_K1_CTAS=$k timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_k$k.json 2> $O/cfg2_k$k.err; done
KSPIDER_CARVEOUT=0 timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_nocarve.json 2> $O/cfg2_nocarve.err
KSPIDER_GRAM=20 timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_g20.json 2> $O/cfg2_g20.err
KSPIDER_GRAM=16 timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_g16.json 2> $O/cfg2_g16.err
for w in cfg3 cfg4 cfg5; do timeout 300 python bench.py --workload $w --no-cpu-baseline --no-e2e > $O/${w}.json 2> $O/${w}.err; done
timeout 300 python bench.py --no-cpu-baseline > $O/cfg2_e2e.json 2> $O/cfg2_e2e.err
tail -n 3 $O/pytest.log
