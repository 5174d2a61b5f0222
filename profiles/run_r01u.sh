#!/bin/bash
# K4 capped at 64 registers (48 used for one weight plane) so K1 CTAs fit beside it: sweep of K1 CTAs per SM
O=gpurun_out/r01u; mkdir -p $O
( timeout 600 python -m pytest tests/test_gpu_gram.py -m gpu -x -q ) > $O/pytest_gram.log 2>&1; echo "rc=$?" >> $O/pytest_gram.log
for k in 8 3 4 5 6; do KSPIDER_K1_CTAS=$k timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_k1ctas$k.json 2> $O/cfg2_k1ctas$k.err; done
for k in 4; do KSPIDER_GRAM=20 KSPIDER_K1_CTAS=$k timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/cfg2_g20_k1ctas$k.json 2> $O/cfg2_g20_k1ctas$k.err; done
for w in cfg3 cfg4 cfg5; do for k in 8 4; do KSPIDER_K1_CTAS=$k timeout 300 python bench.py --workload $w --no-cpu-baseline --no-e2e > $O/${w}_k1ctas$k.json 2> $O/${w}_k1ctas$k.err; done; done
tail -n 3 $O/pytest_gram.log
