#!/bin/bash
# 2-GPU check: the multi-device tests, cfg2 and cfg4 at N=2 through torchrun
O=gpurun_out/r01ad; mkdir -p $O
( timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two or multi or shard or drop" ) > $O/pytest_2gpu.log 2>&1; echo "rc=$?" >> $O/pytest_2gpu.log
for w in cfg2 cfg4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --workload $w > $O/${w}_n2.json 2> $O/${w}_n2.err
done
tail -n 3 $O/pytest_2gpu.log; tail -c 600 $O/cfg2_n2.err
