"""One rank's cluster-only step of a row-block shard, on one GPU (for ncu launch lists of the multi-GPU step):
    python profiles/tools/shard_step.py <workload> <world> <rank> [steps]"""
import os
import sys
import time
sys.path.insert(0, os.getcwd())
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench  # noqa: E402
from kspider_b200 import Index  # noqa: E402

name, world, rank = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
s = bench.make_workload(name)
idx = Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=16)
sh = idx.shard(rank, world, 8) if world > 1 else idx
dev = sh.upload()
dev.set_graph(os.environ.get("KSPIDER_GRAPH", "1") != "0")
cut = bench.DEFAULT_CUTOFF[name]
for _ in range(3):
    dev.cluster(s.n_samples, "max_cont", cut)
t0 = time.perf_counter()
for _ in range(steps):
    lab, ne, nr = dev.cluster(s.n_samples, "max_cont", cut)
print(f"{name} world {world} rank {rank}: {(time.perf_counter() - t0) / steps * 1e3:.3f} ms/step, edges {ne}, rows {nr}, payload {sh.info.payload_bytes / 1e6:.1f} MB")
