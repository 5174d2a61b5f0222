import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
from kspider_b200 import Index, synth
s = synth.cfg4()
idx = Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=16)
N = s.n_samples
whole, ne, nr = idx.upload().cluster(N, "max_cont", 0.3)
print("whole", ne, nr, flush=True)
for world in (int(a) for a in sys.argv[1:]):
    for r in range(world):
        sh = idx.shard(r, world, 8)
        dev = sh.upload()
        for rep in range(3):
            lab, e, n = dev.cluster(N, "max_cont", 0.3)
        print("world", world, "rank", r, "ok", e, n, flush=True)
        dev.close()
