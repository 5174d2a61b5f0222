import sys, time, numpy as np
sys.path.insert(0, '.')
from kspider_b200 import indexing
rng = np.random.default_rng(31)
n_clades, clade, sketch, core, fam = 200, 50, 5000, 3000, 1000
fam_pools = [rng.integers(1, 2**63, size=2000, dtype=np.uint64) for _ in range(n_clades // 20)]
sk = []
for c in range(n_clades):
    cpool = rng.integers(1, 2**63, size=core, dtype=np.uint64)
    for g in range(clade):
        a = cpool[rng.random(core) < 0.9]
        b = fam_pools[c // 20][rng.random(2000) < 0.25]
        p = rng.integers(1, 2**63, size=max(0, sketch - a.size - b.size), dtype=np.uint64)
        sk.append(np.concatenate([a, b, p]))
t0 = time.perf_counter(); col = indexing.build_colors(sk, device=0); t1 = time.perf_counter()
col = indexing.build_colors(sk, device=0); t2 = time.perf_counter()
tot = sum(s.size for s in sk)
print(f"samples {len(sk)} entries {tot} distinct {col.distinct_hashes} colours {col.counts.size} members {col.members.size} device_ms {col.device_ms:.2f} wall_s {t2-t1:.3f} (first {t1-t0:.3f})  entries/s(device) {tot/col.device_ms*1e3:.3e}")
