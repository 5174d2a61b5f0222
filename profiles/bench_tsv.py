"""f2: throughput of the byte-identical TSV writer (ksb_write_tsv) on cfg3's 5 M rows."""
import os, sys, time, tempfile
sys.path.insert(0, '.')
from kspider_b200 import synth, Index
s = synth.cfg3()
idx = Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=16)
res = idx.pairwise(want_records=True)
td = tempfile.mkdtemp()
for th in (1, 4, 16):
    p = os.path.join(td, f"out{th}.tsv")
    t0 = time.perf_counter(); res.write_tsv(p, host_threads=th); t1 = time.perf_counter()
    print(f"threads {th}: {res.info.n_rows} rows, {os.path.getsize(p)/1e6:.0f} MB in {t1-t0:.3f} s = {res.info.n_rows/(t1-t0)/1e6:.1f} M rows/s")
