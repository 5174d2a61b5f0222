"""Where the end-to-end call (ksb_pairwise_cluster on the pinned host index) spends its host time:
    KSPIDER_SETUP_TIMERS=1 python profiles/tools/e2e_trace.py <workload> [calls]
prints the library's set-up phases (stderr) and the wall time of every call."""
import os
import sys
import time
sys.path.insert(0, os.getcwd())
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench  # noqa: E402
from kspider_b200 import Index  # noqa: E402

name = sys.argv[1]
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 6
s = bench.make_workload(name)
idx = Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=16)
cut = bench.DEFAULT_CUTOFF[name]
for k in range(calls):
    sys.stderr.write(f"--- call {k}\n")
    sys.stderr.flush()
    t0 = time.perf_counter()
    lab, ne, nr, inf = idx.pairwise_cluster("max_cont", cut)
    dt = (time.perf_counter() - t0) * 1e3
    t = inf.t
    print(f"call {k}: {dt:.3f} ms  setup {t.host_setup_ms:.3f} step {t.host_accumulate_ms:.3f} free {t.host_free_ms:.3f} | h2d {t.h2d_ms:.3f} "
          f"acc {t.accumulate_ms:.3f} post {t.post_ms:.3f} cluster {t.cluster_ms:.3f}  edges {ne} rows {nr} payload {idx.info.payload_bytes / 1e6:.1f} MB", flush=True)
