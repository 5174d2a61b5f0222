"""Parity at the size of BASELINE.json's configs, row for row (not checksums): the CUDA path through the C-ABI
against the C oracle's BASELINE-size accumulator (ko_pairwise_csr_banded, pinned against the reference-pinned hash
map oracle in tests/test_oracle_pinning.py), integers equal and float32 containments equal by bit pattern; and
the cluster labels of every dist type at cut-offs {0, 0.3, 0.95, 1} (+ one that splits the config non-trivially)
against the reference-pinned cluster oracle (oracle/cluster_oracle.py: kept_mask + connected components).

cfg2  1 000 samples, 10^7 colours (the benched config): also through the FILES -- index files written, the drop-in
      kSpider::pairwise on them, TSV text parsed back, `Clusters` on the TSV;
cfg3  10 000 genomes in clades (dense 400 MB triangle in the reference's terms);
cfg4  100 000 samples, 5*10^7 rows;
cfg5  N = 10 000, 16 giant colours of 4 000 members with weights up to 2^32 - 1 (all four byte planes of K4)."""
import os

import numpy as np
import pytest

from kspider_b200 import synth
from oracle import cluster_oracle, pairwise_oracle as po

pytestmark = pytest.mark.gpu
DIST = {"min_cont": "cmin", "avg_cont": "cavg", "max_cont": "cmax"}
REC = {"min_cont": "min_containment", "avg_cont": "avg_containment", "max_cont": "max_containment"}


@pytest.fixture(scope="module")
def ksb(built_lib):
    from kspider_b200.capi import check
    check(built_lib.ksb_cuda_available())
    import kspider_b200
    return kspider_b200


def oracle_of(s):
    return po.pairwise_csr_banded(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, threads=min(32, os.cpu_count() or 1))


def assert_rows_equal(rec, want):
    assert rec.size == want.src1.size, (rec.size, want.src1.size)
    assert np.array_equal(rec["source_1"], want.src1) and np.array_equal(rec["source_2"], want.src2)
    assert np.array_equal(rec["shared_kmers"], want.shared)
    for f, g in (("min_containment", "cmin"), ("avg_containment", "cavg"), ("max_containment", "cmax")):
        assert np.array_equal(rec[f].view(np.uint32), getattr(want, g).view(np.uint32)), f


def check_clusters(res, n, want, cutoffs):
    """labels (= smallest member of the component) of every dist type against the cluster oracle"""
    a = (want.src1 - 1).astype(np.int64)   # ids are 1..N: compact index = node = id - 1 (ks_clustering.py:99-100)
    b = (want.src2 - 1).astype(np.int64)
    seen_nontrivial = False
    for dist, col in DIST.items():
        for cutoff in cutoffs:
            keep = cluster_oracle.kept_mask(getattr(want, col), float(cutoff) * 100)
            ref = cluster_oracle.min_labels(n, a[keep], b[keep])
            labels, n_edges = res.cluster(n, dist, cutoff)
            assert n_edges == int(keep.sum()), (dist, cutoff)
            assert np.array_equal(labels, ref), (dist, cutoff)
            k = np.unique(ref).size
            seen_nontrivial |= 1 < k < n
    assert seen_nontrivial, "no cut-off produced a partition between one cluster and all singletons"


def check_fused(idx, n, want, cutoffs, **opts):
    """the cluster-only step (ksb_device_cluster: K56 fused scan, no records) on the same index"""
    a, b = (want.src1 - 1).astype(np.int64), (want.src2 - 1).astype(np.int64)
    dev = idx.upload(**opts)
    for dist, cutoff in [("max_cont", c) for c in cutoffs] + [("min_cont", cutoffs[1]), ("avg_cont", cutoffs[2])]:
        keep = cluster_oracle.kept_mask(getattr(want, DIST[dist]), float(cutoff) * 100)
        labels, ne, nr = dev.cluster(n, dist, cutoff)
        assert (ne, nr) == (int(keep.sum()), want.src1.size), (dist, cutoff)
        assert np.array_equal(labels, cluster_oracle.min_labels(n, a[keep], b[keep])), (dist, cutoff)
    dev.close()


def run_config(ksb, s, cutoffs, **opts):
    want = oracle_of(s)
    idx = ksb.Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=min(32, os.cpu_count() or 1))
    assert idx.info.n_updates == s.n_updates
    res = idx.pairwise(want_records=True, **opts)
    assert_rows_equal(res.records(), want)
    check_clusters(res, s.n_samples, want, cutoffs)
    res.close()
    check_fused(idx, s.n_samples, want, cutoffs, **opts)
    return idx, res, want


def test_cfg2_full_size_rows_and_labels(ksb):
    s = synth.zipf_index(1000, 10_000_000, seed=12345, name="cfg2")
    idx, res, want = run_config(ksb, s, (0.0, 0.0257, 0.0258, 0.3, 0.95, 1.0))
    assert idx.info.n_gram_colors > 0 and idx.info.n_small_colors > 0   # K1 and K4 both carried the result


def test_cfg2_full_size_through_the_files(ksb, tmp_path, capfd):
    """the benched config end to end as a user of the reference runs it: index files -> `pairwise` -> TSV -> `cluster`"""
    from kspider_b200 import kSpider_internal
    from kspider_b200.clustering import Clusters
    s = synth.zipf_index(1000, 10_000_000, seed=12345, name="cfg2")
    prefix = str(tmp_path / "cfg2")
    synth.write_index_files(s, prefix, with_names=True)
    kSpider_internal.pairwise(index_prefix=prefix, user_threads=min(32, os.cpu_count() or 1))
    capfd.readouterr()
    want = oracle_of(s)
    rows = open(prefix + "_kSpider_pairwise.tsv").read().splitlines()
    assert rows[0] == "source_1\tsource_2\tshared_kmers\tmin_containment\tavg_containment\tmax_containment"
    assert len(rows) - 1 == want.src1.size
    step = 997   # every 997th row as text (formatting 5e5 rows with the oracle's snprintf takes a while), all rows as numbers
    for k in range(0, want.src1.size, step):
        assert rows[1 + k] == (f"{want.src1[k]}\t{want.src2[k]}\t{want.shared[k]}\t{po.format_float(float(want.cmin[k]))}\t"
                               f"{po.format_float(float(want.cavg[k]))}\t{po.format_float(float(want.cmax[k]))}")
    cols = np.array([r.split("\t") for r in rows[1:]])
    assert np.array_equal(cols[:, 0].astype(np.uint32), want.src1) and np.array_equal(cols[:, 1].astype(np.uint32), want.src2)
    assert np.array_equal(cols[:, 2].astype(np.uint64), want.shared)
    for c, f in ((3, "cmin"), (4, "cavg"), (5, "cmax")):   # "%g" keeps 6 digits: the text is the float32 rounded to them
        assert np.array_equal(cols[:, c].astype(np.float64), np.array(["%g" % float(v) for v in getattr(want, f)]).astype(np.float64))

    class Log:
        def INFO(self, m): pass
        def ERROR(self, m): raise SystemExit(1)
    a, b = (want.src1 - 1).astype(np.int64), (want.src2 - 1).astype(np.int64)
    for dist, cutoff in (("max_cont", 0.0258), ("min_cont", 0.0257), ("avg_cont", 0.3)):
        kcl = Clusters(logger_obj=Log(), index_prefix=prefix, cut_off_threshold=float(cutoff) * 100, dist_type=dist)
        kcl.construct_graph()
        kcl.cluster_graph()
        keep = cluster_oracle.kept_mask(getattr(want, DIST[dist]), float(cutoff) * 100)
        assert kcl.n_edges == int(keep.sum())
        assert np.array_equal(kcl.labels, cluster_oracle.min_labels(1000, a[keep], b[keep]))
        lines = open(kcl.output).read().splitlines()
        assert sorted(n for l in lines for n in l.split(",")) == sorted(f"sample_{i}" for i in range(1, 1001))
        assert len(lines) == np.unique(kcl.labels).size


def test_cfg3_full_size_rows_and_labels(ksb):
    run_config(ksb, synth.cfg3(), (0.0, 0.3, 0.55, 0.95, 1.0))


def test_cfg4_full_size_rows_and_labels(ksb):
    run_config(ksb, synth.cfg4(), (0.0, 0.3, 0.55, 0.95, 1.0))


def test_cfg4_full_size_spill_path_rows(ksb):
    """cfg4 forced onto the K3 sort-reduce path with a 1 GB dense budget -- the same 5*10^7 rows"""
    s = synth.cfg4()
    want = oracle_of(s)
    idx = ksb.Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=min(32, os.cpu_count() or 1))
    res = idx.pairwise(want_records=True, dense_budget_bytes=10 ** 9, spill="always")
    assert_rows_equal(res.records(), want)


def test_cfg5_giant_colours_four_planes(ksb):
    s = synth.giant_index(n_samples=10_000, n_background=1_000_000, n_giant=16, giant_size=4000)
    idx, res, want = run_config(ksb, s, (0.0, 0.3, 0.95, 1.0, 0.000002))
    assert idx.info.n_gram_colors >= 16 and idx.info.gram_planes == 4
