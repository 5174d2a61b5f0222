"""Pins oracle/cluster_oracle.py and oracle/ani_oracle.py against the UNMODIFIED reference Python
(pykSpider/kSpider2/ks_clustering.py:22-137,150-163 and ks_pairwise.py:29-82):
  1. against the committed goldens tests/golden/*.clusters.json / *.ani_k*.tsv, which the unmodified
     commands wrote through oracle/refpy.py (tests/golden/make_golden_py.py) -- runs everywhere;
  2. live, where /root/reference exists (this container), on a fresh seeded case.
CPU only.  The GPU path is compared with the same goldens in tests/test_gpu_cluster_golden.py."""
import glob
import os

import numpy as np
import pytest

from oracle import ani_oracle, cluster_oracle, refpy
from util import GOLD, golden_clusters, partition_of_names, stage_golden_prefix

CLUSTER_CASES = sorted(os.path.basename(p)[:-len(".clusters.json")] for p in glob.glob(os.path.join(GOLD, "*.clusters.json")))
ANI_FILES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "*.ani_k*.tsv")))


def test_fixtures_present():
    assert {"micro_kat", "clades", "mid_colors"} <= set(CLUSTER_CASES)
    assert "clades.ani_k21.tsv" in ANI_FILES and "clades.ani_k1.tsv" in ANI_FILES


@pytest.mark.parametrize("name", CLUSTER_CASES)
def test_cluster_oracle_matches_reference_goldens(name, tmp_path):
    prefix = stage_golden_prefix(name, tmp_path)
    gold = golden_clusters(name)
    ani = os.path.join(GOLD, f"{name}.ani_k21.tsv")
    if os.path.exists(ani):
        os.symlink(ani, prefix + "_kSpider_pairwise.ani_col.tsv")
    assert any(d == "ani" for d, _ in gold) == os.path.exists(ani)
    nontrivial = 0
    for (dist, cutoff), want in gold.items():
        got = partition_of_names(cluster_oracle.cluster(prefix, cutoff, dist))
        assert got == want, (name, dist, cutoff)
        nontrivial += 1 < len(want) < sum(len(c) for c in want)
    if name in ("clades", "mid_colors"):
        assert nontrivial >= 6  # the goldens do exercise partitions between "one cluster" and "all singletons"


@pytest.mark.parametrize("fname", ANI_FILES)
def test_ani_oracle_matches_reference_goldens(fname, tmp_path):
    name, k = fname.split(".ani_k")
    ksize = int(k[:-len(".tsv")])
    prefix = stage_golden_prefix(name, tmp_path, ksize)
    got = ani_oracle.ani_column_text(ani_oracle.ani_column_of_tsv(prefix + "_kSpider_pairwise.tsv", ksize))
    assert got == open(os.path.join(GOLD, fname)).read()


def test_micro_kat_clusters_by_hand():
    """max containments 1-2 0.8, 1-3 0.4, 2-3 0.7; sample 4 isolated."""
    g = golden_clusters("micro_kat")
    assert g[("max_cont", 0.7)] == [[1, 2, 3], [4]]
    assert g[("max_cont", 0.8)] == [[1, 2], [3], [4]]
    assert g[("max_cont", 0.95)] == [[1], [2], [3], [4]]
    assert g[("min_cont", 0.5)] == [[1], [2], [3], [4]]  # mins: 0.4, 0.3, 0.466667


@pytest.mark.skipif(not refpy.have_reference_python(), reason="/root/reference is not on this box")
def test_oracles_match_live_reference(tmp_path):
    """Fresh input: the clades generator with another seed, through the unmodified binary and the unmodified CLI."""
    from oracle import reftools
    import sys
    sys.path.insert(0, GOLD)
    from make_golden import clades_csr, from_csr
    if not reftools.have_ref():
        pytest.skip("oracle/_ref not built")
    totals, counts, colors = from_csr(clades_csr(seed=77, n_clades=5, clade=12, n_core=90, n_family=15, n_noise=40), extra_singletons=False)
    prefix = str(tmp_path / "live")
    reftools.write_raw(prefix + ".raw", totals, counts, colors)
    reftools.run_index_writer_ref(prefix + ".raw", prefix)
    n = len(totals)
    with open(prefix + ".namesMap", "w") as f:
        f.write(f"{n}\n" + "".join(f"{i} s{i}\n" for i in range(1, n + 1)))
    with open(prefix + ".extra", "w") as f:
        f.write("31\n1\n1\n\n")
    out = refpy.pairwise(prefix, 2)   # the reference CLI -> _kSpider_internal shim -> unmodified src/pairwise.cpp
    assert "pairwise hashmap construction" in out.stdout
    refpy.estimate_ani(prefix, 1000)
    want_ani = open(prefix + "_kSpider_pairwise.ani_col.tsv").read()
    assert ani_oracle.ani_column_text(ani_oracle.ani_column_of_tsv(prefix + "_kSpider_pairwise.tsv", 31)) == want_ani
    for dist, cutoff in (("max_cont", 0.85), ("ani", 0.99)):
        path = refpy.cluster(prefix, cutoff, dist)
        assert path == cluster_oracle.output_name(prefix, cutoff)
        want = cluster_oracle.read_clusters_file(path)
        assert 1 < len(want) < n
        assert cluster_oracle.cluster(prefix, cutoff, dist) == want


@pytest.mark.parametrize("name", ["clades", "mid_colors", "small_dense"])
def test_float32_threshold_rule_equals_the_text_rule(name, tmp_path):
    """cluster_oracle.kept_mask (used at BASELINE sizes) against ks_clustering.py's text rule row by row"""
    prefix = stage_golden_prefix(name, tmp_path)
    rows = [l.split("\t") for l in open(prefix + "_kSpider_pairwise.tsv").read().splitlines()[1:]]
    for col in (3, 4, 5):
        text = [r[col] for r in rows]
        f32 = np.array([float(t) for t in text], dtype=np.float64).astype(np.float32)
        assert ["%g" % float(v) for v in f32] == text            # the float32 the text came from
        for cutoff in (0.0, 0.004, 0.3, 0.333333, 0.5, 0.95, 1.0, 2.0):
            cut = float(cutoff) * 100
            want = np.array([not (float(t) * 100 < cut) for t in text])
            assert np.array_equal(cluster_oracle.kept_mask(f32, cut), want), (name, col, cutoff)


def test_min_labels_is_the_partition_of_components():
    rng = np.random.default_rng(5)
    a, b = rng.integers(0, 500, 300), rng.integers(0, 500, 300)
    lab = cluster_oracle.min_labels(500, a, b)
    ref = cluster_oracle.components(500, list(zip(a.tolist(), b.tolist())))
    for g in np.unique(ref):
        mem = np.flatnonzero(ref == g)
        assert np.all(lab[mem] == mem.min())
