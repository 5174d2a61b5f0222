"""GPU drop-ins of `kSpider cluster` and `kSpider pairwise --estimate-ani` against what the UNMODIFIED reference
Python wrote (tests/golden/*.clusters.json, *.ani_k*.tsv: ks_clustering.py / ks_pairwise.py run through
oracle/refpy.py on the unmodified src/pairwise.cpp's TSVs -- tests/golden/make_golden_py.py).

The cluster test executes the body of the reference's command (ks_clustering.py:155-163) verbatim with ONLY the
class swapped, which is the edit INTEGRATION.md section 3 describes."""
import glob
import os

import numpy as np
import pytest

from util import GOLD, golden_clusters, partition_of_names, stage_golden_prefix

pytestmark = pytest.mark.gpu

CLUSTER_CASES = sorted(os.path.basename(p)[:-len(".clusters.json")] for p in glob.glob(os.path.join(GOLD, "*.clusters.json")))
ANI_FILES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "*.ani_k*.tsv")))


@pytest.fixture(scope="module")
def ksb(built_lib):
    from kspider_b200.capi import check
    check(built_lib.ksb_cuda_available())
    import kspider_b200
    return kspider_b200


class Logger:
    """ctx.obj of the reference's CLI (kSpider2/customLogger.py): INFO is silent by default, ERROR exits"""
    def __init__(self):
        self.lines = []

    def INFO(self, msg):
        self.lines.append(msg)

    SUCCESS = WARNING = INFO

    def ERROR(self, msg):
        raise SystemExit(1)


def reference_main_body(Clusters, ctx_obj, index_prefix, cutoff, distance_type):
    """ks_clustering.py:155-163, verbatim but for `ctx.obj` -> ctx_obj and the class being a parameter"""
    cutoff = float(cutoff) * 100
    kCl = Clusters(logger_obj=ctx_obj, index_prefix=index_prefix,
                   cut_off_threshold=cutoff, dist_type=distance_type)
    ctx_obj.INFO("Building the main graph...")
    kCl.construct_graph()
    ctx_obj.INFO("Clustering...")
    kCl.cluster_graph()
    return kCl


def read_partition(path):
    with open(path) as f:
        return partition_of_names(line.rstrip("\n").split(",") for line in f if line.strip())


@pytest.mark.parametrize("name", CLUSTER_CASES)
@pytest.mark.parametrize("source", ["dropin+sidecar", "dropin+text", "reference_tsv"])
def test_cluster_dropin_equals_reference_goldens(ksb, name, source, tmp_path, monkeypatch, capfd):
    """Every dist type x cut-off of the goldens.  source: the TSV on disk is ours with its binary side-car, ours with
    the side-car switched off (text parse), or the REFERENCE's own TSV (no side-car exists)."""
    from kspider_b200 import kSpider_internal
    from kspider_b200.clustering import Clusters
    prefix = stage_golden_prefix(name, tmp_path, with_index_files=True)
    if source != "reference_tsv":
        kSpider_internal.pairwise(index_prefix=prefix, user_threads=2)  # overwrites the staged TSV with ours
        assert os.path.exists(prefix + "_kSpider_pairwise.ksb")
        if source == "dropin+text":
            monkeypatch.setenv("KSPIDER_SIDECAR", "0")
    gold = golden_clusters(name)
    ani = os.path.join(GOLD, f"{name}.ani_k21.tsv")
    if os.path.exists(ani):
        # the reference's ani column lines up with ITS TSV, which is sorted like ours (goldens are stored sorted)
        os.symlink(ani, prefix + "_kSpider_pairwise.ani_col.tsv")
    for (dist, cutoff), want in sorted(gold.items()):
        log = Logger()
        kcl = reference_main_body(Clusters, log, prefix, cutoff, dist)
        assert kcl.output == prefix + f"_kSpider_clusters_{float(cutoff) * 100}%.tsv"
        assert read_partition(kcl.output) == want, (name, source, dist, cutoff)
        assert kcl._table.source == ("sidecar" if source == "dropin+sidecar" and dist != "ani" else "tsv")
        assert log.lines[0] == "Loading TSV pairwise file" and f"number of clusters: {len(want)}" in log.lines
        assert sorted(map(sorted, kcl.connected_components)) == sorted(sorted(i - 1 for i in c) for c in want)
    assert "[i] constructing graph" in capfd.readouterr().out


def test_cluster_dropin_error_behaviour(ksb, tmp_path):
    from kspider_b200.capi import KsbError
    from kspider_b200.clustering import Clusters
    prefix = stage_golden_prefix("micro_kat", tmp_path)
    with pytest.raises(SystemExit):   # ks_clustering.py:37-38 "unknown distance!"
        Clusters(Logger(), prefix, 30.0, "max_containment")
    with pytest.raises(SystemExit):   # :41-43 ani selected without the ani file
        Clusters(Logger(), prefix, 30.0, "ani")
    os.remove(prefix + "_kSpider_seqToKmersNo.tsv")
    with pytest.raises(FileNotFoundError):  # :35,49-50
        Clusters(Logger(), prefix, 30.0, "max_cont")
    # a kept row naming a node the namesMap did not create: rustworkx refuses the edge, so do we
    prefix = stage_golden_prefix("micro_kat", tmp_path)
    open(prefix + ".namesMap", "w").write("2\n1 a\n2 b\n")
    k = Clusters(Logger(), prefix, 0.0, "max_cont")
    k.construct_graph()
    with pytest.raises(KsbError):
        k.cluster_graph()
    k = Clusters(Logger(), prefix, 75.0, "max_cont")  # only row 1-2 (0.8) is kept: fine
    k.construct_graph()
    k.cluster_graph()
    assert open(k.output).read() == "a,b\n"


@pytest.mark.parametrize("fname", ANI_FILES)
@pytest.mark.parametrize("source", ["sidecar", "reference_tsv"])
def test_estimate_ani_dropin_equals_reference_goldens(ksb, fname, source, tmp_path):
    from kspider_b200 import kSpider_internal
    from kspider_b200.ani import estimate_ani
    name, k = fname.split(".ani_k")
    ksize = int(k[:-len(".tsv")])
    prefix = stage_golden_prefix(name, tmp_path, ksize, with_index_files=True)
    if source == "sidecar":
        kSpider_internal.pairwise(index_prefix=prefix, user_threads=2)
    out = estimate_ani(prefix, 1000, user_threads=2)
    got = open(out).read().splitlines()
    want = open(os.path.join(GOLD, fname)).read().splitlines()
    assert got[0] == want[0] == "avg_ani" and len(got) == len(want)
    if ksize == 1:
        assert got == want  # c ** 1.0 == c: no pow, the text round trip alone -- bit-exact
    else:
        g, w = np.array([float(x) for x in got[1:]]), np.array([float(x) for x in want[1:]])
        same = np.count_nonzero(g == w)
        assert same >= 0.99 * g.size, f"{same} of {g.size} identical"
        assert np.max(np.abs(g - w)) <= 2 * 2.0 ** -53  # double-double pow vs glibc pow: at most an ulp of a value < 1
