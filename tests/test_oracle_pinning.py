"""Pins the oracle (oracle/pairwise_oracle.c, oracle/cluster_oracle.py) before anything trusts it:
  1. against the committed golden vectors made by the UNMODIFIED reference (tests/golden/make_golden.py),
  2. against the reference binary itself, live, on fresh seeded inputs (when oracle/_ref is present),
  3. against the only oracle the reference's own tests define: Python set intersection of the
     sketches (test/generate_golden_files.py:40-49) on a synthetic sourmash-style collection.
CPU only."""
import glob
import gzip
import os

import numpy as np
import pytest

from oracle import cluster_oracle, pairwise_oracle as po, reftools
from util import random_csr

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p)[:-len(".raw.npz")] for p in glob.glob(os.path.join(GOLD, "*.raw.npz")))


def golden_rows(name):
    p = os.path.join(GOLD, name + ".pairwise.tsv")
    f = gzip.open(p + ".gz", "rt") if os.path.exists(p + ".gz") else open(p)
    with f:
        header = f.readline()
        return header, f.read().splitlines()


def golden_input(name):
    """The reference's view of a golden input: last record per narrowed colour id, narrowed counts."""
    z = np.load(os.path.join(GOLD, name + ".raw.npz"))
    counts = {int(k) & 0xFFFFFFFF: int(v) & 0xFFFFFFFF for k, v in zip(z["count_ids"], z["count_vals"])}
    weights = np.array([counts.get(int(c) & 0xFFFFFFFF, 0) for c in z["color_ids"]], dtype=np.uint32)
    return dict(offsets=z["offsets"], members=z["members"], weights=weights,
                total_ids=z["total_ids"], total_vals=z["total_vals"])


def oracle_rows(csr):
    r = po.pairwise_csr(**csr)
    return [f"{a}\t{b}\t{s}\t{po.format_float(x)}\t{po.format_float(y)}\t{po.format_float(z)}"
            for a, b, s, x, y, z in zip(r.src1, r.src2, r.shared, r.cmin, r.cavg, r.cmax)]


def test_cases_present():
    assert "micro_kat" in CASES and len(CASES) >= 5


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_golden(name):
    header, rows = golden_rows(name)
    assert header == "source_1\tsource_2\tshared_kmers\tmin_containment\tavg_containment\tmax_containment\n"
    assert oracle_rows(golden_input(name)) == rows


def test_micro_kat_by_hand():
    """values checkable by hand."""
    _, rows = golden_rows("micro_kat")
    assert rows == ["1\t2\t40\t0.4\t0.6\t0.8", "1\t3\t30\t0.3\t0.35\t0.4", "2\t3\t35\t0.466667\t0.583333\t0.7"]


@pytest.mark.parametrize("name", ["micro_kat", "small_dense"])
def test_oracle_file_loader_on_phmap_bytes(name, tmp_path):
    """The oracle's own phmap-dump reader on files phmap itself wrote (committed fixtures)."""
    ix = po.OracleIndex(os.path.join(GOLD, name))
    r = po.pairwise_index(ix)
    rows = [f"{a}\t{b}\t{s}\t{po.format_float(x)}\t{po.format_float(y)}\t{po.format_float(z)}"
            for a, b, s, x, y, z in zip(r.src1, r.src2, r.shared, r.cmin, r.cavg, r.cmax)]
    assert rows == golden_rows(name)[1]
    out = tmp_path / "seq.tsv"
    ix.write_seq_to_kmers(str(out))
    assert out.read_text() == open(os.path.join(GOLD, name + ".seq.tsv")).read()


@pytest.mark.skipif(not reftools.have_ref(), reason="oracle/_ref not built (needs /root/reference at build time)")
@pytest.mark.parametrize("seed,kw", [
    (101, dict(n_samples=30, n_colors=500, s_lo=1, s_hi=6)),
    (102, dict(n_samples=120, n_colors=800, s_lo=2, s_hi=20, zero_every=11)),
    (103, dict(n_samples=90, n_colors=50, s_lo=2, s_hi=5, big=(64, 65, 80), id_mode="sparse")),
])
def test_oracle_matches_live_reference(seed, kw, tmp_path):
    csr = random_csr(seed, **kw)
    off = csr["offsets"].astype(np.int64)
    base = 2 ** 20
    colors = [(base + c, csr["members"][off[c]:off[c + 1]].tolist()) for c in range(off.size - 1)]
    counts = {base + c: int(w) for c, w in enumerate(csr["weights"])}
    totals = {int(k): int(v) for k, v in zip(csr["total_ids"], csr["total_vals"])}
    raw, prefix = str(tmp_path / "in.raw"), str(tmp_path / "ix")
    reftools.write_raw(raw, totals, counts, colors)
    reftools.run_index_writer_ref(raw, prefix)
    reftools.run_pairwise_ref(prefix, threads=3)
    _, ref_rows = reftools.read_pairwise_tsv_sorted(prefix + "_kSpider_pairwise.tsv")
    assert oracle_rows(csr) == ref_rows
    # and through the oracle's file reader
    r = po.pairwise_index(po.OracleIndex(prefix))
    assert r.src1.size == len(ref_rows)


def test_oracle_vs_set_intersection():
    """The reference's own test oracle (test/generate_golden_files.py:40-49): shared_kmers of a
    pair = |mins_a & mins_b|.  Build colours from synthetic sketches the way an indexer would
    (colour = distinct sample set of a hash, count = number of such hashes) and compare."""
    rng = np.random.default_rng(5)
    pools = [rng.choice(10 ** 9, size=400, replace=False) for _ in range(4)]
    sketches = []
    for i in range(24):
        own = rng.choice(10 ** 9, size=150, replace=False) + 2 * 10 ** 9
        parts = [own] + [p[rng.random(p.size) < 0.5] for p in pools if rng.random() < 0.6]
        sketches.append(set(int(h) for h in np.concatenate(parts)))
    by_hash = {}
    for sid, sk in enumerate(sketches, start=1):
        for h in sk:
            by_hash.setdefault(h, []).append(sid)
    colors = {}
    for h, ms in by_hash.items():
        colors[tuple(ms)] = colors.get(tuple(ms), 0) + 1
    keys = list(colors)
    offsets = np.concatenate([[0], np.cumsum([len(k) for k in keys])]).astype(np.uint64)
    csr = dict(offsets=offsets, members=np.concatenate([np.array(k, np.uint32) for k in keys]),
               weights=np.array([colors[k] for k in keys], np.uint32),
               total_ids=np.arange(1, 25, dtype=np.uint32),
               total_vals=np.array([len(s) for s in sketches], np.uint32))
    r = po.pairwise_csr(**csr)
    got = {(int(a), int(b)): int(s) for a, b, s in zip(r.src1, r.src2, r.shared)}
    want = {}
    for a in range(24):
        for b in range(a + 1, 24):
            n = len(sketches[a] & sketches[b])
            if n:
                want[(a + 1, b + 1)] = n
    assert got == want
    # containments to 3 decimals like test/generate_golden_files.py:76-82
    for (a, b), cmin, cmax in zip(zip(r.src1, r.src2), r.cmin, r.cmax):
        la, lb = len(sketches[a - 1]), len(sketches[b - 1])
        assert abs(float(cmin) - want[(a, b)] / max(la, lb)) < 1e-3
        assert abs(float(cmax) - want[(a, b)] / min(la, lb)) < 1e-3


def test_cluster_oracle_threshold_and_components(tmp_path):
    prefix = str(tmp_path / "kat")
    with open(prefix + ".namesMap", "w") as f:
        f.write("4\n1 a\n2 b\n3 c\n4 d\n")
    with open(prefix + "_kSpider_pairwise.tsv", "w") as f:
        f.write(open(os.path.join(GOLD, "micro_kat.pairwise.tsv")).read())
    # max_containment column: 0.8, 0.4, 0.7
    assert cluster_oracle.cluster(prefix, 0.75) == [["a", "b"], ["c"], ["d"]]
    assert cluster_oracle.cluster(prefix, 0.7) == [["a", "b", "c"], ["d"]]
    assert cluster_oracle.cluster(prefix, 0.0) == [["a", "b", "c"], ["d"]]
    assert cluster_oracle.cluster(prefix, 0.5, "min_cont") == [["a"], ["b"], ["c"], ["d"]]
    assert cluster_oracle.output_name(prefix, 0.95).endswith("_kSpider_clusters_95.0%.tsv")


@pytest.mark.parametrize("name", CASES)
def test_banded_oracle_equals_hash_map_oracle_on_goldens(name):
    """ko_pairwise_csr_banded (the BASELINE-size checker) against the pinned ko_pairwise_csr"""
    csr = golden_input(name)
    a, b = po.pairwise_csr(**csr), po.pairwise_csr_banded(**csr, threads=3)
    for f in ("src1", "src2", "shared"):
        assert np.array_equal(getattr(a, f), getattr(b, f)), f
    for f in ("cmin", "cavg", "cmax"):
        assert np.array_equal(getattr(a, f).view(np.uint32), getattr(b, f).view(np.uint32)), f


@pytest.mark.parametrize("seed,kw", [
    (201, dict(n_samples=300, n_colors=4000, s_lo=1, s_hi=40, big=(70, 150, 290), zero_every=97)),
    (202, dict(n_samples=70, n_colors=900, s_lo=1, s_hi=9, id_mode="sparse", zero_every=5)),
    (203, dict(n_samples=1000, n_colors=20000, s_lo=2, s_hi=6)),
])
def test_banded_oracle_equals_hash_map_oracle_on_seeded_inputs(seed, kw):
    csr = random_csr(seed, **kw)
    a, b = po.pairwise_csr(**csr), po.pairwise_csr_banded(**csr, threads=4)
    assert a.src1.size == b.src1.size
    for f in ("src1", "src2", "shared"):
        assert np.array_equal(getattr(a, f), getattr(b, f)), f
    for f in ("cmin", "cavg", "cmax"):
        assert np.array_equal(getattr(a, f).view(np.uint32), getattr(b, f).view(np.uint32)), f
