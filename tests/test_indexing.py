"""Colour construction (BASELINE configs[0]: sourmash signatures -> index -> pairwise).

CPU part: the restatement of the reference's incremental colouring loop (oracle/indexing_oracle.py)
reaches the end state its definition gives.  GPU part: ksb_colors_build equals that end state, and
the whole chain — synthetic .sig files -> sourmash_sigs_indexing -> pairwise — is checked against the
reference's own golden procedure (test/generate_golden_files.py: set intersections, containments to
3 decimals; compared like test/validate.py) and, row for row, against the UNMODIFIED reference
pairwise binary reading the files we wrote."""
import gzip
import json
import os

import numpy as np
import pytest

from oracle import indexing_oracle as io


def _sketches(seed, n=14, pools=3, pool_size=300, own=60, dup=False, empty=()):
    rng = np.random.default_rng(seed)
    ps = [rng.choice(10 ** 12, size=pool_size, replace=False) for _ in range(pools)]
    out = []
    for i in range(n):
        if i in empty:
            out.append(np.zeros(0, dtype=np.uint64))
            continue
        mine = rng.choice(10 ** 12, size=own, replace=False) + 2 * 10 ** 12
        parts = [mine] + [p[rng.random(p.size) < 0.5] for p in ps if rng.random() < 0.7]
        h = np.concatenate(parts).astype(np.uint64)
        if dup:
            h = np.concatenate([h, h[: h.size // 3]])
        rng.shuffle(h)
        out.append(h)
    return out


@pytest.mark.parametrize("seed,kw", [(1, {}), (2, dict(dup=True)), (3, dict(n=30, pools=6)), (4, dict(empty=(0, 5)))])
def test_incremental_restatement_reaches_the_defined_end_state(seed, kw):
    sk = [[int(v) for v in s] for s in _sketches(seed, **kw)]
    legend, cc, totals = io.incremental(sk)
    assert io.end_state(legend, cc) == io.by_definition(sk)
    assert [totals[g] for g in range(1, len(sk) + 1)] == [len(s) for s in sk]
    # colorsCount of real colours never goes negative and adds up to the number of distinct hashes; key 0
    # ("no colour yet") is decremented once per new hash by the reference (:247, a wrapped u64 in its file)
    assert all(v >= 0 for c, v in cc.items() if c != 0)
    assert sum(v for c, v in cc.items() if c != 0) == len(set(h for s in sk for h in s))


def _as_dict(colors):
    out = {}
    off = colors.offsets.astype(np.int64)
    for c in range(off.size - 1):
        key = tuple(int(v) + 1 for v in colors.members[off[c]:off[c + 1]])
        assert list(key) == sorted(set(key))
        out[key] = out.get(key, 0) + int(colors.counts[c])
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("seed,kw", [(11, {}), (12, dict(dup=True)), (13, dict(n=40, pools=8, pool_size=2000)),
                                     (14, dict(empty=(0, 3, 13))), (15, dict(n=1)), (16, dict(n=300, pools=2, own=5, pool_size=50))])
def test_gpu_colour_table_equals_the_end_state(built_lib, seed, kw):
    from kspider_b200 import indexing
    sk = _sketches(seed, **kw)
    col = indexing.build_colors(sk, device=0)
    want = io.by_definition([[int(v) for v in s] for s in sk])
    assert _as_dict(col) == want
    assert col.distinct_hashes == sum(want.values())
    assert [int(v) for v in col.totals] == [s.size for s in sk]


@pytest.mark.gpu
def test_gpu_colour_table_edge_cases(built_lib):
    from kspider_b200 import indexing
    assert indexing.build_colors([], device=0).counts.size == 0
    assert indexing.build_colors([np.zeros(0, np.uint64)] * 3, device=0).counts.size == 0
    full = np.array([0, 1, 2 ** 64 - 1, 2 ** 63], dtype=np.uint64)   # extreme hash values, one shared set
    col = indexing.build_colors([full, full[::-1].copy(), full], device=0)
    assert _as_dict(col) == {(1, 2, 3): 4}


def _write_sigs(d, sketches, ksize=25):
    names = []
    for i, s in enumerate(sketches):
        name = f"genome_{i:03d}"
        doc = [{"class": "sourmash_signature", "signatures": [
            {"ksize": ksize + 6, "mins": [1, 2, 3]},                      # another k: must be skipped
            {"ksize": ksize, "mins": [int(v) for v in s], "molecule": "dna"}]}]
        path = os.path.join(d, name + ".sig")
        if i % 5 == 4:   # sourmash also writes gzip-compressed .sig files; the reference reads them through zstr
            with gzip.open(path, "wt") as f:
                json.dump(doc, f)
        else:
            with open(path, "w") as f:
                json.dump(doc, f)
        names.append(name)
    with open(os.path.join(d, "README.txt"), "w") as f:                   # not a signature: ignored
        f.write("x")
    return names


@pytest.mark.gpu
def test_signatures_to_pairwise_against_the_reference_golden_procedure(built_lib, tmp_path, monkeypatch, capfd):
    """BASELINE configs[0]."""
    from kspider_b200 import kSpider_internal
    from kspider_b200.clustering import cluster
    sk = _sketches(21, n=24, pools=5, pool_size=800, own=200)
    sigs = tmp_path / "sigs"
    sigs.mkdir()
    names = _write_sigs(str(sigs), sk)
    monkeypatch.chdir(tmp_path)
    kSpider_internal.sourmash_sigs_indexing(str(sigs) + "/", 25)
    prefix = str(tmp_path / "sigs")
    for suf in ("_color_to_sources.bin", "_color_count.bin", "_groupID_to_kmerCount.bin", ".namesMap", ".extra"):
        assert os.path.getsize(prefix + suf) > 0
    kSpider_internal.pairwise(index_prefix=prefix, user_threads=2)
    capfd.readouterr()
    # ---- golden, test/generate_golden_files.py
    mins = {n: set(int(v) for v in s) for n, s in zip(names, sk)}
    id_to_name = {}
    with open(prefix + ".namesMap") as f:
        assert int(f.readline()) == len(names)
        for line in f:
            i, nm = line.split()
            id_to_name[int(i)] = nm
    assert sorted(id_to_name.values()) == sorted(names)
    # lengths (validate.py checks seqToKmersNo against golden_sig_to_len)
    with open(prefix + "_kSpider_seqToKmersNo.tsv") as f:
        assert f.readline() == "ID\tseq\tkmers\n"
        lens = {id_to_name[int(l.split("\t")[1])]: int(l.split("\t")[2]) for l in f}
    assert lens == {n: len(m) for n, m in mins.items()}
    got = {}
    with open(prefix + "_kSpider_pairwise.tsv") as f:
        assert f.readline() == "source_1\tsource_2\tshared_kmers\tmin_containment\tavg_containment\tmax_containment\n"
        for line in f:
            a, b, sh, cmin, cavg, cmax = line.rstrip("\n").split("\t")
            got[frozenset((id_to_name[int(a)], id_to_name[int(b)]))] = (int(sh), float(cmin), float(cavg), float(cmax))
    want = {}
    for i, a in enumerate(names):
        for b in names[i + 1:]:
            common = len(mins[a] & mins[b])
            if common:
                mx = common / min(len(mins[a]), len(mins[b]))
                mn = common / max(len(mins[a]), len(mins[b]))
                want[frozenset((a, b))] = (common, mn, (mx + mn) / 2, mx)
    assert set(got) == set(want)
    for k, (sh, mn, av, mx) in want.items():
        g = got[k]
        assert g[0] == sh
        assert abs(g[1] - mn) < 1e-3 and abs(g[2] - av) < 1e-3 and abs(g[3] - mx) < 1e-3
    # ---- the unmodified reference binary on the files we wrote: identical TSV rows
    from oracle import reftools
    if reftools.have_ref():
        ours = reftools.read_pairwise_tsv_sorted(prefix + "_kSpider_pairwise.tsv")
        ref_prefix = str(tmp_path / "ref")
        for suf in ("_color_to_sources.bin", "_color_count.bin", "_groupID_to_kmerCount.bin"):
            os.link(prefix + suf, ref_prefix + suf)
        reftools.run_pairwise_ref(ref_prefix, threads=2, timeout=600)
        assert reftools.read_pairwise_tsv_sorted(ref_prefix + "_kSpider_pairwise.tsv") == ours
    # ---- cluster on top (single linkage at 20 % max containment), all samples present exactly once
    out = cluster(prefix, 0.2, "max_cont")
    assert os.path.basename(out) == "sigs_kSpider_clusters_20.0%.tsv"
    seen = [n for line in open(out) for n in line.strip().split(",")]
    assert sorted(seen) == sorted(names)


@pytest.mark.gpu
def test_mixed_directory_numbers_samples_like_the_reference(built_lib, tmp_path, monkeypatch):
    """src/sourmash_indexing.cpp:85-117 gives an id (and a namesMap line) to EVERY *.sig and *.gz file, in glob order,
    but parses only *.sig (:152); a .sig without a signature of the requested ksize gets no k-mer count (:176-178,282-285)."""
    from kspider_b200 import kSpider_internal
    sk = _sketches(31, n=6, pools=2, pool_size=200, own=40)
    d = tmp_path / "mix"
    d.mkdir()
    _write_sigs(str(d), sk)                                     # genome_000.sig .. genome_005.sig (+ README.txt)
    with gzip.open(str(d / "genome_002b.sig.gz"), "wt") as f:   # extension "gz": named "genome_002b.sig", numbered, never parsed
        json.dump([{"signatures": [{"ksize": 25, "mins": [1, 2, 3]}]}], f)
    with open(str(d / "genome_004b.sig"), "w") as f:            # parsed, but holds no ksize-25 signature
        json.dump([{"signatures": [{"ksize": 31, "mins": [7, 8]}]}], f)
    monkeypatch.chdir(tmp_path)
    kSpider_internal.sourmash_sigs_indexing(str(d), 25)
    prefix = str(tmp_path / "mix")
    lines = open(prefix + ".namesMap").read().splitlines()
    want_names = ["genome_000", "genome_001", "genome_002", "genome_002b.sig", "genome_003", "genome_004", "genome_004b", "genome_005"]
    assert lines[0] == "8" and [l.split() for l in lines[1:]] == [[str(i + 1), n] for i, n in enumerate(want_names)]
    kSpider_internal.pairwise(index_prefix=prefix, user_threads=2)
    seq = [l.split("\t") for l in open(prefix + "_kSpider_seqToKmersNo.tsv").read().splitlines()[1:]]
    got = {int(r[1]): int(r[2]) for r in seq}
    assert got == {1: sk[0].size, 2: sk[1].size, 3: sk[2].size, 5: sk[3].size, 6: sk[4].size, 8: sk[5].size}  # ids 4 and 7: no count
    pairs = {(int(l.split("\t")[0]), int(l.split("\t")[1])) for l in open(prefix + "_kSpider_pairwise.tsv").read().splitlines()[1:]}
    assert pairs and all(a not in (4, 7) and b not in (4, 7) for a, b in pairs)
