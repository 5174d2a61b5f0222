"""Shared helpers of the test-suite: seeded random indexes and result comparison."""
from __future__ import annotations

import gzip
import json
import os
import shutil

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def random_csr(seed: int, n_samples: int, n_colors: int, s_lo: int = 1, s_hi: int = 12, id_mode: str = "dense",
               w_hi: int = 1000, zero_every: int = 0, big: tuple = ()):
    """Random colour table.  id_mode: "dense" ids 1..N, "sparse" arbitrary u32 ids.
    big = extra colour sizes appended (e.g. (70, 300, 1500)).  Members within a colour are shuffled
    (files store them in hash-slot order).  Returns dict with offsets/members/weights/total_ids/total_vals."""
    rng = np.random.default_rng(seed)
    if id_mode == "dense":
        ids = np.arange(1, n_samples + 1, dtype=np.uint32)
    else:
        ids = np.sort(rng.choice(np.arange(1, 2 ** 32 - 1, 65537, dtype=np.uint64), size=n_samples, replace=False)).astype(np.uint32)
    sizes = list(rng.integers(s_lo, s_hi + 1, size=n_colors)) + [min(b, n_samples) for b in big]
    sizes = [int(min(s, n_samples)) for s in sizes]
    offsets = np.zeros(len(sizes) + 1, dtype=np.uint64)
    members = []
    for c, s in enumerate(sizes):
        m = rng.choice(n_samples, size=s, replace=False)
        members.append(ids[m])
        offsets[c + 1] = offsets[c] + np.uint64(s)
    members = np.concatenate(members).astype(np.uint32) if members else np.empty(0, np.uint32)
    weights = rng.integers(1, w_hi + 1, size=len(sizes), dtype=np.int64).astype(np.uint32)
    if zero_every:
        weights[::zero_every] = 0
    totals = rng.integers(1, 10 ** 6, size=n_samples, dtype=np.int64).astype(np.uint32)
    return dict(offsets=offsets, members=members, weights=weights, total_ids=ids.copy(), total_vals=totals)


def oracle_records(csr):
    """Oracle result as a structured array sorted by (source_1, source_2) — the oracle sorts."""
    from oracle import pairwise_oracle as po
    from kspider_b200.api import RECORD_DTYPE
    r = po.pairwise_csr(csr["offsets"], csr["members"], csr["weights"], csr["total_ids"], csr["total_vals"])
    out = np.zeros(r.src1.size, dtype=RECORD_DTYPE)
    out["source_1"], out["source_2"], out["shared_kmers"] = r.src1, r.src2, r.shared
    out["min_containment"], out["avg_containment"], out["max_containment"] = r.cmin, r.cavg, r.cmax
    return out


def sort_records(rec: np.ndarray) -> np.ndarray:
    order = np.lexsort((rec["source_2"], rec["source_1"]))
    return rec[order]


def assert_records_equal(got: np.ndarray, want: np.ndarray) -> None:
    """Bit-exact: integers equal, float32 containments compared by bit pattern."""
    got, want = sort_records(got), sort_records(want)
    assert got.size == want.size, f"row count {got.size} != {want.size}"
    for f in ("source_1", "source_2", "shared_kmers"):
        bad = np.nonzero(got[f] != want[f])[0]
        assert bad.size == 0, f"{f} differs at row {bad[:5]}: {got[f][bad[:5]]} vs {want[f][bad[:5]]}"
    for f in ("min_containment", "avg_containment", "max_containment"):
        g, w = got[f].view(np.uint32), want[f].view(np.uint32)
        bad = np.nonzero(g != w)[0]
        assert bad.size == 0, f"{f} bits differ at row {bad[:5]}: {got[f][bad[:5]]} vs {want[f][bad[:5]]}"


def canonical_partition(labels: np.ndarray):
    groups = {}
    for i, l in enumerate(labels.tolist()):
        groups.setdefault(l, []).append(i)
    return sorted(tuple(g) for g in groups.values())


def stage_golden_prefix(name: str, td: str, ksize: int = 21, with_index_files: bool = False) -> str:
    """Index prefix under `td` holding what `kSpider cluster` / `pairwise --estimate-ani` read for golden case
    `name`: the reference's pairwise TSV and seqToKmersNo.tsv (tests/golden, made by the unmodified
    src/pairwise.cpp), a namesMap with names s<id> in id order and .extra (first line = kSize).
    with_index_files: also the three phmap index files (so that the product can run on the same prefix)."""
    prefix = os.path.join(str(td), name)
    src = os.path.join(GOLD, name + ".pairwise.tsv")
    if os.path.exists(src + ".gz"):
        with gzip.open(src + ".gz", "rb") as f, open(prefix + "_kSpider_pairwise.tsv", "wb") as g:
            shutil.copyfileobj(f, g)
    else:
        shutil.copy(src, prefix + "_kSpider_pairwise.tsv")
    shutil.copy(os.path.join(GOLD, name + ".seq.tsv"), prefix + "_kSpider_seqToKmersNo.tsv")
    ids = np.sort(np.load(os.path.join(GOLD, name + ".raw.npz"))["total_ids"])
    assert ids[0] == 1 and ids[-1] == ids.size, "the reference's cluster command needs ids 1..N"
    with open(prefix + ".namesMap", "w") as f:
        f.write(f"{ids.size}\n")
        for i in ids:
            f.write(f"{i} s{i}\n")
    with open(prefix + ".extra", "w") as f:
        f.write(f"{ksize}\n1\n1\n\n")
    if with_index_files:
        for suf in ("_color_to_sources.bin", "_color_count.bin", "_groupID_to_kmerCount.bin"):
            shutil.copy(os.path.join(GOLD, name + suf), prefix + suf)
    return prefix


def golden_clusters(name: str) -> dict:
    """{(dist_type, cutoff): partition as sorted lists of sample ids} written by the UNMODIFIED
    ks_clustering.py (tests/golden/make_golden_py.py)."""
    with open(os.path.join(GOLD, name + ".clusters.json")) as f:
        raw = json.load(f)
    return {(k.split("|")[0], float(k.split("|")[1])): [list(c) for c in v] for k, v in raw.items()}


def partition_of_names(clusters) -> list:
    """clusters of names s<id> -> sorted lists of ids (the goldens' canonical form)"""
    return sorted(sorted(int(n[1:]) for n in c) for c in clusters)
