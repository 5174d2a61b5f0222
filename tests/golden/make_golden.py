"""Generates the golden fixtures of tests/golden/ by running the UNMODIFIED reference
(oracle/_ref/pairwise_ref = /root/reference/src/pairwise.cpp, built by oracle/Makefile) on seeded
inputs written with the reference's own vendored phmap (oracle/_ref/index_writer_ref).

Run in the build container (needs oracle/_ref/, i.e. /root/reference at build time):
    python tests/golden/make_golden.py
Each case <name> leaves:
    <name>.raw.npz        the input (CSR + totals + colour ids/counts as written to the files)
    <name>.pairwise.tsv[.gz]  the reference's TSV, rows sorted by (source_1, source_2)
    <name>.seq.tsv        the reference's _kSpider_seqToKmersNo.tsv, verbatim
    <name>_*.bin          the three index files as phmap itself dumped them (small cases only)
"""
import gzip
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import reftools  # noqa: E402
from util import random_csr  # noqa: E402


def micro_kat():
    """SURVEY Appendix C.2 — hand-checkable."""
    totals = {1: 100, 2: 50, 3: 75, 4: 3}
    colors = [(1, [1]), (2, [2]), (3, [3]), (8, [4]), (5, [1, 2]), (6, [1, 2, 3]), (7, [2, 3])]
    counts = {1: 60, 2: 10, 3: 40, 4: 0, 5: 10, 6: 30, 7: 5, 8: 3}
    return totals, counts, colors


def from_csr(csr, extra_singletons=True, wide_ids=False):
    """dict form for reftools.write_raw: colour ids N+1.., optional singleton colours, optional
    ids/counts above 2^32 to exercise the reference's u32 narrowing (src/pairwise.cpp:109,118-120)."""
    totals = {int(k): int(v) for k, v in zip(csr["total_ids"], csr["total_vals"])}
    colors, counts = [], {}
    off = csr["offsets"].astype(np.int64)
    base = int(csr["total_ids"].max()) + 1
    for c in range(off.size - 1):
        cid = base + c
        w = int(csr["weights"][c])
        if wide_ids and c % 3 == 0:
            w += (c % 5 + 1) << 32  # high bits are dropped by the reference
        colors.append((cid, csr["members"][off[c]:off[c + 1]].tolist()))
        counts[cid] = w
    if extra_singletons:
        for sid in csr["total_ids"][: 50]:
            colors.append((int(sid), [int(sid)]))
            counts[int(sid)] = 1
    return totals, counts, colors


def clades_csr(seed=15, n_clades=8, clade=15, n_core=220, n_family=40, n_noise=150):
    """Clade-structured colour table with ids 1..N and totals = the sample's own k-mer count (sum of the
    weights of every colour containing it + a private remainder), so every containment lies in [0, 1] and
    spreads over it: the case behind the cluster / ANI goldens (make_golden_py.py)."""
    rng = np.random.default_rng(seed)
    n = n_clades * clade
    cols, w = [], []
    for c in range(n_core):
        k = c % n_clades
        keep = rng.random(clade) < rng.uniform(0.55, 1.0)
        m = np.flatnonzero(keep) + k * clade
        if m.size >= 1:
            cols.append(m); w.append(int(rng.integers(1, 400)))
    for c in range(n_family):
        k = int(rng.integers(0, n_clades - 1))
        m = np.flatnonzero(rng.random(2 * clade) < 0.5) + k * clade
        if m.size >= 1:
            cols.append(m); w.append(int(rng.integers(1, 60)))
    for c in range(n_noise):
        cols.append(rng.choice(n, size=int(rng.integers(1, 5)), replace=False)); w.append(int(rng.integers(0, 8)))
    ids = np.arange(1, n + 1, dtype=np.uint32)
    tot = np.zeros(n, dtype=np.int64)
    for m, ww in zip(cols, w):
        tot[m] += ww
    private = rng.integers(0, 300, size=n)
    private[::9] = 0          # some samples are fully covered by shared colours: containment exactly 1 happens
    tot += private
    tot[tot == 0] = 1
    sizes = np.array([m.size for m in cols], dtype=np.uint64)
    members = np.concatenate([ids[rng.permutation(m)] for m in cols]).astype(np.uint32)
    return dict(offsets=np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64), members=members,
                weights=np.array(w, dtype=np.uint32), total_ids=ids, total_vals=tot.astype(np.uint32))


CASES = {
    "micro_kat": lambda: micro_kat(),
    "small_dense": lambda: from_csr(random_csr(11, 40, 300, 1, 9)),
    "sparse_ids_zero_w": lambda: from_csr(random_csr(12, 60, 400, 1, 12, id_mode="sparse", zero_every=7)),
    "mid_colors": lambda: from_csr(random_csr(13, 400, 200, 2, 40, big=(65, 66, 130, 257, 399))),
    "wide_counts": lambda: from_csr(random_csr(14, 50, 200, 2, 8), wide_ids=True),
    "clades": lambda: from_csr(clades_csr(), extra_singletons=False),
}
KEEP_BINS = set(CASES)


def main():
    assert reftools.have_ref(), "oracle/_ref is not built: make -C oracle ref"
    for name, make in CASES.items():
        totals, counts, colors = make()
        with tempfile.TemporaryDirectory() as td:
            raw = os.path.join(td, "in.raw")
            prefix = os.path.join(td, name)
            reftools.write_raw(raw, totals, counts, colors)
            reftools.run_index_writer_ref(raw, prefix)
            reftools.run_pairwise_ref(prefix, threads=2)
            header, rows = reftools.read_pairwise_tsv_sorted(prefix + "_kSpider_pairwise.tsv")
            big = len(rows) > 5000  # large goldens are stored gzip-compressed
            out = os.path.join(HERE, name + ".pairwise.tsv" + (".gz" if big else ""))
            with (gzip.open(out, "wt") if big else open(out, "w")) as f:
                f.write(header)
                f.write("\n".join(rows) + ("\n" if rows else ""))
            shutil.copy(prefix + "_kSpider_seqToKmersNo.tsv", os.path.join(HERE, name + ".seq.tsv"))
            if name in KEEP_BINS:
                for suf in ("_color_to_sources.bin", "_color_count.bin", "_groupID_to_kmerCount.bin"):
                    shutil.copy(prefix + suf, os.path.join(HERE, name + suf))
        sizes = np.array([len(m) for _, m in colors], dtype=np.uint64)
        np.savez_compressed(
            os.path.join(HERE, name + ".raw.npz"),
            total_ids=np.array(list(totals.keys()), dtype=np.uint32),
            total_vals=np.array(list(totals.values()), dtype=np.uint32),
            count_ids=np.array(list(counts.keys()), dtype=np.uint64),
            count_vals=np.array(list(counts.values()), dtype=np.uint64),
            color_ids=np.array([c for c, _ in colors], dtype=np.uint64),
            offsets=np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64),
            members=np.concatenate([np.asarray(m, dtype=np.uint32) for _, m in colors]))
        print(name, len(rows), "rows")


if __name__ == "__main__":
    main()
