"""Generates the kProcessor `colored_kDataFrame` fixtures with the REFERENCE's own code:
oracle/_ref/ckf_ref = /root/reference/lib/kProcessor/src/colorTable.cpp (intVectorsTable save / load, unmodified) +
the vendored phmap (kDataFramePHMAP's map type), built by oracle/Makefile.

    python tests/golden/make_golden_ckf.py        (build container only: needs oracle/_ref)
leaves, for the case `ckf_small`:
    ckf_small.phmap, ckf_smallcolors.intvectors, ckf_smallcolorTableConfig.json, ckf_small.namesMap, ckf_small.extra
                             the file set colored_kDataFrame::save(prefix) writes (real bytes / real text)
    ckf_small.dump.txt       what the reference's loaders read back: per colour id, #k-mers and members; per sample total
    ckf_small.pairwise.tsv   the UNMODIFIED src/pairwise.cpp on the three .bin files that state corresponds to
                             (written by oracle/_ref/index_writer_ref from the dump), rows sorted
"""
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import reftools  # noqa: E402

CKF_REF = os.path.join(reftools.REF_DIR, "ckf_ref")


def write_ckf_raw(path, ksize, kmers, colors, names):
    with open(path, "wb") as f:
        f.write(struct.pack("<QQ", 0x3130464B43425343, ksize))
        f.write(struct.pack("<Q", len(kmers)))
        for h, c in kmers:
            f.write(struct.pack("<QQ", h, c))
        f.write(struct.pack("<Q", len(colors)))
        for cid, members in colors:
            f.write(struct.pack("<QI", cid, len(members)))
            f.write(np.asarray(members, dtype="<u4").tobytes())
        f.write(struct.pack("<Q", len(names)))
        for sid, name in names:
            b = name.encode()
            f.write(struct.pack("<II", sid, len(b)) + b)


def parse_dump(path):
    colors, totals = [], {}
    for line in open(path):
        t = line.split()
        if t[0] == "C":
            colors.append((int(t[1]), int(t[2]), [int(x) for x in t[4:]]))
            assert len(t) - 4 == int(t[3])
        elif t[0] == "T":
            totals[int(t[1])] = int(t[2])
    return colors, totals


def make_case(seed=41, n_samples=60, n_colors=320):
    rng = np.random.default_rng(seed)
    colors, kmers = [], []
    used = set()
    for c in range(n_colors):
        cid = int(rng.integers(1, 1 << 40)) if c % 7 == 0 else 1000 + c   # some ids far above 2^32
        s = int(rng.integers(1, 13)) if c % 29 else int(rng.integers(30, n_samples))
        members = (rng.choice(n_samples, size=s, replace=False) + 1).tolist()
        colors.append((cid, members))
        n_k = 0 if c % 11 == 5 else int(rng.integers(1, 40))   # some colours have no k-mer left: count 0
        for _ in range(n_k):
            h = int(rng.integers(0, 1 << 62))
            while h in used:
                h = int(rng.integers(0, 1 << 62))
            used.add(h)
            kmers.append((h, cid))
    # src/pairwise.cpp:109 narrows colour ids to u32 when it reads the .bin files: keep them distinct after narrowing
    assert len({cid & 0xFFFFFFFF for cid, _ in colors}) == len(colors)
    names = [(i, f"genome_{i}") for i in range(1, n_samples + 3)]   # two names no k-mer reached
    return kmers, colors, names


def main():
    assert os.access(CKF_REF, os.X_OK), "oracle/_ref/ckf_ref is not built: make -C oracle ref"
    name = "ckf_small"
    kmers, colors, names = make_case()
    prefix = os.path.join(HERE, name)
    with tempfile.TemporaryDirectory() as td:
        raw = os.path.join(td, "in.raw")
        write_ckf_raw(raw, 21, kmers, colors, names)
        subprocess.run([CKF_REF, "write", raw, prefix], check=True)
        subprocess.run([CKF_REF, "dump", prefix, prefix + ".dump.txt"], check=True)
        dcolors, totals = parse_dump(prefix + ".dump.txt")
        # the .bin state the same index corresponds to -> the unmodified pairwise pass
        for sid, _ in names:
            totals.setdefault(sid, 0)
        bprefix = os.path.join(td, name)
        reftools.write_raw(os.path.join(td, "bins.raw"), totals, {cid: cnt for cid, cnt, _ in dcolors},
                           [(cid, m) for cid, _, m in dcolors])
        reftools.run_index_writer_ref(os.path.join(td, "bins.raw"), bprefix)
        reftools.run_pairwise_ref(bprefix, threads=2)
        header, rows = reftools.read_pairwise_tsv_sorted(bprefix + "_kSpider_pairwise.tsv")
        with open(prefix + ".pairwise.tsv", "w") as f:
            f.write(header + "\n".join(rows) + "\n")
    print(name, len(kmers), "k-mers", len(dcolors), "colours", len(rows), "rows")


if __name__ == "__main__":
    main()
