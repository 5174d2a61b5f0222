"""Colour construction for the pairwise path: host mirror of the reference's
`kSpider::sourmash_sigs_indexing(sigs_dir, kSize)` (include/kSpider.hpp:18,
src/sourmash_indexing.cpp:50-327) on top of the GPU builder `ksb_colors_build` (csrc/colors.cu).

What is reproduced: the END STATE the pairwise pass reads — per hash the exact set of samples that
contain it, per set the number of hashes (colorsCount), per sample mins.size() — and the file set
(`<prefix>_color_to_sources.bin`, `_color_count.bin`, `_groupID_to_kmerCount.bin` as phmap dumps,
`.namesMap`, `.extra`).  Sample ids are 1.. in glob (sorted) order of the signature files like the
reference's groupID counter (:74, :107-116); the singleton colour of sample g keeps id g (:108-112);
multi-member colours get fresh ids above N (the reference's ids depend on its processing order and
id recycling, :214-222, and influence no output of the pairwise path).
"""
from __future__ import annotations

import ctypes as C
import glob
import gzip
import json
import os
from dataclasses import dataclass
from typing import Optional, List, Sequence

import numpy as np

from .api import Index, _view, write_index_files
from .capi import check, lib


@dataclass
class Colors:
    offsets: np.ndarray   # uint64 [C+1]
    members: np.ndarray   # uint32 sample POSITIONS (0..N-1), ascending inside a colour
    counts: np.ndarray    # uint64 [C]
    totals: np.ndarray    # uint64 [N]
    distinct_hashes: int
    device_ms: float


def build_colors(sketches: Sequence[np.ndarray], device: int = -1) -> Colors:
    """ksb_colors_build on the concatenated hash lists (one uint64 array per sample)."""
    n = len(sketches)
    offsets = np.zeros(n + 1, dtype=np.uint64)
    if n:
        offsets[1:] = np.cumsum([int(np.asarray(s).size) for s in sketches], dtype=np.uint64)
    hashes = (np.concatenate([np.ascontiguousarray(s, dtype=np.uint64).ravel() for s in sketches])
              if n and int(offsets[-1]) else np.zeros(0, dtype=np.uint64))
    h = C.c_void_p()
    check(lib().ksb_colors_build(n, offsets.ctypes.data, hashes.ctypes.data if hashes.size else None, int(device), C.byref(h)))
    try:
        L = lib()
        nc = int(L.ksb_colors_count(h))
        off = _view(L.ksb_colors_offsets(h), nc + 1, np.uint64).copy()
        mem = _view(L.ksb_colors_members(h), int(off[-1]), np.uint32).copy() if nc and int(off[-1]) else np.zeros(0, np.uint32)
        cnt = _view(L.ksb_colors_counts(h), nc, np.uint64).copy() if nc else np.zeros(0, np.uint64)
        tot = _view(L.ksb_colors_totals(h), n, np.uint64).copy() if n else np.zeros(0, np.uint64)
        return Colors(off, mem, cnt, tot, int(L.ksb_colors_distinct_hashes(h)), float(L.ksb_colors_device_ms(h)))
    finally:
        lib().ksb_colors_free(h)


def index_from_colors(colors: Colors, sample_ids: np.ndarray, host_threads: int = 1) -> Index:
    """The parsed-index state the pairwise pass starts from (counts and totals narrowed to u32 exactly
    as src/pairwise.cpp:118-120,166-170 narrows what it reads from the files)."""
    ids = np.ascontiguousarray(sample_ids, dtype=np.uint32)
    return Index.from_csr(colors.offsets, ids[colors.members], (colors.counts & np.uint64(0xFFFFFFFF)).astype(np.uint32),
                          ids, (colors.totals & np.uint64(0xFFFFFFFF)).astype(np.uint32), host_threads=host_threads)


def write_index(prefix: str, colors: Colors, sample_ids: np.ndarray, names: Sequence[str], ksize: int,
                has_total: Optional[np.ndarray] = None) -> None:
    """The reference's index file set for `prefix` (src/sourmash_indexing.cpp:281-327).  has_total[s] False: sample s
    gets no entry in _groupID_to_kmerCount.bin (the reference only records a k-mer count for a signature it parsed
    with the requested ksize, :205,282-285)."""
    ids = np.ascontiguousarray(sample_ids, dtype=np.uint32)
    n = ids.size
    sizes = np.diff(colors.offsets).astype(np.int64)
    color_ids = np.zeros(sizes.size, dtype=np.uint64)
    single = sizes == 1
    first = colors.members[colors.offsets[:-1][single].astype(np.int64)] if single.any() else np.zeros(0, np.uint32)
    color_ids[single] = ids[first]                       # the sample's own colour keeps the sample's id
    base = int(ids.max()) + 1 if n else 1
    color_ids[~single] = base + np.arange(int((~single).sum()), dtype=np.uint64)
    # colorsCount also holds every sample's own colour, with 0 when no hash is private to it (:112)
    have = set(int(v) for v in color_ids[single])
    extra = np.array([int(v) for v in ids if int(v) not in have], dtype=np.uint64)
    count_ids = np.concatenate([color_ids, extra])
    count_vals = np.concatenate([colors.counts.astype(np.uint64), np.zeros(extra.size, dtype=np.uint64)])
    keep = np.ones(n, dtype=bool) if has_total is None else np.asarray(has_total, dtype=bool)
    write_index_files(prefix, color_ids, colors.offsets, ids[colors.members], count_ids, count_vals,
                      ids[keep], (colors.totals & np.uint64(0xFFFFFFFF)).astype(np.uint32)[keep])
    with open(prefix + ".namesMap", "w") as f:           # :313-319
        f.write(f"{n}\n")
        for i, nm in zip(ids, names):
            f.write(f"{int(i)} {nm}\n")
    with open(prefix + ".extra", "w") as f:              # :322-327 (only the kSize line is ever read back)
        f.write(f"{int(ksize)}\n3\n1\nmumur_hasher\n")


def read_signature_mins(path: str, ksize: int):
    """mins of the first signature with the requested ksize (json[0]["signatures"], :165-176, :274 break)."""
    opener = gzip.open if open(path, "rb").read(2) == b"\x1f\x8b" else open
    with opener(path, "rt") as f:
        doc = json.load(f)
    for sig in doc[0]["signatures"]:
        if sig.get("ksize") == ksize:
            return np.array(sig["mins"], dtype=np.uint64)
    return None


def sourmash_sigs_indexing(sigs_dir: str, kSize: int, device: int = -1, out_dir: str = ".") -> str:
    """Mirror of kSpider_internal.sourmash_sigs_indexing(sigs_dir, kSize): writes the index file set
    with prefix basename(sigs_dir) (in the working directory like the reference, or `out_dir`)."""
    sigs_dir = sigs_dir.rstrip("/")
    prefix = os.path.join(out_dir, os.path.basename(sigs_dir))
    # First loop of the reference (:85-117): EVERY file whose last extension is "sig" or "gz" names a group (the file name
    # without that extension) and gets the next id, in glob(3) order = byte order of the paths; a name seen twice keeps
    # its id.  Second loop (:140-275): only files whose last extension is "sig" are parsed (plain or gzip JSON, zstr).
    names: List[str] = []
    index_of = {}
    sketches: List[np.ndarray] = []
    parsed: List[bool] = []
    for path in sorted(glob.glob(sigs_dir + "/*")):
        root, ext = os.path.splitext(path)
        if ext not in (".sig", ".gz"):
            continue
        name = os.path.basename(root)
        if name not in index_of:
            index_of[name] = len(names)
            names.append(name)
            sketches.append(np.zeros(0, dtype=np.uint64))
            parsed.append(False)
        if ext != ".sig":
            continue
        mins = read_signature_mins(path, int(kSize))
        if mins is not None:                                 # no signature of this ksize: no k-mer count for the sample (:176-178)
            sketches[index_of[name]] = mins
            parsed[index_of[name]] = True
    ids = np.arange(1, len(names) + 1, dtype=np.uint32)      # groupID counter starts at 1 (:74)
    colors = build_colors(sketches, device=device)
    write_index(prefix, colors, ids, names, int(kSize), has_total=np.array(parsed, dtype=bool))
    return prefix
