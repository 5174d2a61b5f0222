"""Drop-in for `kSpider pairwise --estimate-ani` (pykSpider/kSpider2/ks_pairwise.py:29-82) on the GPU.

The reference re-reads its pairwise TSV line by line in Python and calls sourmash for every row ("does not
allow multithreading", :39-40).  Here the min / max containment columns of the TSV ON DISK are read -- from the
binary side-car our pairwise wrote beside it when it still belongs to that TSV, else by parsing the text
(multithreaded) -- K8 turns them into the avg_ani column on the device (csrc/kernels.cu `k8_ani_cols`), and the
column is written in the row order of the TSV, as the reference does: line k of the column file belongs to
line k of the TSV, whichever tool wrote it.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .api import write_ani_column
from .capi import check, lib
from .clustering import _Table


def estimate_ani(index_prefix: str, sourmash_scale: int, user_threads: int = 1, device: int = -1) -> str:
    """Mirrors `kSpider pairwise -i PREFIX --estimate-ani -s SCALE` and its error behaviour: the
    pairwise TSV must exist (:34-35), a scale must be given (:42-43; it does not enter the value),
    kSize is the first line of `<prefix>.extra` (:47-48), every sample id of the TSV must have a row in
    `<prefix>_kSpider_seqToKmersNo.tsv` (:55-60,78-79: KeyError otherwise).  A row whose printed containment is
    > 1, inf or nan raises ValueError, as sourmash's ANIResult does.  Returns the output path."""
    pairwise_file = index_prefix + "_kSpider_pairwise.tsv"
    ani_col = index_prefix + "_kSpider_pairwise.ani_col.tsv"
    seqToKmers_tsv = f"{index_prefix}_kSpider_seqToKmersNo.tsv"
    if not os.path.exists(pairwise_file):
        raise FileNotFoundError("Please, run the same command without --estimate-ani first, then run this command.")
    if not sourmash_scale:
        raise ValueError("estimating ANI requires to provide --scale value")
    with open(f"{index_prefix}.extra") as extra:
        ksize = int(next(extra))
    known = set()
    with open(seqToKmers_tsv) as KMER_COUNT:
        next(KMER_COUNT)
        for line in KMER_COUNT:
            seq_ID, _no_of_kmers = tuple(line.strip().split('\t')[1:])
            known.add(int(seq_ID))
    table = _Table.read(index_prefix, 3, 5, threads=max(1, int(user_threads)))
    try:
        n = C.c_uint64()
        s1, s2 = C.c_void_p(), C.c_void_p()
        check(lib().ksb_table_get(table._h, C.byref(n), C.byref(s1), C.byref(s2), None, None))
        if n.value:
            ids = np.unique(np.concatenate([np.frombuffer((C.c_uint32 * n.value).from_address(p.value), dtype=np.uint32)
                                            for p in (s1, s2)]))
            missing = [int(i) for i in ids.tolist() if i not in known]
            if missing:
                raise KeyError(missing[0])  # id_to_kmer_count[id] in the reference
        ani = np.empty(n.value, dtype=np.float64)
        bad = C.c_uint64()
        check(lib().ksb_ani_table(table._h, ksize, int(device), ani.ctypes.data, C.byref(bad)))
    finally:
        table.close()
    if bad.value:
        raise ValueError(f"Error: {bad.value} row(s) hold a containment outside [0, 1]; distance value and p_threshold "
                         "value must both be between 0 and 1.")
    write_ani_column(ani, ani_col, user_threads)
    return ani_col
