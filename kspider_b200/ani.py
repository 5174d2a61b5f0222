"""`kSpider pairwise --estimate-ani` on the GPU (pykSpider/kSpider2/ks_pairwise.py:29-82).

The reference re-reads its own pairwise TSV line by line in Python and calls sourmash for every row
("does not allow multithreading", :42).  Here the pairwise result is recomputed on the device (a few
milliseconds), K8 turns the records into the avg_ani column (csrc/kernels.cu `k8_ani`), and the
column is written in the row order of the EXISTING `<prefix>_kSpider_pairwise.tsv` — whichever tool
wrote it (the reference's rows come in hash-map order, ours sorted) — so line k of the column file
belongs to line k of the TSV, as in the reference.
"""
from __future__ import annotations

import os

import numpy as np

from .api import Index, write_ani_column


def estimate_ani(index_prefix: str, sourmash_scale: int, user_threads: int = 1, **opts) -> str:
    """Mirrors `kSpider pairwise -i PREFIX --estimate-ani -s SCALE` and its error behaviour: the
    pairwise TSV must exist (:35-36), a scale must be given (:44-45; it does not enter the value),
    kSize is the first line of `<prefix>.extra` (:47-48).  A row whose printed containment is > 1,
    inf or nan raises ValueError, as sourmash's ANIResult does.  Returns the output path."""
    pairwise_file = index_prefix + "_kSpider_pairwise.tsv"
    ani_col = index_prefix + "_kSpider_pairwise.ani_col.tsv"
    if not os.path.exists(pairwise_file):
        raise FileNotFoundError("Please, run the same command without --estimate-ani first, then run this command.")
    if not sourmash_scale:
        raise ValueError("estimating ANI requires to provide --scale value")
    with open(f"{index_prefix}.extra") as extra:
        ksize = int(next(extra))
    index = Index.load(index_prefix, user_threads)
    try:
        result = index.pairwise(want_records=True, **opts)
        try:
            ani, n_bad = result.ani_column(ksize)
            rec = result.records()
        finally:
            result.close()
    finally:
        index.close()
    if n_bad:
        raise ValueError(f"Error: {n_bad} row(s) hold a containment outside [0, 1]; distance value and p_threshold "
                         "value must both be between 0 and 1.")
    # row order of the TSV on disk: records are sorted by (source_1, source_2)
    import pandas as pd
    tsv = pd.read_csv(pairwise_file, sep="\t", usecols=[0, 1], dtype=np.uint64)
    want = (tsv.iloc[:, 0].to_numpy() << np.uint64(32)) | tsv.iloc[:, 1].to_numpy()
    have = (rec["source_1"].astype(np.uint64) << np.uint64(32)) | rec["source_2"].astype(np.uint64)
    pos = np.searchsorted(have, want)
    if want.size != have.size or (pos >= have.size).any() or not np.array_equal(have[np.minimum(pos, have.size - 1)], want):
        raise RuntimeError(f"{pairwise_file} does not hold the pairs of this index (stale file?)")
    write_ani_column(ani[pos], ani_col, user_threads)
    return ani_col
