"""Mirror of the reference's SWIG module for the hot path.

Reference: src/swig_interfaces/kSpider_internal.i:11 declares
    void pairwise(string index_prefix, int user_threads);
built with `-keyword` (setup.py:140-146), so both positional and keyword calls work:
    kSpider_internal.pairwise(index_prefix, user_threads)          (ks_pairwise.py:27)
    kSpider_internal.pairwise(index_prefix=..., user_threads=...)  (test/kspider_run.py:4)
Install this file as `_kSpider_internal` on sys.path and the reference's
`kSpider pairwise` command runs on the GPU unchanged.
"""
from __future__ import annotations

from .capi import KsbError, lib


def pairwise(index_prefix: str, user_threads: int = 1) -> None:
    """Reads the three index files of `index_prefix`, writes `<prefix>_kSpider_seqToKmersNo.tsv`
    and `<prefix>_kSpider_pairwise.tsv`.  Raises RuntimeError on failure (the reference would
    abort on an assert or propagate a C++ exception through SWIG)."""
    L = lib()
    rc = L.ksb_dropin_pairwise(str(index_prefix).encode(), int(user_threads))
    if rc != 0:
        raise KsbError(rc, (L.ksb_last_error() or b"").decode(errors="replace"))


def sourmash_sigs_indexing(sigs_dir: str, kSize: int) -> None:
    """kSpider_internal.sourmash_sigs_indexing(sigs_dir, kSize) (kSpider_internal.i:18,
    src/sourmash_indexing.cpp:50-327): index file set with prefix basename(sigs_dir) in the working
    directory; the colour table is built on the GPU (csrc/colors.cu)."""
    from .indexing import sourmash_sigs_indexing as _impl
    _impl(str(sigs_dir), int(kSize))
