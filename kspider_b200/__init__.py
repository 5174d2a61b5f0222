"""kspider_b200 — B200-native (sm_100a CUDA) replacement of kSpider's pairwise hot path.

Product = kspider_b200/libkspider_b200.so (C-ABI in include/kspider_b200.h).  This package is the
Python host side mirroring the reference's own interfaces for the path:
  kSpider_internal.pairwise(index_prefix, user_threads)   <- src/swig_interfaces/kSpider_internal.i:11
  clustering.Clusters / clustering.cluster(...)            <- pykSpider/kSpider2/ks_clustering.py
"""
from . import capi  # noqa: F401
from .api import (Device, Index, Result, bench_red, cluster_merge, format_float, threshold_float,  # noqa: F401
                  write_index_files)

__all__ = ["Index", "Device", "Result", "capi", "threshold_float", "format_float", "cluster_merge",
           "write_index_files", "bench_red"]
