"""ctypes binding of libkspider_b200.so (include/kspider_b200.h).

The library is the product; this module only declares its prototypes.  There is no Python or CPU
implementation behind it: if the shared library is missing, importing anything that computes
raises ImportError, and on a box without an sm_100 device every compute call returns KSB_ERR_CUDA.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkspider_b200.so")

KSB_OK = 0
KSB_ERR_IO, KSB_ERR_FORMAT, KSB_ERR_NOMEM, KSB_ERR_CUDA, KSB_ERR_ARG, KSB_ERR_BUDGET = -1, -2, -3, -4, -5, -6


class KsbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libkspider_b200 error {code}: {msg}")
        self.code = code


class IndexInfo(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "n_samples", "n_colors_file", "n_colors", "n_members", "n_updates", "n_small_colors",
        "n_mid_colors", "n_gram_colors", "gram_min_size", "gram_planes", "n_gram_updates", "n_zero_colors", "max_color_size", "payload_bytes", "triangle_cells",
        "n_small_members", "n_small_updates", "n_gram_members", "n_gram_windows", "n_gram_macs")] + \
        [("pinned", C.c_uint32), ("gram_form", C.c_uint32), ("n_gram_cells", C.c_uint64)]


class Opts(C.Structure):
    _fields_ = [("device", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32), ("want_records", C.c_int32),
                ("dense_budget_bytes", C.c_uint64), ("spill_mode", C.c_uint64), ("reserved", C.c_uint64 * 3)]


class Record(C.Structure):
    _fields_ = [("source_1", C.c_uint32), ("source_2", C.c_uint32), ("shared_kmers", C.c_uint64),
                ("min_containment", C.c_float), ("avg_containment", C.c_float), ("max_containment", C.c_float),
                ("pad", C.c_uint32)]


class Timings(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("h2d_ms", "zero_ms", "small_ms", "mid_ms", "zero_touch_ms",
                                         "accumulate_ms", "post_ms", "d2h_ms", "cluster_ms", "gram_ms", "spill_emit_ms", "spill_sort_ms",
                                         "host_setup_ms", "host_accumulate_ms", "host_post_ms", "host_free_ms")]


class ResultInfo(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("n_rows", "n_updates", "row_begin", "row_end", "owned_cells",
                                          "gpu_launches")] + [("t", Timings)]


_PROTOS = {
    # name: (restype, argtypes)
    "ksb_last_error": (C.c_char_p, []),
    "ksb_abi_version": (C.c_int, []),
    "ksb_cuda_available": (C.c_int, []),
    "ksb_index_load": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]),
    "ksb_index_load_colored": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "ksb_index_from_csr": (C.c_int, [C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                     C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]),
    "ksb_index_get_info": (C.c_int, [C.c_void_p, C.POINTER(IndexInfo)]),
    "ksb_index_sample_ids": (C.c_void_p, [C.c_void_p]),
    "ksb_index_totals": (C.c_void_p, [C.c_void_p]),
    "ksb_index_payload": (C.c_void_p, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "ksb_index_descriptor": (C.c_void_p, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "ksb_index_row_cuts": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "ksb_index_bands": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ksb_blob_row_cuts": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]),
    "ksb_index_shard": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "ksb_index_from_blob": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p)]),
    "ksb_write_seq_to_kmers": (C.c_int, [C.c_void_p, C.c_char_p]),
    "ksb_index_free": (None, [C.c_void_p]),
    "ksb_upload": (C.c_int, [C.c_void_p, C.POINTER(Opts), C.POINTER(C.c_void_p)]),
    "ksb_adopt_device_payload": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.POINTER(Opts),
                                           C.POINTER(C.c_void_p)]),
    "ksb_accumulate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "ksb_device_triangle": (C.c_void_p, [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                         C.POINTER(C.c_uint64)]),
    "ksb_device_read_triangle": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "ksb_device_cell_bits": (C.c_int, [C.c_void_p]),
    "ksb_device_is_spill": (C.c_int, [C.c_void_p]),
    "ksb_device_set_overlap": (C.c_int, [C.c_void_p, C.c_int]),
    "ksb_device_run_spill": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "ksb_device_n_passes": (C.c_int, [C.c_void_p]),
    "ksb_device_set_pass": (C.c_int, [C.c_void_p, C.c_int]),
    "ksb_device_get_info": (C.c_int, [C.c_void_p, C.POINTER(ResultInfo)]),
    "ksb_postpass": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]),
    "ksb_device_cluster": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "ksb_device_cluster_enqueue": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_int]),
    "ksb_device_cluster_collect": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "ksb_device_set_graph": (C.c_int, [C.c_void_p, C.c_int]),
    "ksb_device_free": (None, [C.c_void_p]),
    "ksb_pairwise": (C.c_int, [C.c_void_p, C.POINTER(Opts), C.POINTER(C.c_void_p)]),
    "ksb_pairwise_cluster": (C.c_int, [C.c_void_p, C.POINTER(Opts), C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64),
                                       C.POINTER(C.c_uint64), C.POINTER(ResultInfo)]),
    "ksb_result_get_info": (C.c_int, [C.c_void_p, C.POINTER(ResultInfo)]),
    "ksb_result_records": (C.c_void_p, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "ksb_write_tsv": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int]),
    "ksb_append_tsv": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int]),
    "ksb_write_tsv_records": (C.c_int, [C.c_void_p, C.c_uint64, C.c_char_p, C.c_int, C.c_int]),
    "ksb_cluster": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.POINTER(C.c_uint64)]),
    "ksb_cluster_pct": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.POINTER(C.c_uint64)]),
    "ksb_cluster_pct_device": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.POINTER(C.c_uint64)]),
    "ksb_threshold_float_pct": (C.c_float, [C.c_double]),
    "ksb_cluster_merge": (C.c_int, [C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ksb_cluster_merge_device": (C.c_int, [C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ksb_threshold_float": (C.c_float, [C.c_double]),
    "ksb_result_free": (None, [C.c_void_p]),
    "ksb_dropin_pairwise": (C.c_int, [C.c_char_p, C.c_int]),
    "ksb_colors_build": (C.c_int, [C.c_uint64, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]),
    "ksb_colors_count": (C.c_uint64, [C.c_void_p]),
    "ksb_colors_distinct_hashes": (C.c_uint64, [C.c_void_p]),
    "ksb_colors_offsets": (C.c_void_p, [C.c_void_p]),
    "ksb_colors_members": (C.c_void_p, [C.c_void_p]),
    "ksb_colors_counts": (C.c_void_p, [C.c_void_p]),
    "ksb_colors_totals": (C.c_void_p, [C.c_void_p]),
    "ksb_colors_device_ms": (C.c_float, [C.c_void_p]),
    "ksb_colors_free": (None, [C.c_void_p]),
    "ksb_format_float": (C.c_int, [C.c_float, C.c_char_p]),
    "ksb_ani_column": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]),
    "ksb_format_double_repr": (C.c_int, [C.c_double, C.c_char_p]),
    "ksb_write_ani_column": (C.c_int, [C.c_void_p, C.c_uint64, C.c_char_p, C.c_int]),
    "ksb_table_read_tsv": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]),
    "ksb_table_read_sidecar": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "ksb_table_get": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "ksb_table_free": (None, [C.c_void_p]),
    "ksb_write_sidecar": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int]),
    "ksb_write_sidecar_records": (C.c_int, [C.c_void_p, C.c_uint64, C.c_char_p, C.c_uint64, C.c_uint64, C.c_int]),
    "ksb_sidecar_matches": (C.c_int, [C.c_char_p, C.c_char_p]),
    "ksb_sidecar_stamp": (C.c_int, [C.c_char_p, C.c_char_p]),
    "ksb_cluster_table": (C.c_int, [C.c_void_p, C.c_uint64, C.c_double, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]),
    "ksb_ani_table": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]),
    "ksb_cluster_edges_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_double,
                                            C.c_void_p, C.c_void_p, C.c_void_p]),
    "ksb_write_index_files": (C.c_int, [C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64,
                                        C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    "ksb_bench_imma": (C.c_int, [C.c_uint32, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_double)]),
    "ksb_bench_red": (C.c_int, [C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float)]),
}

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is not built (run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "or `make -C kspider_b200/csrc`); kspider_b200 has no Python/CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(L, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def exported_names():
    return sorted(_PROTOS)


def check(rc: int) -> None:
    if rc != KSB_OK:
        raise KsbError(rc, (lib().ksb_last_error() or b"").decode(errors="replace"))
