"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/_build/libksb_oracle.so (pairwise_oracle.c).

Parity status: pinned against the unmodified reference binary (tests/test_oracle_pinning.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "libksb_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB_PATH) or \
            os.path.getmtime(LIB_PATH) < os.path.getmtime(os.path.join(HERE, "pairwise_oracle.c")):
        subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp, u64, cp = C.c_void_p, C.c_uint64, C.c_char_p
        L.ko_index_load.argtypes = [cp, C.POINTER(vp)]
        L.ko_index_free.argtypes = [vp]
        L.ko_pairwise.argtypes = [vp, C.POINTER(vp)]
        L.ko_pairwise_csr.argtypes = [u64, vp, vp, vp, u64, vp, vp, C.POINTER(vp)]
        L.ko_pairwise_csr_banded.argtypes = [u64, vp, vp, vp, u64, vp, vp, C.c_int, C.POINTER(vp)]
        L.ko_result_free.argtypes = [vp]
        L.ko_write_tsv.argtypes = [vp, cp]
        L.ko_write_seq_to_kmers.argtypes = [vp, cp]
        L.ko_format_float.argtypes = [C.c_float, C.c_char_p, C.c_size_t]
        for name in ("ko_index_n_colors", "ko_index_n_members", "ko_index_n_totals", "ko_result_n_rows"):
            getattr(L, name).argtypes = [vp]
            getattr(L, name).restype = u64
        for name in ("ko_index_color_ids", "ko_index_offsets", "ko_index_members", "ko_index_weights",
                     "ko_index_total_ids", "ko_index_total_vals", "ko_result_src1", "ko_result_src2",
                     "ko_result_shared", "ko_result_cmin", "ko_result_cavg", "ko_result_cmax"):
            getattr(L, name).argtypes = [vp]
            getattr(L, name).restype = vp
        _lib = L
    return _lib


def _arr(ptr, n, dtype):
    if n == 0:
        return np.empty(0, dtype=dtype)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).copy()


class OracleIndex:
    """Index as the reference sees it after src/pairwise.cpp:95-121,166-170."""

    def __init__(self, prefix: str):
        L = lib()
        h = C.c_void_p()
        rc = L.ko_index_load(prefix.encode(), C.byref(h))
        if rc != 0:
            raise RuntimeError(f"oracle: ko_index_load({prefix}) failed rc={rc}")
        self._h = h
        nc, nm, nt = L.ko_index_n_colors(h), L.ko_index_n_members(h), L.ko_index_n_totals(h)
        self.color_ids = _arr(L.ko_index_color_ids(h), nc, np.uint32)
        self.offsets = _arr(L.ko_index_offsets(h), nc + 1, np.uint64)
        self.members = _arr(L.ko_index_members(h), nm, np.uint32)
        self.weights = _arr(L.ko_index_weights(h), nc, np.uint32)
        self.total_ids = _arr(L.ko_index_total_ids(h), nt, np.uint32)
        self.total_vals = _arr(L.ko_index_total_vals(h), nt, np.uint32)

    def write_seq_to_kmers(self, path: str):
        assert lib().ko_write_seq_to_kmers(self._h, path.encode()) == 0

    def __del__(self):
        try:
            lib().ko_index_free(self._h)
        except Exception:
            pass


class OracleResult:
    def __init__(self, h):
        L = lib()
        self._h = h
        n = L.ko_result_n_rows(h)
        self.src1 = _arr(L.ko_result_src1(h), n, np.uint32)
        self.src2 = _arr(L.ko_result_src2(h), n, np.uint32)
        self.shared = _arr(L.ko_result_shared(h), n, np.uint64)
        self.cmin = _arr(L.ko_result_cmin(h), n, np.float32)
        self.cavg = _arr(L.ko_result_cavg(h), n, np.float32)
        self.cmax = _arr(L.ko_result_cmax(h), n, np.float32)

    def write_tsv(self, path: str):
        assert lib().ko_write_tsv(self._h, path.encode()) == 0

    def __del__(self):
        try:
            lib().ko_result_free(self._h)
        except Exception:
            pass


def pairwise_index(ix: OracleIndex) -> OracleResult:
    h = C.c_void_p()
    rc = lib().ko_pairwise(ix._h, C.byref(h))
    if rc != 0:
        raise RuntimeError(f"oracle: ko_pairwise rc={rc}")
    return OracleResult(h)


def pairwise_csr(offsets, members, weights, total_ids, total_vals) -> OracleResult:
    """offsets u64[C+1], members u32[M] (any order within a colour), weights u32[C];
    totals as parallel u32 arrays (sample ids are arbitrary u32)."""
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    members = np.ascontiguousarray(members, dtype=np.uint32)
    weights = np.ascontiguousarray(weights, dtype=np.uint32)
    total_ids = np.ascontiguousarray(total_ids, dtype=np.uint32)
    total_vals = np.ascontiguousarray(total_vals, dtype=np.uint32)
    h = C.c_void_p()
    rc = lib().ko_pairwise_csr(offsets.size - 1, offsets.ctypes.data, members.ctypes.data,
                               weights.ctypes.data, total_ids.size, total_ids.ctypes.data,
                               total_vals.ctypes.data, C.byref(h))
    if rc != 0:
        raise RuntimeError(f"oracle: ko_pairwise_csr rc={rc}")
    return OracleResult(h)


def pairwise_csr_banded(offsets, members, weights, total_ids, total_vals, threads: int = 0) -> OracleResult:
    """ko_pairwise_csr_banded: the same rows for indexes of BASELINE size (all host threads, banded accumulator)."""
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    members = np.ascontiguousarray(members, dtype=np.uint32)
    weights = np.ascontiguousarray(weights, dtype=np.uint32)
    total_ids = np.ascontiguousarray(total_ids, dtype=np.uint32)
    total_vals = np.ascontiguousarray(total_vals, dtype=np.uint32)
    h = C.c_void_p()
    rc = lib().ko_pairwise_csr_banded(offsets.size - 1, offsets.ctypes.data, members.ctypes.data,
                                      weights.ctypes.data, total_ids.size, total_ids.ctypes.data,
                                      total_vals.ctypes.data, int(threads) or (os.cpu_count() or 1), C.byref(h))
    if rc != 0:
        raise RuntimeError(f"oracle: ko_pairwise_csr_banded rc={rc}")
    return OracleResult(h)


def format_float(v: float) -> str:
    buf = C.create_string_buffer(64)
    lib().ko_format_float(C.c_float(v), buf, 64)
    return buf.value.decode()
