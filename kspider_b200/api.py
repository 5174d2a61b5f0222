"""Thin object layer over the C-ABI: Index (host CSR), Device (resident CSR + owned triangle
block), Result (records / labels).  Everything that computes happens inside libkspider_b200.so."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import capi
from .capi import IndexInfo, Opts, Record, ResultInfo, check, lib

RECORD_DTYPE = np.dtype([("source_1", "<u4"), ("source_2", "<u4"), ("shared_kmers", "<u8"),
                         ("min_containment", "<f4"), ("avg_containment", "<f4"), ("max_containment", "<f4"),
                         ("pad", "<u4")])
assert RECORD_DTYPE.itemsize == C.sizeof(Record) == 32

DIST_COL = {"min_cont": 3, "avg_cont": 4, "max_cont": 5}  # ks_clustering.py:12-17 ("ani": 6 is a next-row)


def _view(ptr: int, n: int, dtype) -> np.ndarray:
    dt = np.dtype(dtype)
    if n == 0 or not ptr:
        return np.empty(0, dtype=dt)
    buf = (C.c_char * (n * dt.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dt)


def make_opts(device: int = -1, rank: int = 0, world: int = 1, want_records: bool = True,
              dense_budget_bytes: int = 0, spill: str = "auto") -> Opts:
    o = Opts()
    o.device, o.rank, o.world = device, rank, world
    o.want_records = 2 if want_records == "host" else (1 if want_records else 0)  # "host": host records only (ksb_opts.want_records = 2)
    o.dense_budget_bytes = int(dense_budget_bytes)
    o.spill_mode = {"auto": 0, "always": 1, "never": 2}[spill]
    return o


class Index:
    """Parsed index = the state of the reference after src/pairwise.cpp:194."""

    def __init__(self, handle: C.c_void_p):
        self._h = handle

    @classmethod
    def load(cls, index_prefix: str, host_threads: int = 1) -> "Index":
        h = C.c_void_p()
        check(lib().ksb_index_load(str(index_prefix).encode(), int(host_threads), C.byref(h)))
        return cls(h)

    @classmethod
    def load_colored(cls, index_prefix: str, host_threads: int = 1, device: int = -1) -> "Index":
        """From a kProcessor colored_kDataFrame file set (<prefix>.phmap, <prefix>colors.intvectors, <prefix>.namesMap):
        ksb_index_load_colored -- per-colour counts by a histogram of the k-mer map on the device."""
        h = C.c_void_p()
        check(lib().ksb_index_load_colored(str(index_prefix).encode(), int(host_threads), int(device), C.byref(h)))
        return cls(h)

    @classmethod
    def from_csr(cls, offsets, members, weights, total_ids, total_vals, host_threads: int = 1) -> "Index":
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        members = np.ascontiguousarray(members, dtype=np.uint32)
        weights = np.ascontiguousarray(weights, dtype=np.uint32)
        total_ids = np.ascontiguousarray(total_ids, dtype=np.uint32)
        total_vals = np.ascontiguousarray(total_vals, dtype=np.uint32)
        if offsets.size < 1 or weights.size != offsets.size - 1 or total_ids.size != total_vals.size:
            raise ValueError("inconsistent CSR arrays")
        h = C.c_void_p()
        check(lib().ksb_index_from_csr(offsets.size - 1, offsets.ctypes.data, members.ctypes.data,
                                       weights.ctypes.data, total_ids.size, total_ids.ctypes.data,
                                       total_vals.ctypes.data, int(host_threads), C.byref(h)))
        return cls(h)

    @property
    def info(self) -> IndexInfo:
        inf = IndexInfo()
        check(lib().ksb_index_get_info(self._h, C.byref(inf)))
        return inf

    @property
    def sample_ids(self) -> np.ndarray:
        return _view(lib().ksb_index_sample_ids(self._h), self.info.n_samples, np.uint32).copy()

    @property
    def totals(self) -> np.ndarray:
        return _view(lib().ksb_index_totals(self._h), self.info.n_samples, np.uint32).copy()

    def payload(self) -> np.ndarray:
        """The binned CSR blob (view, valid while the Index lives)."""
        n = C.c_uint64()
        p = lib().ksb_index_payload(self._h, C.byref(n))
        return _view(p, n.value, np.uint8)

    def descriptor(self) -> bytes:
        n = C.c_uint64()
        p = lib().ksb_index_descriptor(self._h, C.byref(n))
        return C.string_at(p, n.value)

    def bands(self):
        """(lo[N], hi[N], band_cells): the columns every row of the accumulator stores."""
        lo, hi, n = C.c_void_p(), C.c_void_p(), C.c_uint64()
        check(lib().ksb_index_bands(self._h, C.byref(lo), C.byref(hi), C.byref(n)))
        ns = int(self.info.n_samples)
        return _view(lo.value, ns, np.uint32).copy(), _view(hi.value, ns, np.uint32).copy(), n.value

    def row_cuts(self, world: int) -> np.ndarray:
        cuts = np.empty(world + 1, dtype=np.uint64)
        check(lib().ksb_index_row_cuts(self._h, int(world), cuts.ctypes.data))
        return cuts

    def shard(self, rank: int, world: int, host_threads: int = 1) -> "Index":
        """Row-block shard for `rank` of `world`: only the colours (cut down to the members) that the
        owner of that row block needs, re-binned (ksb_index_shard).  A Device made from it owns those rows."""
        h = C.c_void_p()
        check(lib().ksb_index_shard(self._h, int(rank), int(world), int(host_threads), C.byref(h)))
        return Index(h)

    @classmethod
    def from_blob(cls, descriptor: bytes, payload: np.ndarray) -> "Index":
        """Host index of THIS process from another rank's (descriptor(), payload()) — ksb_index_from_blob."""
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        buf = C.create_string_buffer(bytes(descriptor), len(descriptor))
        h = C.c_void_p()
        check(lib().ksb_index_from_blob(C.cast(buf, C.c_void_p), len(descriptor), payload.ctypes.data, payload.size, C.byref(h)))
        return cls(h)

    def write_seq_to_kmers(self, path: str) -> None:
        check(lib().ksb_write_seq_to_kmers(self._h, str(path).encode()))

    def upload(self, **kw) -> "Device":
        o = make_opts(**kw)
        h = C.c_void_p()
        check(lib().ksb_upload(self._h, C.byref(o), C.byref(h)))
        return Device(h)

    def pairwise(self, **kw) -> "Result":
        o = make_opts(**kw)
        h = C.c_void_p()
        check(lib().ksb_pairwise(self._h, C.byref(o), C.byref(h)))
        return Result(h)

    def pairwise_cluster(self, dist_type: str = "max_cont", cutoff01: float = 0.0, cut_pct=None, d_labels_ptr: int = 0,
                         to_host: bool = True, **kw):
        """ksb_pairwise_cluster: upload + cluster-only step + release.  Returns (labels[N] or None, kept edges,
        touched pairs, info).  d_labels_ptr: device uint32[N] that also receives the labels."""
        o = make_opts(**{**kw, "want_records": False})
        labels = np.empty(int(self.info.n_samples), dtype=np.uint32) if to_host else None
        ne, nr, inf = C.c_uint64(), C.c_uint64(), ResultInfo()
        cut = float(cutoff01) * 100 if cut_pct is None else float(cut_pct)
        check(lib().ksb_pairwise_cluster(self._h, C.byref(o), DIST_COL[dist_type], cut, labels.ctypes.data if to_host else None,
                                         C.c_void_p(d_labels_ptr) if d_labels_ptr else None, C.byref(ne), C.byref(nr), C.byref(inf)))
        return labels, ne.value, nr.value, inf

    def close(self) -> None:
        if self._h:
            lib().ksb_index_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Device:
    def __init__(self, handle: C.c_void_p, keepalive=None):
        self._h = handle
        self._keep = keepalive  # e.g. the torch tensor holding an adopted payload
        self._first_steps_left = 2  # distributed.cluster_step: steps during which the ranks agree on a repeat

    @classmethod
    def adopt(cls, descriptor: bytes, d_payload_ptr: int, payload_bytes: int, keepalive=None, **kw) -> "Device":
        o = make_opts(**kw)
        h = C.c_void_p()
        buf = C.create_string_buffer(descriptor, len(descriptor))
        check(lib().ksb_adopt_device_payload(buf, len(descriptor), C.c_void_p(d_payload_ptr), payload_bytes,
                                             C.byref(o), C.byref(h)))
        return cls(h, keepalive)

    @property
    def n_passes(self) -> int:
        return lib().ksb_device_n_passes(self._h)

    def set_pass(self, p: int) -> None:
        check(lib().ksb_device_set_pass(self._h, p))

    def set_overlap(self, on: bool) -> None:
        """K4 beside K1/K2 on a second stream (default) or one kernel after the other (per-kernel timings)."""
        check(lib().ksb_device_set_overlap(self._h, 1 if on else 0))

    def accumulate(self, stream: int = 0, sync: bool = True) -> None:
        check(lib().ksb_accumulate(self._h, C.c_void_p(stream) if stream else None, 1 if sync else 0))

    @property
    def info(self) -> ResultInfo:
        inf = ResultInfo()
        check(lib().ksb_device_get_info(self._h, C.byref(inf)))
        return inf

    def triangle(self):
        """(device pointer, cells, row_begin, row_end) of the current pass."""
        c, rb, re = C.c_uint64(), C.c_uint64(), C.c_uint64()
        p = lib().ksb_device_triangle(self._h, C.byref(c), C.byref(rb), C.byref(re))
        return p, c.value, rb.value, re.value

    @property
    def cell_bits(self) -> int:
        """32 when the index proves that no cell of the accumulator can exceed 2^32 - 1, else 64 (ksb_device_cell_bits)."""
        return int(lib().ksb_device_cell_bits(self._h))

    def read_triangle(self) -> np.ndarray:
        _, cells, _, _ = self.triangle()
        out = np.empty(cells, dtype=np.uint64)
        check(lib().ksb_device_read_triangle(self._h, out.ctypes.data, cells))
        return out

    def postpass(self, want_records: bool = True, into: Optional["Result"] = None) -> "Result":
        h = into._h if into is not None else C.c_void_p()
        check(lib().ksb_postpass(self._h, 1 if want_records else 0, C.byref(h)))
        if into is not None:
            return into
        return Result(h)

    @property
    def is_spill(self) -> bool:
        return bool(lib().ksb_device_is_spill(self._h))

    def run_all_passes(self, want_records: bool = True) -> "Result":
        if self.is_spill:  # K3: no dense passes, one sort-reduce run
            h = C.c_void_p()
            check(lib().ksb_device_run_spill(self._h, 1 if want_records else 0, C.byref(h)))
            return Result(h)
        res = None
        for p in range(self.n_passes):
            self.set_pass(p)
            self.accumulate(sync=True)
            res = self.postpass(want_records, into=res)
        return res

    def cluster(self, n_samples: int, dist_type: str = "max_cont", cutoff01: float = 0.0, cut_pct=None, stream: int = 0,
                d_labels_ptr: int = 0, to_host: bool = True):
        """Cluster-only step (ksb_device_cluster): accumulate + one fused scan per pass + unions; no records.
        Returns (labels[N] or None, kept edges, touched pairs).  d_labels_ptr: device uint32[N] that also receives
        the labels (multi-GPU merge); stream: a cudaStream_t to enqueue on (0 = the library's)."""
        labels = np.empty(n_samples, dtype=np.uint32) if to_host else None
        ne, nr = C.c_uint64(), C.c_uint64()
        cut = float(cutoff01) * 100 if cut_pct is None else float(cut_pct)
        check(lib().ksb_device_cluster(self._h, DIST_COL[dist_type], cut, C.c_void_p(stream) if stream else None,
                                       labels.ctypes.data if to_host else None,
                                       C.c_void_p(d_labels_ptr) if d_labels_ptr else None, C.byref(ne), C.byref(nr)))
        return labels, ne.value, nr.value

    def set_graph(self, on: bool) -> None:
        """the cluster-only step as one CUDA graph launch (default) or kernel by kernel (keeps the per-kernel timers)"""
        check(lib().ksb_device_set_graph(self._h, 1 if on else 0))

    def cluster_enqueue(self, dist_type: str, cutoff01: float, stream: int = 0, d_labels_ptr: int = 0, host_labels: bool = False) -> None:
        check(lib().ksb_device_cluster_enqueue(self._h, DIST_COL[dist_type], float(cutoff01) * 100, C.c_void_p(stream) if stream else None,
                                               C.c_void_p(d_labels_ptr) if d_labels_ptr else None, 1 if host_labels else 0))

    def cluster_collect(self, n_samples: int = 0, want_labels: bool = False):
        """(labels or None, kept edges, touched pairs); raises KsbError(KSB_ERR_BUDGET) when the step must be repeated"""
        labels = np.empty(n_samples, dtype=np.uint32) if want_labels else None
        ne, nr = C.c_uint64(), C.c_uint64()
        check(lib().ksb_device_cluster_collect(self._h, labels.ctypes.data if want_labels else None, C.byref(ne), C.byref(nr)))
        return labels, ne.value, nr.value

    def close(self) -> None:
        if self._h:
            lib().ksb_device_free(self._h)
            self._h = None
            self._keep = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Result:
    def __init__(self, handle: C.c_void_p):
        self._h = handle

    @property
    def info(self) -> ResultInfo:
        inf = ResultInfo()
        check(lib().ksb_result_get_info(self._h, C.byref(inf)))
        return inf

    def records(self) -> np.ndarray:
        n = C.c_uint64()
        p = lib().ksb_result_records(self._h, C.byref(n))
        return _view(p, n.value, RECORD_DTYPE).copy()

    def write_tsv(self, path: str, host_threads: int = 1, append: bool = False) -> None:
        fn = lib().ksb_append_tsv if append else lib().ksb_write_tsv
        check(fn(self._h, str(path).encode(), int(host_threads)))

    def cluster(self, n_samples: int, dist_type: str = "max_cont", cutoff01: float = 0.0, cut_pct=None):
        """labels[N] (label = smallest compact sample index of the component), number of kept edges.
        cut_pct (the reference's cut_off_threshold = float(cutoff)*100) overrides cutoff01."""
        labels = np.empty(n_samples, dtype=np.uint32)
        ne = C.c_uint64()
        if cut_pct is None:
            check(lib().ksb_cluster(self._h, DIST_COL[dist_type], float(cutoff01), labels.ctypes.data, C.byref(ne)))
        else:
            check(lib().ksb_cluster_pct(self._h, DIST_COL[dist_type], float(cut_pct), labels.ctypes.data, C.byref(ne)))
        return labels, ne.value

    def ani_column(self, ksize: int):
        """avg_ani of every record, in record order (K8 on the result's device; ks_pairwise.py:29-82).
        Returns (float64[n_rows], number of rows for which the reference would raise)."""
        n = int(self.info.n_rows)
        out = np.empty(n, dtype=np.float64)
        bad = C.c_uint64()
        check(lib().ksb_ani_column(self._h, int(ksize), out.ctypes.data, C.byref(bad)))
        return out, bad.value

    def cluster_into(self, d_labels_ptr: int, dist_type: str = "max_cont", cutoff01: float = 0.0) -> int:
        """Same as cluster(), labels written to DEVICE memory at `d_labels_ptr` (uint32[N] on this result's
        device; e.g. a torch tensor's data_ptr()).  Returns the number of kept edges."""
        ne = C.c_uint64()
        check(lib().ksb_cluster_pct_device(self._h, DIST_COL[dist_type], float(cutoff01) * 100.0, C.c_void_p(d_labels_ptr), C.byref(ne)))
        return ne.value

    def close(self) -> None:
        if self._h:
            lib().ksb_result_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def threshold_float(cutoff01: float) -> float:
    return float(lib().ksb_threshold_float(float(cutoff01)))


def format_float(v: float) -> str:
    buf = C.create_string_buffer(32)
    n = lib().ksb_format_float(C.c_float(v), buf)
    return buf.raw[:n].decode()


def format_double_repr(v: float) -> str:
    buf = C.create_string_buffer(40)
    n = lib().ksb_format_double_repr(C.c_double(v), buf)
    return buf.raw[:n].decode()


def write_ani_column(ani: np.ndarray, path: str, host_threads: int = 1) -> None:
    a = np.ascontiguousarray(ani, dtype=np.float64)
    check(lib().ksb_write_ani_column(a.ctypes.data, a.size, str(path).encode(), int(host_threads)))


def cluster_merge(labels_a: np.ndarray, labels_b: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(labels_a, dtype=np.uint32)
    b = np.ascontiguousarray(labels_b, dtype=np.uint32)
    out = np.empty_like(a)
    check(lib().ksb_cluster_merge(a.size, a.ctypes.data, b.ctypes.data, out.ctypes.data))
    return out


def write_index_files(prefix: str, color_ids, offsets, members, count_ids, count_vals, total_ids, total_vals) -> None:
    color_ids = np.ascontiguousarray(color_ids, dtype=np.uint64)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    members = np.ascontiguousarray(members, dtype=np.uint32)
    count_ids = np.ascontiguousarray(count_ids, dtype=np.uint64)
    count_vals = np.ascontiguousarray(count_vals, dtype=np.uint64)
    total_ids = np.ascontiguousarray(total_ids, dtype=np.uint32)
    total_vals = np.ascontiguousarray(total_vals, dtype=np.uint32)
    check(lib().ksb_write_index_files(str(prefix).encode(), color_ids.size, color_ids.ctypes.data, offsets.ctypes.data,
                                      members.ctypes.data, count_ids.size, count_ids.ctypes.data, count_vals.ctypes.data,
                                      total_ids.size, total_ids.ctypes.data, total_vals.ctypes.data))


def bench_red(cells: int, n_ops: int, cell_bits: int = 64, pattern: int = 0, reps: int = 5) -> float:
    ms = C.c_float()
    check(lib().ksb_bench_red(int(cells), int(n_ops), cell_bits, pattern, reps, C.byref(ms)))
    return ms.value


def bench_imma(iters: int = 2048, reps: int = 5):
    """(best ms, u8 MACs per launch) of the int8 tcgen05 peak kernel in K4's instruction shape"""
    ms, macs = C.c_float(), C.c_double()
    check(lib().ksb_bench_imma(int(iters), int(reps), C.byref(ms), C.byref(macs)))
    return ms.value, macs.value
