"""Drop-in for `kSpider cluster` (pykSpider/kSpider2/ks_clustering.py) with the threshold and the connected
components on the GPU.

`Clusters` has the reference's constructor and methods, so that in ks_clustering.py:150-163 only the class
changes: `Clusters(logger_obj=ctx.obj, index_prefix=..., cut_off_threshold=..., dist_type=...)`,
`.construct_graph()`, `.cluster_graph()`.  Same inputs -- `<prefix>.namesMap`, `<prefix>_kSpider_seqToKmersNo.tsv`,
`<prefix>_kSpider_pairwise.tsv` and, for dist type "ani", `<prefix>_kSpider_pairwise.ani_col.tsv` -- same
output file name and format (`<prefix>_kSpider_clusters_{cut_off_threshold}%.tsv`, one comma-joined line of
names per component, singletons included, :31-32,129-137), same messages.

construct_graph() reads the rows of the pairwise TSV ON DISK (whichever tool wrote it): from the binary
side-car `<prefix>_kSpider_pairwise.ksb` when our pairwise wrote one beside a TSV of exactly this size
(ksb_sidecar_matches), otherwise by parsing the text (multithreaded, ksb_table_read_tsv).  cluster_graph() runs
ksb_cluster_table: node = id - 1, row dropped iff float(text) * 100 < cut_off_threshold, lock-free union-find.

Differences, all outside what the file contract defines: line order and member order inside a line are
rustworkx-internal in the reference, here clusters come in ascending order of their smallest node and
members in ascending node order; the reference loses one kept edge whenever a 10 M-edge batch fills up
(:107-113), here every kept edge counts; `names_map` / `seq_to_kmers` are per instance,
not class attributes shared by every instance (:19-20).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Set

import numpy as np

from .api import Index, Result
from .capi import check, lib


def _host_threads() -> int:
    v = os.environ.get("KSPIDER_THREADS")
    return max(1, int(v)) if v else max(1, min(32, os.cpu_count() or 1))


class _Table:
    """Parsed rows of a pairwise TSV / side-car (ksb_table)."""

    def __init__(self, handle: C.c_void_p, source: str):
        self._h = handle
        self.source = source  # "sidecar" or "tsv"

    @classmethod
    def read(cls, index_prefix: str, dist_col: int, dist_col2: int = 0, threads: Optional[int] = None) -> "_Table":
        tsv = index_prefix + "_kSpider_pairwise.tsv"
        side = index_prefix + "_kSpider_pairwise.ksb"
        threads = threads or _host_threads()
        h = C.c_void_p()
        use_side = dist_col != 6 and os.environ.get("KSPIDER_SIDECAR", "1") != "0" and \
            lib().ksb_sidecar_matches(side.encode(), tsv.encode()) == 1
        if use_side:
            check(lib().ksb_table_read_sidecar(side.encode(), dist_col, dist_col2, threads, C.byref(h)))
            return cls(h, "sidecar")
        ani = (index_prefix + "_kSpider_pairwise.ani_col.tsv").encode() if dist_col == 6 else None
        check(lib().ksb_table_read_tsv(tsv.encode(), dist_col, dist_col2, ani, threads, C.byref(h)))
        return cls(h, "tsv")

    @property
    def n_rows(self) -> int:
        n = C.c_uint64()
        check(lib().ksb_table_get(self._h, C.byref(n), None, None, None, None))
        return n.value

    def close(self) -> None:
        if self._h:
            lib().ksb_table_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Clusters:

    distance_to_col = {
        "min_cont": 3,
        "avg_cont": 4,
        "max_cont": 5,
        "ani": 6,
    }

    def __init__(self, logger_obj, index_prefix, cut_off_threshold, dist_type):
        """ks_clustering.py:22-47.  cut_off_threshold is in PERCENT (main() passes float(cutoff) * 100, :157)."""
        self.seq_to_kmers = dict()
        self.names_map = dict()
        self.index_prefix = index_prefix
        self.Logger = logger_obj
        self.dist_type = dist_type
        self.edges_batch_number = 10_000_000  # kept for attribute parity; no edge is dropped at a batch boundary here
        self.names_file = index_prefix + ".namesMap"
        self.cut_off_threshold = cut_off_threshold
        self.seqToKmers_file = index_prefix + "_kSpider_seqToKmersNo.tsv"
        self.pairwise_file = index_prefix + "_kSpider_pairwise.tsv"
        self.output = index_prefix + \
            f"_kSpider_clusters_{cut_off_threshold}%.tsv"
        self.shared_kmers_threshold = 200
        self.Logger.INFO("Loading TSV pairwise file")
        self.load_seq_to_kmers(self.seqToKmers_file)
        self.tsv_get_namesmap()
        if dist_type not in self.distance_to_col:
            logger_obj.ERROR("unknown distance!")
        self.dist_col = self.distance_to_col[dist_type]

        if dist_type == "ani":
            if not os.path.exists(self.index_prefix + "_kSpider_pairwise.ani_col.tsv"):
                logger_obj.ERROR(f"ANI was selected, but the ani file {self.index_prefix}_kSpider_pairwise.ani_col.tsv was not found!")

        # the reference's graph: one node per namesMap entry, node index = position (rx.PyGraph.add_nodes_from, :45-47)
        self.nodes_indeces = list(range(len(self.names_map)))
        self.device = int(os.environ.get("KSPIDER_DEVICE", "-1"))
        self._table: Optional[_Table] = None
        self.labels: Optional[np.ndarray] = None
        self.n_edges = 0
        self.connected_components: List[Set[int]] = []

    def load_seq_to_kmers(self, tsv):
        with open(tsv) as KMER_COUNT:
            next(KMER_COUNT)
            for line in KMER_COUNT:
                seq_ID, no_of_kmers = tuple(line.strip().split('\t')[1:])
                self.seq_to_kmers[int(seq_ID)] = int(no_of_kmers)

    def tsv_get_namesmap(self):
        with open(self.names_file, 'r') as namesMap:
            next(namesMap)  # skip the header
            for row in namesMap:
                row = row.strip().split()
                self.names_map[int(row[0])] = row[1]

    def construct_graph(self):
        """ks_clustering.py:63-117: the rows of the pairwise TSV (+ the ani column) become the edge table."""
        print("[i] constructing graph")
        if self._table is not None:
            self._table.close()
        self._table = _Table.read(self.index_prefix, self.dist_col)

    def cluster_graph(self):
        """ks_clustering.py:119-137: connected components (on the GPU), log lines, cluster file."""
        if self._table is None:
            raise RuntimeError("construct_graph() has not been called")
        n_nodes = len(self.names_map)
        labels = np.empty(n_nodes, dtype=np.uint32)
        ne = C.c_uint64()
        check(lib().ksb_cluster_table(self._table._h, n_nodes, float(self.cut_off_threshold), self.device,
                                      labels.ctypes.data, C.byref(ne)))
        self.labels, self.n_edges = labels, ne.value
        self.connected_components = components_of_labels(labels)
        self.Logger.INFO(
            f"number of clusters: {len(self.connected_components)}")
        retworkx_export = self.index_prefix + \
            f"_kSpider_graph_{self.cut_off_threshold}%.json"
        self.Logger.INFO(f"writing {retworkx_export}")
        with open(self.output, 'w') as CLUSTERS:
            for component in self.connected_components:
                named_component = [self.names_map[node + 1]
                                   for node in sorted(component)]
                CLUSTERS.write(','.join(named_component) + '\n')


def components_of_labels(labels: np.ndarray) -> List[Set[int]]:
    """label per node (= smallest node of its component) -> list of sets of node indices, ascending by label"""
    order = np.argsort(labels, kind="stable")
    cuts = np.flatnonzero(np.diff(labels[order])) + 1
    return [set(int(v) for v in grp) for grp in np.split(order, cuts)] if labels.size else []


class _QuietLogger:
    """What ctx.obj is in the reference (kSpider2/customLogger.py) with -q unset: silent INFO, fatal ERROR."""

    def INFO(self, msg):
        pass

    SUCCESS = WARNING = INFO

    def ERROR(self, msg):
        import sys
        print(f"[ERROR] {msg}", file=sys.stderr)
        sys.exit(1)


def cluster(index_prefix: str, cutoff: float = 0.0, distance_type: str = "max_cont", logger_obj=None) -> str:
    """The body of `kSpider cluster -i prefix -c cutoff -d distance_type` (ks_clustering.py:155-163).
    Returns the path of the cluster file."""
    cutoff = float(cutoff) * 100
    kCl = Clusters(logger_obj=logger_obj or _QuietLogger(), index_prefix=index_prefix,
                   cut_off_threshold=cutoff, dist_type=distance_type)
    kCl.construct_graph()
    kCl.cluster_graph()
    return kCl.output


def cluster_result(index: Index, result: Result, names_map: dict, output: str, cutoff: float = 0.0,
                   distance_type: str = "max_cont") -> List[List[int]]:
    """Same file from a LIVE pairwise result (records still on the device, nothing read back from disk):
    K6 + K7 behind ksb_cluster.  Components as lists of ORIGINAL sample ids; every name is a node."""
    ids = index.sample_ids
    labels, _ = result.cluster(ids.size, distance_type, cut_pct=float(cutoff) * 100)
    in_index = dict(zip(ids.tolist(), labels.tolist()))
    groups: dict = {}
    nxt = int(ids.size)
    for sid in names_map:
        if sid in in_index:
            key = in_index[sid]
        else:  # a name without any index entry: isolated node
            key = nxt
            nxt += 1
        groups.setdefault(key, []).append(sid)
    comps = sorted((sorted(g) for g in groups.values()), key=lambda g: g[0])
    with open(output, "w") as out:
        for comp in comps:
            out.write(",".join(names_map[i] for i in comp) + "\n")
    return comps
