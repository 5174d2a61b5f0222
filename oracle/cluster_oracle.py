"""TEST INFRASTRUCTURE ONLY — CPU restatement of `kSpider cluster`
(pykSpider/kSpider2/ks_clustering.py:22-137,150-163).

Parity status: PINNED against the unmodified reference for everything ks_clustering.py itself does —
tests/test_cluster_pinning.py compares this restatement with (1) the committed goldens
tests/golden/*.clusters.json, written by the UNMODIFIED ks_clustering.py run through oracle/refpy.py on
the unmodified src/pairwise.cpp's TSVs (every dist type incl. ani, nine cut-offs, five cases), and (2)
the reference run live on fresh inputs where /root/reference exists.  The one thing still "unpinned" is
rustworkx.connected_components itself (third-party Rust module, listed unpinned in setup.py:203-210,
not vendored, not installed here; oracle/py_shim/rustworkx.py stands in for it with scipy): connected
components of an undirected graph are unique as a PARTITION, which is what is compared.  The
threshold arithmetic (text -> Python float -> *100 -> `<`) is plain Python and is followed exactly.

Deliberately NOT restated (reference defect): ks_clustering.py:107-113 drops
the current edge whenever a 10 M-edge batch fills up; which edge is lost depends on the hash-map
row order of the TSV, so it is not reproducible.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np
from scipy.sparse import coo_matrix
from scipy.sparse.csgraph import connected_components

DISTANCE_TO_COL = {"min_cont": 3, "avg_cont": 4, "max_cont": 5, "ani": 6}  # ks_clustering.py:12-17


def load_names_map(path: str) -> Dict[int, str]:
    """ks_clustering.py:56-61 — skip the count header, `int(id) -> name`, dict keeps file order."""
    names: Dict[int, str] = {}
    with open(path) as f:
        next(f)
        for row in f:
            row = row.strip().split()
            names[int(row[0])] = row[1]
    return names


def edges_from_tsv(pairwise_tsv: str, dist_col: int, cutoff01: float, ani_col_tsv: str | None = None) -> List[Tuple[int, int]]:
    """ks_clustering.py:97-105,157 — cut_off_threshold = float(cutoff)*100 ;
    distance = float(text)*100 ; the row is skipped iff distance < cut_off_threshold
    (so NaN distances are KEPT).  dist_col 6 (:70-83): the distance is the same row of the one-column
    `_kSpider_pairwise.ani_col.tsv`, float(line) * 100.0."""
    cut = float(cutoff01) * 100
    edges = []
    with open(pairwise_tsv) as f:
        next(f)
        ani = None
        if dist_col == 6:
            ani = open(ani_col_tsv)
            next(ani)
        for row in f:
            row = row.strip().split("\t")
            seq1 = int(row[0]) - 1
            seq2 = int(row[1]) - 1
            distance = float(next(ani).strip()) * 100.0 if ani is not None else float(row[dist_col]) * 100
            if distance < cut:
                continue
            edges.append((seq1, seq2))
        if ani is not None:
            ani.close()
    return edges


def components(n_nodes: int, edges: List[Tuple[int, int]]) -> np.ndarray:
    """Stand-in for rx.connected_components (ks_clustering.py:120): label per node index."""
    if edges:
        e = np.asarray(edges, dtype=np.int64)
        g = coo_matrix((np.ones(len(e), dtype=np.int8), (e[:, 0], e[:, 1])), shape=(n_nodes, n_nodes))
    else:
        g = coo_matrix((n_nodes, n_nodes), dtype=np.int8)
    _, labels = connected_components(g, directed=False)
    return labels


def cluster(index_prefix: str, cutoff01: float, dist_type: str = "max_cont") -> List[List[str]]:
    """Whole command: returns the partition as sorted lists of names (canonical form: members
    sorted, clusters sorted) — line/member order in the reference's file is rustworkx-internal."""
    names = load_names_map(index_prefix + ".namesMap")
    node_ids = list(names.keys())  # node index = position in file (ks_clustering.py:45-47)
    edges = edges_from_tsv(index_prefix + "_kSpider_pairwise.tsv", DISTANCE_TO_COL[dist_type], cutoff01,
                           index_prefix + "_kSpider_pairwise.ani_col.tsv")
    labels = components(len(node_ids), edges)
    groups: Dict[int, List[str]] = {}
    for node, lab in enumerate(labels):
        groups.setdefault(int(lab), []).append(names[node + 1])  # names_map[node + 1] (:135)
    return canonical(list(groups.values()))


def canonical(clusters: List[List[str]]) -> List[List[str]]:
    return sorted(sorted(c) for c in clusters)


def read_clusters_file(path: str) -> List[List[str]]:
    with open(path) as f:
        return canonical([line.rstrip("\n").split(",") for line in f if line.rstrip("\n")])


def output_name(index_prefix: str, cutoff01: float) -> str:
    """ks_clustering.py:31-32,157 — f"_kSpider_clusters_{float(cutoff)*100}%.tsv"."""
    return index_prefix + f"_kSpider_clusters_{float(cutoff01) * 100}%.tsv"


# ---- the same decisions on float32 columns (for BASELINE-size results, where formatting 5e7 rows in Python
# ---- is not an option): pinned against the text rule above in tests/test_cluster_pinning.py

def threshold_float32(cut_pct: float) -> float:
    """Smallest float32 c >= 0 whose printed text the reference keeps: float("%g" % c) * 100 >= cut_pct
    (ks_clustering.py:101-105 on src/pairwise.cpp:266-273's text).  "%g" and float() are monotone, so the
    kept set is {c >= c*} (plus NaN): binary search over the bit patterns +0 .. +inf.  cut_pct above every
    finite value returns inf (only inf / NaN rows survive)."""
    def kept(bits: int) -> bool:
        c = float(np.array([bits], dtype=np.uint32).view(np.float32)[0])
        return not (float("%g" % c) * 100 < cut_pct)
    lo, hi = 0, 0x7F800000
    if kept(lo):
        return 0.0
    while hi - lo > 1:
        mid = (lo + hi) // 2
        if kept(mid):
            hi = mid
        else:
            lo = mid
    return float(np.array([hi], dtype=np.uint32).view(np.float32)[0])


def kept_mask(col: np.ndarray, cut_pct: float) -> np.ndarray:
    """rows of a float32 containment column the reference keeps at cut_off_threshold = cut_pct"""
    cstar = np.float32(threshold_float32(cut_pct))
    return ~(np.asarray(col, dtype=np.float32) < cstar)


def min_labels(n_nodes: int, a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """label[v] = smallest node of v's component (the form ksb_cluster returns) for the edge list (a, b)"""
    g = coo_matrix((np.ones(a.size, dtype=np.int8), (a, b)), shape=(n_nodes, n_nodes))
    _, lab = connected_components(g, directed=False)
    first = np.full(lab.max() + 1 if lab.size else 0, n_nodes, dtype=np.int64)
    np.minimum.at(first, lab, np.arange(n_nodes))
    return first[lab].astype(np.uint32)
