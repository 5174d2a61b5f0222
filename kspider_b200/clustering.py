"""Mirror of `kSpider cluster` (pykSpider/kSpider2/ks_clustering.py) with the threshold and the
connected components computed on the GPU (K6 + K7 behind ksb_cluster).

Same inputs (`<prefix>.namesMap`, and either a live pairwise Result or the binary records), same
output file name and format (`<prefix>_kSpider_clusters_{cutoff*100}%.tsv`, one comma-joined
line of names per component, singletons included — ks_clustering.py:31-32,129-137).  Line order
and member order inside a line are rustworkx-internal in the reference; here clusters are written
in ascending order of their smallest sample id and members in ascending id order.
"""
from __future__ import annotations

from typing import Dict, List

import numpy as np

from .api import DIST_COL, Index, Result


class Clusters:
    distance_to_col = {"min_cont": 3, "avg_cont": 4, "max_cont": 5}  # "ani" (col 6) is not on this path

    def __init__(self, index_prefix: str, cut_off_threshold: float, dist_type: str = "max_cont"):
        """cut_off_threshold is in PERCENT like the reference's constructor argument
        (main() passes float(cutoff) * 100, ks_clustering.py:157)."""
        if dist_type not in self.distance_to_col:
            raise ValueError("unknown distance!")
        self.index_prefix = index_prefix
        self.dist_type = dist_type
        self.cut_off_threshold = cut_off_threshold
        self.names_file = index_prefix + ".namesMap"
        self.output = index_prefix + f"_kSpider_clusters_{cut_off_threshold}%.tsv"
        self.names_map: Dict[int, str] = {}
        self.tsv_get_namesmap()
        self.labels = None
        self.n_edges = 0

    def tsv_get_namesmap(self) -> None:
        with open(self.names_file, "r") as f:
            next(f)
            for row in f:
                row = row.strip().split()
                self.names_map[int(row[0])] = row[1]

    def cluster_result(self, index: Index, result: Result) -> List[List[int]]:
        """Threshold + CC of `result` on the device; returns components as lists of ORIGINAL ids."""
        ids = index.sample_ids
        labels, self.n_edges = result.cluster(ids.size, self.dist_type, cut_pct=self.cut_off_threshold)
        return self._components(ids, labels)

    def _components(self, ids: np.ndarray, labels: np.ndarray) -> List[List[int]]:
        self.labels = labels
        in_index = dict(zip(ids.tolist(), labels.tolist()))
        groups: Dict[int, List[int]] = {}
        nxt = int(ids.size)
        for sid in self.names_map:  # every name is a node (ks_clustering.py:45-47)
            if sid in in_index:
                key = in_index[sid]
            else:  # a name without any index entry: isolated node
                key = nxt
                nxt += 1
            groups.setdefault(key, []).append(sid)
        comps = [sorted(g) for g in groups.values()]
        comps.sort(key=lambda g: g[0])
        return comps

    def write(self, components: List[List[int]]) -> None:
        with open(self.output, "w") as out:
            for comp in components:
                out.write(",".join(self.names_map[i] for i in comp) + "\n")


def cluster(index_prefix: str, cutoff: float = 0.0, distance_type: str = "max_cont", host_threads: int = 1,
            index: Index | None = None, result: Result | None = None, **opts) -> str:
    """`kSpider cluster -i prefix -c cutoff -d distance_type` (ks_clustering.py:150-163) fed from the
    index files: runs the pairwise pass on the device when no live result is given, clusters on the
    device and writes the cluster file.  Returns the output path."""
    cut = float(cutoff) * 100
    kcl = Clusters(index_prefix, cut, distance_type)
    own_index = index is None
    if index is None:
        index = Index.load(index_prefix, host_threads)
    own_result = result is None
    if result is None:
        result = index.pairwise(want_records=False, **opts)
    comps = kcl.cluster_result(index, result)
    kcl.write(comps)
    if own_result:
        result.close()
    if own_index:
        index.close()
    return kcl.output
