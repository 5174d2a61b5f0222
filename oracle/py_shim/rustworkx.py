"""TEST INFRASTRUCTURE ONLY — stand-in for the `rustworkx` module, so that the UNMODIFIED
/root/reference/pykSpider/kSpider2/ks_clustering.py can be imported and run in this image
(rustworkx is third-party, unpinned in the reference's setup.py:203-210, absent here).

Only the surface ks_clustering.py touches exists (:45-47, 89-95, 111-120):
    PyGraph(), PyGraph.add_nodes_from(list) -> node indices, PyGraph.add_edges_from([(a, b, w), ...]),
    connected_components(graph) -> list[set[int]]
with rustworkx's documented semantics: nodes are indexed 0.. in insertion order, an edge names node
INDICES and an index that does not exist is an error, parallel edges are allowed, and the components of an
undirected graph are returned as sets of node indices.  The partition is computed by
scipy.sparse.csgraph.connected_components; the ORDER of the components (and of the members inside a
set) is rustworkx-internal and is not reproduced -- compare partitions.
What this pins: everything ks_clustering.py does around the call (text -> float -> x100 -> `<`, the
10 M-edge batching with its dropped edge, node = id - 1, names_map[node + 1], the output file).  What it
cannot pin: rustworkx's own connected-components code.
"""
import numpy as np
from scipy.sparse import coo_matrix
from scipy.sparse.csgraph import connected_components as _cc


class PyGraph:
    def __init__(self, multigraph=True):
        self._nodes = []
        self._a = []
        self._b = []

    def add_nodes_from(self, payloads):
        first = len(self._nodes)
        self._nodes.extend(payloads)
        return list(range(first, len(self._nodes)))

    def add_edges_from(self, edges):
        n = len(self._nodes)
        if not edges:
            return []
        e = np.asarray([(t[0], t[1]) for t in edges], dtype=np.int64)
        if e.min() < 0 or e.max() >= n:
            raise IndexError("One of the endpoints of the edge does not exist in graph")
        self._a.append(e[:, 0])
        self._b.append(e[:, 1])
        return list(range(len(edges)))

    def num_nodes(self):
        return len(self._nodes)

    def num_edges(self):
        return int(sum(a.size for a in self._a))


def connected_components(graph):
    n = len(graph._nodes)
    if n == 0:
        return []
    if graph._a:
        a = np.concatenate(graph._a)
        b = np.concatenate(graph._b)
        g = coo_matrix((np.ones(a.size, dtype=np.int8), (a, b)), shape=(n, n))
    else:
        g = coo_matrix((n, n), dtype=np.int8)
    _, labels = _cc(g, directed=False)
    order = np.argsort(labels, kind="stable")
    cuts = np.flatnonzero(np.diff(labels[order])) + 1
    return [set(int(v) for v in grp) for grp in np.split(order, cuts)]
