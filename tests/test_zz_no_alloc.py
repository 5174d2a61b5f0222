"""ksb_cluster_edges_noalloc -- the allocation-free variant of ksb_cluster_table's device step (no counterpart in the
reference; static __device__ scratch, include/kspider_b200.h) -- against ksb_cluster_table itself, through the C-ABI,
on two calls in a row that share the scratch.  Collected last on purpose."""
import ctypes as C
import os

import numpy as np
import pytest

from kspider_b200.api import RECORD_DTYPE
from kspider_b200.capi import KSB_ERR_ARG, check

pytestmark = pytest.mark.gpu

MAX_NODES = 1 << 20  # KSB_NOALLOC_MAX_NODES


def edge_table(seed, n_nodes, n_rows, clique=0):
    """Random rows (ids 1-based, source_1 < source_2) whose containments straddle the cut-off; optionally a clique so
    that hook-min and the unions meet on the same nodes."""
    rng = np.random.default_rng(seed)
    a = rng.integers(1, n_nodes, n_rows, dtype=np.uint32)
    b = a + rng.integers(1, 40, n_rows).astype(np.uint32)
    b = np.minimum(b, np.uint32(n_nodes))
    keep = a < b
    a, b = a[keep], b[keep]
    if clique:
        i, j = np.triu_indices(clique, 1)
        a = np.concatenate([a, (i + 1).astype(np.uint32)])
        b = np.concatenate([b, (j + 1).astype(np.uint32)])
    cont = rng.random(a.size, dtype=np.float32)
    cont[rng.integers(0, a.size, 5)] = np.float32("nan")  # NaN rows are kept by the reference's test
    return a, b, cont


def via_cluster_table(lib, tmp_path, tag, a, b, cont, n_nodes, cut_pct):
    """The existing entry point: side-car -> ksb_table -> ksb_cluster_table."""
    rec = np.zeros(a.size, dtype=RECORD_DTYPE)
    rec["source_1"], rec["source_2"], rec["max_containment"] = a, b, cont
    side = str(tmp_path / f"{tag}.ksb")
    check(lib.ksb_write_sidecar_records(rec.ctypes.data, rec.size, side.encode(), 0, 0, 0))
    h = C.c_void_p()
    check(lib.ksb_table_read_sidecar(side.encode(), 5, 0, 2, C.byref(h)))
    labels = np.empty(n_nodes, dtype=np.uint32)
    ne = C.c_uint64()
    try:
        check(lib.ksb_cluster_table(h, n_nodes, cut_pct, 0, labels.ctypes.data, C.byref(ne)))
    finally:
        lib.ksb_table_free(h)
    return labels, ne.value


def via_noalloc(lib, torch, a, b, cont, n_nodes, cut_pct):
    dev = torch.device("cuda:0")
    # every buffer the call touches exists BEFORE it; the call itself must not allocate
    d_a = torch.from_numpy(a.view(np.int32)).to(dev)
    d_b = torch.from_numpy(b.view(np.int32)).to(dev)
    d_c = torch.from_numpy(cont).to(dev)
    d_labels = torch.full((n_nodes,), -1, dtype=torch.int32, device=dev)
    d_cnt = torch.full((2,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    st = torch.cuda.current_stream().cuda_stream
    before = torch.cuda.mem_get_info()[0]
    check(lib.ksb_cluster_edges_noalloc(d_a.data_ptr(), d_b.data_ptr(), d_c.data_ptr(), a.size, n_nodes, cut_pct,
                                        C.c_void_p(st), d_labels.data_ptr(), d_cnt.data_ptr()))
    torch.cuda.synchronize()
    after = torch.cuda.mem_get_info()[0]
    cnt = d_cnt.cpu().numpy().view(np.uint64)
    return d_labels.cpu().numpy().view(np.uint32), int(cnt[0]), int(cnt[1]), before - after


def test_noalloc_matches_cluster_table_on_two_calls_in_a_row(built_lib, tmp_path):
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    lib = built_lib
    # set-up call (resolves the scratch address on this device): no rows, every node its own component
    d_l = torch.empty(8, dtype=torch.int32, device="cuda:0")
    check(lib.ksb_cluster_edges_noalloc(None, None, None, 0, 8, 50.0, None, d_l.data_ptr(), None))
    torch.cuda.synchronize()
    assert d_l.cpu().tolist() == list(range(8))
    # two calls in a row on different inputs: the second is LARGER in nodes and smaller in edges than the first, and
    # the first call again afterwards, so stale parents or counters of either would show
    cases = [
        dict(seed=11, n_nodes=50_000, n_rows=400_000, clique=300, cut_pct=97.0),
        dict(seed=12, n_nodes=200_000, n_rows=150_000, clique=0, cut_pct=35.5),
        dict(seed=11, n_nodes=50_000, n_rows=400_000, clique=300, cut_pct=97.0),
    ]
    for k, cs in enumerate(cases):
        a, b, cont = edge_table(cs["seed"], cs["n_nodes"], cs["n_rows"], cs["clique"])
        want, want_ne = via_cluster_table(lib, tmp_path, f"t{k}", a, b, cont, cs["n_nodes"], cs["cut_pct"])
        got, ne, bad, grew = via_noalloc(lib, torch, a, b, cont, cs["n_nodes"], cs["cut_pct"])
        assert bad == 0xFFFFFFFF
        assert ne == want_ne and ne > 0
        assert np.array_equal(got, want)  # integer labels: bit-exact, the tolerance of every cluster test here
        assert 1 < np.unique(got).size < cs["n_nodes"]
        assert grew <= 0, f"device memory in use grew by {grew} bytes across the call"


def test_noalloc_limits_and_foreign_nodes(built_lib):
    torch = pytest.importorskip("torch")
    torch.cuda.set_device(0)
    lib = built_lib
    d = torch.zeros(4, dtype=torch.int32, device="cuda:0")
    rc = lib.ksb_cluster_edges_noalloc(None, None, None, 0, MAX_NODES + 1, 0.0, None, d.data_ptr(), None)
    assert rc == KSB_ERR_ARG and b"KSB_NOALLOC_MAX_NODES" in lib.ksb_last_error()
    # a kept row naming a node outside the graph is skipped and reported in counters[1]
    a = np.array([1, 2, 7], dtype=np.uint32)
    b = np.array([2, 3, 9], dtype=np.uint32)
    c = np.array([0.9, 0.1, 0.9], dtype=np.float32)
    d_a, d_b, d_c = (torch.from_numpy(x.view(np.int32) if x.dtype == np.uint32 else x).cuda() for x in (a, b, c))
    d_l = torch.empty(4, dtype=torch.int32, device="cuda:0")
    d_n = torch.empty(2, dtype=torch.int64, device="cuda:0")
    check(lib.ksb_cluster_edges_noalloc(d_a.data_ptr(), d_b.data_ptr(), d_c.data_ptr(), 3, 4, 50.0, None, d_l.data_ptr(), d_n.data_ptr()))
    torch.cuda.synchronize()
    assert d_l.cpu().tolist() == [0, 0, 2, 3]
    assert d_n.cpu().tolist()[1] == 2
