"""ksb_device_cluster -- the cluster-only step (K56 fused scan: containment + threshold + kept pairs only, then
unions over the edge list) -- against the record path (ksb_postpass + ksb_cluster) and the reference-pinned
cluster oracle: every dist type, several cut-offs, zero-weight colours, several dense passes, row blocks,
an edge list that is too short at first."""
import numpy as np
import pytest

from oracle import cluster_oracle
from util import oracle_records, random_csr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ksb(built_lib):
    from kspider_b200.capi import check
    check(built_lib.ksb_cuda_available())
    import kspider_b200
    return kspider_b200


def spread_totals(csr, seed, lo=300, hi=4000):
    csr["total_vals"] = np.random.default_rng(seed).integers(lo, hi, size=csr["total_ids"].size).astype(np.uint32)
    return csr


def oracle_labels(csr, ids, dist, cutoff):
    want = oracle_records(csr)
    col = {"min_cont": "min_containment", "avg_cont": "avg_containment", "max_cont": "max_containment"}[dist]
    keep = cluster_oracle.kept_mask(want[col], float(cutoff) * 100)
    a = np.searchsorted(ids, want["source_1"][keep])
    b = np.searchsorted(ids, want["source_2"][keep])
    return cluster_oracle.min_labels(ids.size, a, b), int(keep.sum()), want.size


CASES = [
    ("plain", dict(seed=41, n_samples=300, n_colors=600, s_lo=2, s_hi=5), {}),
    ("zero_weights_sparse_ids", dict(seed=42, n_samples=200, n_colors=900, s_lo=1, s_hi=7, zero_every=5, id_mode="sparse"), {}),
    ("mid_and_gram", dict(seed=43, n_samples=400, n_colors=300, s_lo=2, s_hi=30, big=(70, 150, 390)), {}),
    ("many_passes", dict(seed=44, n_samples=500, n_colors=2500, s_lo=2, s_hi=6), dict(dense_budget_bytes=120_000)),
]


@pytest.mark.parametrize("name,kw,opts", CASES, ids=[c[0] for c in CASES])
def test_fused_step_equals_record_path_and_oracle(ksb, name, kw, opts):
    csr = spread_totals(random_csr(**kw), kw["seed"])
    idx = ksb.Index.from_csr(**csr)
    n = int(idx.info.n_samples)
    ids = idx.sample_ids
    dev = idx.upload(**opts)
    if name == "many_passes":
        assert dev.n_passes > 3
    res = dev.run_all_passes(want_records=False)
    for dist in ("min_cont", "avg_cont", "max_cont"):
        for cutoff in (0.0, 0.05, 0.2, 0.6, 1.0):
            labels, ne, nr = dev.cluster(n, dist, cutoff)
            rec_labels, rec_ne = res.cluster(n, dist, cutoff)
            want, want_ne, want_rows = oracle_labels(csr, ids, dist, cutoff)
            assert (ne, nr) == (want_ne, want_rows) and ne == rec_ne, (dist, cutoff)
            assert np.array_equal(labels, want) and np.array_equal(labels, rec_labels), (dist, cutoff)
    assert dev.info.gpu_launches > 0


def test_fused_step_grows_a_short_edge_list(ksb, monkeypatch):
    csr = spread_totals(random_csr(seed=45, n_samples=300, n_colors=3000, s_lo=2, s_hi=6), 45)
    idx = ksb.Index.from_csr(**csr)
    monkeypatch.setenv("KSPIDER_EDGE_CAP", "7")
    dev = idx.upload()
    labels, ne, nr = dev.cluster(300, "max_cont", 0.0)   # every touched pair is kept: far more than 7 edges
    want, want_ne, want_rows = oracle_labels(csr, idx.sample_ids, "max_cont", 0.0)
    assert ne == want_ne > 7 and nr == want_rows and np.array_equal(labels, want)
    labels2, ne2, _ = dev.cluster(300, "max_cont", 0.3)  # the grown list is kept for the next step
    want2, want_ne2, _ = oracle_labels(csr, idx.sample_ids, "max_cont", 0.3)
    assert ne2 == want_ne2 and np.array_equal(labels2, want2)


def test_fused_step_on_row_blocks_merges_to_the_whole(ksb):
    csr = spread_totals(random_csr(seed=46, n_samples=400, n_colors=1500, s_lo=2, s_hi=5), 46)
    idx = ksb.Index.from_csr(**csr)
    whole, ne, nr = idx.upload().cluster(400, "max_cont", 0.15)
    parts, edges, rows = [], 0, 0
    for r in range(3):
        lab, e, n = idx.shard(r, 3).upload().cluster(400, "max_cont", 0.15)
        parts.append(lab); edges += e; rows += n
    merged = ksb.cluster_merge(ksb.cluster_merge(parts[0], parts[1]), parts[2])
    assert (edges, rows) == (ne, nr) and np.array_equal(merged, whole)
    want, _, _ = oracle_labels(csr, idx.sample_ids, "max_cont", 0.15)
    assert np.array_equal(whole, want)


def test_spill_device_refuses(ksb):
    from kspider_b200.capi import KsbError
    csr = random_csr(seed=47, n_samples=100, n_colors=300, s_lo=2, s_hi=5)
    dev = ksb.Index.from_csr(**csr).upload(spill="always", dense_budget_bytes=50_000_000)
    with pytest.raises(KsbError):
        dev.cluster(100, "max_cont", 0.0)


def test_graph_replay_equals_kernel_by_kernel(ksb):
    """the step as one CUDA graph launch (captured once per column / threshold, replayed afterwards) against the
    same step enqueued kernel by kernel; a change of threshold re-captures"""
    csr = spread_totals(random_csr(seed=48, n_samples=400, n_colors=1500, s_lo=2, s_hi=30, big=(70, 150, 390)), 48)
    idx = ksb.Index.from_csr(**csr)
    dev = idx.upload()
    for cutoff in (0.1, 0.4, 0.1):
        dev.set_graph(False)
        want = dev.cluster(400, "max_cont", cutoff)
        dev.set_graph(True)
        launches0 = dev.info.gpu_launches
        for rep in range(3):   # capture, then two replays
            got = dev.cluster(400, "max_cont", cutoff)
            assert np.array_equal(got[0], want[0]) and got[1:] == want[1:], (cutoff, rep)
        assert dev.info.gpu_launches - launches0 >= 3 * 5   # replays count their kernels
    lab, ne, nr = dev.cluster(400, "min_cont", 0.1)
    ref, ref_ne, ref_rows = oracle_labels(csr, idx.sample_ids, "min_cont", 0.1)
    assert (ne, nr) == (ref_ne, ref_rows) and np.array_equal(lab, ref)


def test_end_to_end_call_reuses_its_device_handle(ksb, monkeypatch):
    """ksb_pairwise_cluster keeps the device handle of its last call in the index and only uploads the blob again: calls
    in a row with other thresholds and columns, other options (which replace the kept handle), a dense budget that needs
    several passes, row blocks, and a different index in between all give the oracle's labels; so does the call with the
    cache switched off."""
    csr = spread_totals(random_csr(seed=61, n_samples=380, n_colors=2200, s_lo=2, s_hi=9, zero_every=19, big=(80, 200, 370)), 61)
    other = spread_totals(random_csr(seed=62, n_samples=150, n_colors=500, s_lo=2, s_hi=6), 62)
    idx = ksb.Index.from_csr(**csr)
    idx2 = ksb.Index.from_csr(**other)
    n, ids = int(idx.info.n_samples), idx.sample_ids

    def call(ix, c, dist, cutoff, **kw):
        labels, ne, nr, info = ix.pairwise_cluster(dist, cutoff, **kw)
        want, want_ne, want_rows = oracle_labels(c, ix.sample_ids, dist, cutoff)
        assert (ne, nr) == (want_ne, want_rows) and np.array_equal(labels, want), (dist, cutoff, kw)
        return info

    first = call(idx, csr, "max_cont", 0.05)
    again = call(idx, csr, "max_cont", 0.05)
    assert again.t.host_setup_ms < first.t.host_setup_ms       # nothing is planned or allocated the second time
    for dist in ("min_cont", "avg_cont", "max_cont"):
        for cutoff in (0.0, 0.2, 1.0):
            call(idx, csr, dist, cutoff)
    call(idx2, other, "max_cont", 0.1)                          # another index has its own handle
    call(idx, csr, "max_cont", 0.3, dense_budget_bytes=120_000)  # other options: the kept handle is replaced
    call(idx, csr, "max_cont", 0.3, dense_budget_bytes=120_000)
    call(idx, csr, "min_cont", 0.1)
    # row blocks: each call returns the labels of its own rows' edges; merged they are the whole partition
    parts = [idx.pairwise_cluster("max_cont", 0.05, rank=r, world=2)[0] for r in range(2)]
    merged = ksb.cluster_merge(parts[0], parts[1]) if hasattr(ksb, "cluster_merge") else None
    if merged is not None:
        assert np.array_equal(merged, oracle_labels(csr, ids, "max_cont", 0.05)[0])
    monkeypatch.setenv("KSPIDER_E2E_CACHE", "0")                # read once per process: a fresh index still answers the same
    idx3 = ksb.Index.from_csr(**csr)
    call(idx3, csr, "max_cont", 0.05)
    idx.close(); idx2.close(); idx3.close()
