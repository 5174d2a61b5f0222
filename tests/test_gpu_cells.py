"""Width of the accumulator's cells, -m gpu.  A cell is a sum of weights of colours that hold its row, so the largest
per-sample sum of weights (the loader's max_row_sum) bounds every cell: below 2^32 the device keeps u32 cells (half the
HBM traffic of the clear, of K4's reds and of the post-pass scan), otherwise u64.  Both widths must give the
reference's u64 sums (src/pairwise.cpp:222-225) bit for bit, on every kernel family, also right at the bound."""
import os

import numpy as np
import pytest

from util import assert_records_equal, oracle_records, random_csr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ksb(built_lib):
    from kspider_b200.capi import check
    check(built_lib.ksb_cuda_available())
    import kspider_b200
    return kspider_b200


def _bound_csr(total, n_extra=40):
    """Samples 1..3 share colours whose weights add up to `total` exactly (so cell (1,2) = total and sample 1's sum of
    weights = total); a few light colours around them.  Members deliberately unsorted."""
    k = -(-total // (2 ** 32 - 1)) + 2                      # colour weights are u32: enough colours to reach the total
    parts = [total // k] * (k - 1)
    parts.append(total - sum(parts))
    assert all(0 < p < 2 ** 32 for p in parts) and sum(parts) == total
    cols = [[3, 1, 2], [2, 3, 1], [1, 3, 2]] * k
    cols = cols[:k]
    weights = list(parts)
    rng = np.random.default_rng(7)
    for _ in range(n_extra):
        k = int(rng.integers(2, 7))
        cols.append(list(rng.choice(np.arange(4, 60), size=k, replace=False)))
        weights.append(int(rng.integers(1, 1000)))
    offsets = np.zeros(len(cols) + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum([len(c) for c in cols])
    ids = np.arange(1, 60, dtype=np.uint32)
    return dict(offsets=offsets, members=np.concatenate(cols).astype(np.uint32), weights=np.array(weights, dtype=np.uint32),
                total_ids=ids, total_vals=np.full(ids.size, 4_000_000_000, dtype=np.uint32))


@pytest.mark.parametrize("gram", ["off", "2"])
@pytest.mark.parametrize("total,bits", [(2 ** 32 - 1, 32), (2 ** 32, 64), (3 * (2 ** 32 - 1), 64)])
def test_cell_width_follows_the_bound_and_sums_are_exact(ksb, monkeypatch, total, bits, gram):
    monkeypatch.setenv("KSPIDER_GRAM", gram)   # "2": the heavy colours go through K4 (four weight byte planes)
    csr = _bound_csr(total)
    idx = ksb.Index.from_csr(**csr)
    assert idx.upload().cell_bits == 64          # a block this small is L2-resident: u64 cells unless asked
    monkeypatch.setenv("KSPIDER_CELL_BITS", "32")  # u32 cells whenever the bound allows them
    dev = idx.upload()
    assert dev.cell_bits == bits
    rec = dev.run_all_passes().records()
    assert_records_equal(rec, oracle_records(csr))
    top = rec[(rec["source_1"] == 1) & (rec["source_2"] == 2)]
    assert top.size == 1 and int(top["shared_kmers"][0]) == total
    dev.accumulate()
    assert int(dev.read_triangle().max()) == total


@pytest.mark.parametrize("cell_bits", ["32", "64"])
def test_both_cell_widths_on_every_kernel_family(ksb, monkeypatch, cell_bits):
    """K1 (small colours, both cell-index variants), K2 (mid colours), Kz (zero weights), K4 (forced), several dense
    passes, row blocks, the record path (K5) and the cluster-only step (K56) -- with u32 cells (what an index of light
    weights gets) and with u64 cells forced on the same index."""
    from oracle import cluster_oracle
    monkeypatch.setenv("KSPIDER_CELL_BITS", cell_bits)
    a = random_csr(seed=91, n_samples=700, n_colors=5000, s_lo=1, s_hi=64, w_hi=1000, big=(100, 400, 700), zero_every=13)
    want = oracle_records(a)
    n = 700
    for gram in ("off", "20"):
        monkeypatch.setenv("KSPIDER_GRAM", gram)
        idx = ksb.Index.from_csr(**a, host_threads=3)
        dev = idx.upload()
        assert dev.cell_bits == int(cell_bits)
        assert_records_equal(dev.run_all_passes().records(), want)
        dev2 = idx.upload(dense_budget_bytes=200_000)
        assert dev2.n_passes >= 3
        assert_records_equal(dev2.run_all_passes().records(), want)
        parts = [idx.pairwise(rank=r, world=3).records() for r in range(3)]
        assert_records_equal(np.concatenate(parts), want)
        for k1 in ("32", "64"):   # K1's 32-bit and 64-bit cell-index kernels
            monkeypatch.setenv("KSPIDER_K1_CELLS", k1)
            assert_records_equal(idx.pairwise().records(), want)
        monkeypatch.delenv("KSPIDER_K1_CELLS")
        ids = idx.sample_ids
        for dist, col in (("max_cont", "max_containment"), ("min_cont", "min_containment"), ("avg_cont", "avg_containment")):
            cutoff = float(np.quantile(want[col][np.isfinite(want[col])], 0.98))
            keep = cluster_oracle.kept_mask(want[col], cutoff * 100)
            ref = cluster_oracle.min_labels(n, np.searchsorted(ids, want["source_1"][keep]), np.searchsorted(ids, want["source_2"][keep]))
            for d in (dev, dev2):
                labels, ne, nr = d.cluster(n, dist, cutoff)
                assert (ne, nr) == (int(keep.sum()), want.size) and np.array_equal(labels, ref)
