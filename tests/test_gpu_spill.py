"""K3 (sort-reduce spill path) parity, -m gpu: the emit -> radix sort -> reduce-by-key -> merge
pipeline must give the oracle's rows bit for bit, including rows whose only colours have weight 0
(src/pairwise.cpp:222-225 inserts them), on one device, per row block, and with several chunks."""
import os

import numpy as np
import pytest

from util import assert_records_equal, oracle_records, random_csr, sort_records

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ksb(built_lib):
    from kspider_b200.capi import check
    check(built_lib.ksb_cuda_available())
    import kspider_b200
    return kspider_b200


CASES = [
    ("small_sizes", dict(seed=61, n_samples=200, n_colors=6000, s_lo=1, s_hi=64)),
    ("mid_and_zero", dict(seed=62, n_samples=1500, n_colors=300, s_lo=2, s_hi=20, big=(255, 257, 513, 1025, 1499), zero_every=7)),
    ("sparse_ids_big_w", dict(seed=63, n_samples=400, n_colors=3000, s_lo=1, s_hi=40, id_mode="sparse", w_hi=2 ** 32 - 1, big=(100, 399))),
    ("pow2_samples", dict(seed=64, n_samples=256, n_colors=2000, s_lo=2, s_hi=12)),
]


@pytest.mark.parametrize("name,kw", CASES, ids=[c[0] for c in CASES])
def test_spill_bit_exact_vs_oracle(ksb, name, kw):
    csr = random_csr(**kw)
    idx = ksb.Index.from_csr(**csr, host_threads=3)
    res = idx.pairwise(device=0, spill="always")
    got = res.records()
    assert_records_equal(got, oracle_records(csr))
    assert np.array_equal(got, sort_records(got)), "the final run is in (source_1, source_2) order"
    t = res.info.t
    assert t.spill_emit_ms > 0 and t.spill_sort_ms > 0 and res.info.n_updates == idx.info.n_updates


def test_spill_many_chunks_and_row_blocks(ksb):
    """~4.6 M updates with the minimum chunk capacity (2^20 entries): several chunks and run merges;
    then the same through 3 row blocks; then with the Gram-routed colours emitted by k3_emit_csr."""
    csr = random_csr(seed=65, n_samples=900, n_colors=9000, s_lo=2, s_hi=60, big=(400, 899), zero_every=31)
    want = oracle_records(csr)
    idx = ksb.Index.from_csr(**csr)
    assert idx.info.n_updates > 4 * (1 << 20)
    dev = idx.upload(spill="always", dense_budget_bytes=1)  # capacity clamps to 2^20 entries
    assert dev.is_spill
    res = dev.run_all_passes()
    assert_records_equal(res.records(), want)
    assert res.info.gpu_launches >= 12
    parts = [idx.pairwise(rank=r, world=3, spill="always", dense_budget_bytes=1).records() for r in range(3)]
    assert_records_equal(np.concatenate(parts), want)
    old = os.environ.get("KSPIDER_GRAM")
    os.environ["KSPIDER_GRAM"] = "30"
    try:
        idx2 = ksb.Index.from_csr(**csr)
        assert idx2.info.n_gram_colors > 0
        assert_records_equal(idx2.pairwise(spill="always", dense_budget_bytes=1).records(), want)
    finally:
        if old is None:
            os.environ.pop("KSPIDER_GRAM", None)
        else:
            os.environ["KSPIDER_GRAM"] = old


def test_spill_auto_when_budget_needs_many_passes(ksb):
    csr = random_csr(seed=66, n_samples=700, n_colors=3000, s_lo=1, s_hi=30)
    idx = ksb.Index.from_csr(**csr)
    dev = idx.upload(dense_budget_bytes=60_000)  # 244 650 cells * 8 B / 60 kB > 16 passes -> spill
    assert dev.is_spill
    assert_records_equal(dev.run_all_passes().records(), oracle_records(csr))
    dev2 = idx.upload(dense_budget_bytes=60_000, spill="never")
    assert not dev2.is_spill and dev2.n_passes > 16
    # cluster works on spill results like on dense ones
    res = idx.pairwise(spill="always")
    full, ne = idx.pairwise(spill="never").cluster(int(idx.info.n_samples), "max_cont", 0.0)
    lab, ne2 = res.cluster(int(idx.info.n_samples), "max_cont", 0.0)
    assert ne == ne2 and np.array_equal(full, lab)
