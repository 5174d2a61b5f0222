"""Shared helpers of the test-suite: seeded random indexes and result comparison."""
from __future__ import annotations

import numpy as np


def random_csr(seed: int, n_samples: int, n_colors: int, s_lo: int = 1, s_hi: int = 12, id_mode: str = "dense",
               w_hi: int = 1000, zero_every: int = 0, big: tuple = ()):
    """Random colour table.  id_mode: "dense" ids 1..N, "sparse" arbitrary u32 ids.
    big = extra colour sizes appended (e.g. (70, 300, 1500)).  Members within a colour are shuffled
    (files store them in hash-slot order).  Returns dict with offsets/members/weights/total_ids/total_vals."""
    rng = np.random.default_rng(seed)
    if id_mode == "dense":
        ids = np.arange(1, n_samples + 1, dtype=np.uint32)
    else:
        ids = np.sort(rng.choice(np.arange(1, 2 ** 32 - 1, 65537, dtype=np.uint64), size=n_samples, replace=False)).astype(np.uint32)
    sizes = list(rng.integers(s_lo, s_hi + 1, size=n_colors)) + [min(b, n_samples) for b in big]
    sizes = [int(min(s, n_samples)) for s in sizes]
    offsets = np.zeros(len(sizes) + 1, dtype=np.uint64)
    members = []
    for c, s in enumerate(sizes):
        m = rng.choice(n_samples, size=s, replace=False)
        members.append(ids[m])
        offsets[c + 1] = offsets[c] + np.uint64(s)
    members = np.concatenate(members).astype(np.uint32) if members else np.empty(0, np.uint32)
    weights = rng.integers(1, w_hi + 1, size=len(sizes), dtype=np.int64).astype(np.uint32)
    if zero_every:
        weights[::zero_every] = 0
    totals = rng.integers(1, 10 ** 6, size=n_samples, dtype=np.int64).astype(np.uint32)
    return dict(offsets=offsets, members=members, weights=weights, total_ids=ids.copy(), total_vals=totals)


def oracle_records(csr):
    """Oracle result as a structured array sorted by (source_1, source_2) — the oracle sorts."""
    from oracle import pairwise_oracle as po
    from kspider_b200.api import RECORD_DTYPE
    r = po.pairwise_csr(csr["offsets"], csr["members"], csr["weights"], csr["total_ids"], csr["total_vals"])
    out = np.zeros(r.src1.size, dtype=RECORD_DTYPE)
    out["source_1"], out["source_2"], out["shared_kmers"] = r.src1, r.src2, r.shared
    out["min_containment"], out["avg_containment"], out["max_containment"] = r.cmin, r.cavg, r.cmax
    return out


def sort_records(rec: np.ndarray) -> np.ndarray:
    order = np.lexsort((rec["source_2"], rec["source_1"]))
    return rec[order]


def assert_records_equal(got: np.ndarray, want: np.ndarray) -> None:
    """Bit-exact: integers equal, float32 containments compared by bit pattern."""
    got, want = sort_records(got), sort_records(want)
    assert got.size == want.size, f"row count {got.size} != {want.size}"
    for f in ("source_1", "source_2", "shared_kmers"):
        bad = np.nonzero(got[f] != want[f])[0]
        assert bad.size == 0, f"{f} differs at row {bad[:5]}: {got[f][bad[:5]]} vs {want[f][bad[:5]]}"
    for f in ("min_containment", "avg_containment", "max_containment"):
        g, w = got[f].view(np.uint32), want[f].view(np.uint32)
        bad = np.nonzero(g != w)[0]
        assert bad.size == 0, f"{f} bits differ at row {bad[:5]}: {got[f][bad[:5]]} vs {want[f][bad[:5]]}"


def canonical_partition(labels: np.ndarray):
    groups = {}
    for i, l in enumerate(labels.tolist()):
        groups.setdefault(l, []).append(i)
    return sorted(tuple(g) for g in groups.values())
