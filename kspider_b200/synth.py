"""Seeded synthetic indexes of the BASELINE.json shapes, as CSR arrays and,
on request, as the reference's own index files (real phmap dump bytes via ksb_write_index_files).

Sample ids are 1..N like the reference's indexers produce; every index also carries the N
singleton colours (colour id = sample id) when written to files.  Multi-member colour ids start
at N+1.  Weights are >= 1 unless `zero_weight_every` is set.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict

import numpy as np


@dataclass
class SynthIndex:
    name: str
    n_samples: int
    offsets: np.ndarray      # u64[C+1]
    members: np.ndarray      # u32[M], ORIGINAL ids (1..N), unsorted within a colour is allowed
    weights: np.ndarray      # u32[C]
    total_ids: np.ndarray    # u32[N]
    total_vals: np.ndarray   # u32[N]
    meta: Dict = field(default_factory=dict)

    @property
    def n_colors(self) -> int:
        return int(self.weights.size)

    @property
    def n_updates(self) -> int:
        s = np.diff(self.offsets).astype(np.int64)
        return int((s * (s - 1) // 2).sum())


def _distinct_rows(rng: np.random.Generator, n_rows: int, s: int, n: int) -> np.ndarray:
    """[n_rows, s] sorted rows of distinct values in [0, n): draw, then redraw duplicates until none."""
    a = rng.integers(0, n, size=(n_rows, s), dtype=np.int32)
    a.sort(axis=1)
    for _ in range(200):
        dup = np.zeros(a.shape, dtype=bool)
        dup[:, 1:] = a[:, 1:] == a[:, :-1]
        k = int(dup.sum())
        if k == 0:
            return a
        rows = np.nonzero(dup.any(axis=1))[0]
        sub = a[rows]
        d = dup[rows]
        sub[d] = rng.integers(0, n, size=int(d.sum()), dtype=np.int32)
        sub.sort(axis=1)
        a[rows] = sub
    raise RuntimeError("could not de-duplicate rows (s too close to n?)")


def zipf_sizes(n_colors: int, s_min: int = 2, s_max: int = 64, alpha: float = 1.5) -> np.ndarray:
    """Deterministic class counts: n_colors split over s in [s_min, s_max] with P(s) ~ s^-alpha."""
    s = np.arange(s_min, s_max + 1, dtype=np.float64)
    p = s ** (-alpha)
    p /= p.sum()
    cnt = np.floor(p * n_colors).astype(np.int64)
    cnt[0] += n_colors - cnt.sum()
    return cnt


def zipf_index(n_samples: int = 1000, n_colors: int = 10_000_000, s_max: int = 64, alpha: float = 1.5,
               seed: int = 12345, w_max: int = 100, singleton_count: int = 1000, name: str = "cfg2") -> SynthIndex:
    """cfg 2: "kProcessor-style" colour table, Zipf colour sizes in [2, s_max], members uniform
    without replacement, w ~ U{1..w_max}, total[i] = sum of w over colours containing i + singleton_count."""
    rng = np.random.default_rng(seed)
    cnt = zipf_sizes(n_colors, 2, s_max, alpha)
    sizes = np.repeat(np.arange(2, s_max + 1, dtype=np.int64), cnt)
    perm = rng.permutation(n_colors)
    # members class by class, then scattered into a random colour order (files have no size order)
    offsets = np.zeros(n_colors + 1, dtype=np.uint64)
    sizes_perm = np.empty(n_colors, dtype=np.int64)
    sizes_perm[perm] = sizes
    np.cumsum(sizes_perm, out=offsets[1:].view(np.int64))
    members = np.empty(int(offsets[-1]), dtype=np.uint32)
    weights = rng.integers(1, w_max + 1, size=n_colors, dtype=np.int64).astype(np.uint32)
    totals = np.full(n_samples, singleton_count, dtype=np.int64)
    at = 0
    for si, s in enumerate(range(2, s_max + 1)):
        n = int(cnt[si])
        if n == 0:
            continue
        rows = _distinct_rows(rng, n, s, n_samples)
        cols = perm[at:at + n]
        pos = offsets[cols].astype(np.int64)[:, None] + np.arange(s, dtype=np.int64)[None, :]
        members[pos.ravel()] = (rows + 1).astype(np.uint32).ravel()
        totals += np.bincount(rows.ravel(), weights=np.repeat(weights[cols].astype(np.float64), s),
                              minlength=n_samples).astype(np.int64)
        at += n
    return SynthIndex(name, n_samples, offsets, members, weights,
                      np.arange(1, n_samples + 1, dtype=np.uint32), (totals & 0xFFFFFFFF).astype(np.uint32),
                      {"seed": seed, "alpha": alpha, "s_max": s_max})


def _colors_from_incidence(inc: np.ndarray, first_id: int):
    """inc: bool [hashes, genomes].  Distinct non-trivial member sets -> (sizes, members, weights)."""
    packed = np.packbits(inc, axis=1)
    uniq, counts = np.unique(packed, axis=0, return_counts=True)
    bits = np.unpackbits(uniq, axis=1)[:, :inc.shape[1]].astype(bool)
    sizes = bits.sum(axis=1)
    keep = sizes >= 2
    bits, counts, sizes = bits[keep], counts[keep], sizes[keep]
    _, g = np.nonzero(bits)
    return sizes.astype(np.int64), (g + first_id).astype(np.uint32), counts.astype(np.uint32)


def fracminhash_index(n_clades: int = 200, clade_size: int = 50, sketch: int = 5000, core: int = 3000,
                      core_p: float = 0.9, family_clades: int = 20, family_pool: int = 2000, family_p: float = 0.5,
                      seed: int = 31, name: str = "cfg3") -> SynthIndex:
    """cfg 3 / cfg 4: genomes in clades; a sketch = clade-core hashes (each kept w.p. core_p) +
    hashes of a `family_clades`-clade pool (w.p. family_p) + private hashes up to `sketch`.
    Colour = distinct member set of a hash, w = number of hashes with that set."""
    rng = np.random.default_rng(seed)
    n = n_clades * clade_size
    sizes_l, members_l, weights_l = [], [], []
    per_sample = np.zeros(n, dtype=np.int64)
    for c in range(n_clades):
        inc = rng.random((core, clade_size)) < core_p
        per_sample[c * clade_size:(c + 1) * clade_size] += inc.sum(axis=0)
        s, m, w = _colors_from_incidence(inc, c * clade_size + 1)
        sizes_l.append(s); members_l.append(m); weights_l.append(w)
    fam_genomes = family_clades * clade_size
    for f in range((n_clades + family_clades - 1) // family_clades):
        g0 = f * fam_genomes
        g1 = min(n, g0 + fam_genomes)
        inc = rng.random((family_pool, g1 - g0)) < family_p
        per_sample[g0:g1] += inc.sum(axis=0)
        s, m, w = _colors_from_incidence(inc, g0 + 1)
        sizes_l.append(s); members_l.append(m); weights_l.append(w)
    sizes = np.concatenate(sizes_l)
    offsets = np.zeros(sizes.size + 1, dtype=np.uint64)
    np.cumsum(sizes, out=offsets[1:].view(np.int64))
    totals = np.maximum(per_sample, sketch)  # private hashes fill the sketch up
    return SynthIndex(name, n, offsets, np.concatenate(members_l), np.concatenate(weights_l),
                      np.arange(1, n + 1, dtype=np.uint32), totals.astype(np.uint32),
                      {"seed": seed, "n_clades": n_clades, "clade_size": clade_size, "sketch": sketch})


def cfg3(seed: int = 31) -> SynthIndex:
    return fracminhash_index(seed=seed, name="cfg3")


def cfg4(seed: int = 100) -> SynthIndex:
    return fracminhash_index(n_clades=2000, clade_size=50, sketch=1000, core=600, family_pool=400, seed=seed,
                             name="cfg4")


def giant_index(n_samples: int = 10_000, n_background: int = 1_000_000, n_giant: int = 16, giant_size: int = 4000,
                seed: int = 5, name: str = "cfg5") -> SynthIndex:
    """cfg 5: Zipf background + n_giant colours of giant_size members, w ~ U{1..2^32-1}."""
    bg = zipf_index(n_samples, n_background, seed=seed, name=name)
    rng = np.random.default_rng(seed + 1)
    rows = np.stack([np.sort(rng.choice(n_samples, size=giant_size, replace=False)) for _ in range(n_giant)]) \
        if n_giant else np.empty((0, giant_size), dtype=np.int64)
    gw = rng.integers(1, 2 ** 32, size=n_giant, dtype=np.int64).astype(np.uint32)
    offsets = np.concatenate([bg.offsets, bg.offsets[-1] + np.uint64(giant_size) * np.arange(1, n_giant + 1, dtype=np.uint64)])
    members = np.concatenate([bg.members, (rows + 1).astype(np.uint32).ravel()])
    weights = np.concatenate([bg.weights, gw])
    totals = bg.total_vals.astype(np.int64)
    for r, w in zip(rows, gw):
        totals[r] += int(w)
    return SynthIndex(name, n_samples, offsets, members, weights, bg.total_ids,
                      (totals & 0xFFFFFFFF).astype(np.uint32), {"seed": seed, "n_giant": n_giant, "giant_size": giant_size})


def write_index_files(ix: SynthIndex, prefix: str, with_names: bool = True) -> None:
    """The three .bin files (+ .namesMap) the way an indexer leaves them: N singleton colours
    (id = sample id, count = 0 after all its hashes moved on... any value; pairwise ignores them)
    followed by the multi-member colours with ids N+1.."""
    from .api import write_index_files as _w
    n, c = ix.n_samples, ix.n_colors
    color_ids = np.concatenate([np.arange(1, n + 1, dtype=np.uint64), np.arange(n + 1, n + 1 + c, dtype=np.uint64)])
    offsets = np.concatenate([np.arange(0, n + 1, dtype=np.uint64), ix.offsets[1:] + np.uint64(n)])
    members = np.concatenate([np.arange(1, n + 1, dtype=np.uint32), ix.members])
    count_vals = np.concatenate([np.ones(n, dtype=np.uint64), ix.weights.astype(np.uint64)])
    _w(prefix, color_ids, offsets, members, color_ids, count_vals, ix.total_ids, ix.total_vals)
    if with_names:
        with open(prefix + ".namesMap", "w") as f:
            f.write(f"{n}\n")
            for i in range(1, n + 1):
                f.write(f"{i} sample_{i}\n")
