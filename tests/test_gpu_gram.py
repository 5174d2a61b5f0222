"""K4 (int8 tcgen05 Gram path) parity, -m gpu.  The routing rule is forced with KSPIDER_GRAM=<min size>
so that the same seeded indexes the atomic kernels are checked on go through the tensor-core path;
results must stay bit-exact against the oracle (integers, float32 bit patterns)."""
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


@pytest.fixture
def gram_env():
    old = os.environ.get("KSPIDER_GRAM")

    def set_(v):
        os.environ["KSPIDER_GRAM"] = str(v)
    yield set_
    if old is None:
        os.environ.pop("KSPIDER_GRAM", None)
    else:
        os.environ["KSPIDER_GRAM"] = old


CASES = [
    # name, gram min size, random_csr kwargs
    ("tiny_one_tile", 2, dict(seed=41, n_samples=40, n_colors=300, s_lo=1, s_hi=12, w_hi=200)),
    ("exact_128", 2, dict(seed=42, n_samples=128, n_colors=500, s_lo=2, s_hi=30, w_hi=255)),
    ("three_tiles_1plane", 2, dict(seed=43, n_samples=300, n_colors=4000, s_lo=1, s_hi=40, w_hi=255, big=(70, 150, 299))),
    ("two_planes", 2, dict(seed=44, n_samples=200, n_colors=2000, s_lo=2, s_hi=20, w_hi=65535)),
    ("four_planes", 2, dict(seed=45, n_samples=260, n_colors=1500, s_lo=2, s_hi=64, w_hi=2 ** 32 - 1, big=(200, 260))),
    ("mixed_routing", 20, dict(seed=46, n_samples=700, n_colors=5000, s_lo=1, s_hi=64, w_hi=1000, big=(100, 400, 700), zero_every=13)),
    ("sparse_ids", 2, dict(seed=47, n_samples=400, n_colors=1000, s_lo=1, s_hi=40, id_mode="sparse", big=(100, 399))),
    ("many_kblocks", 2, dict(seed=48, n_samples=150, n_colors=40000, s_lo=2, s_hi=6, w_hi=100)),
    ("whole_population", 2, dict(seed=49, n_samples=1100, n_colors=30, s_lo=2, s_hi=10, w_hi=2 ** 32 - 1, big=(1100, 1100, 1099))),
]


@pytest.mark.parametrize("name,gmin,kw", CASES, ids=[c[0] for c in CASES])
def test_gram_bit_exact_vs_oracle(ksb, gram_env, name, gmin, kw):
    gram_env(gmin)
    csr = random_csr(**kw)
    idx = ksb.Index.from_csr(**csr, host_threads=3)
    inf = idx.info
    assert inf.gram_min_size >= gmin and inf.n_gram_colors > 0
    res = idx.pairwise(device=0)
    got = res.records()
    want = oracle_records(csr)
    assert_records_equal(got, want)
    assert np.array_equal(got, sort_records(got))
    assert res.info.t.gram_ms > 0.0 and res.info.n_updates == inf.n_updates


def test_gram_off_and_auto(ksb, gram_env):
    csr = random_csr(seed=50, n_samples=300, n_colors=2000, s_lo=2, s_hi=40, big=(250,))
    gram_env("off")
    assert ksb.Index.from_csr(**csr).info.n_gram_colors == 0
    gram_env("auto")
    idx = ksb.Index.from_csr(**csr)
    # whatever auto decides, the result is the oracle's
    assert_records_equal(idx.pairwise().records(), oracle_records(csr))


@pytest.mark.parametrize("world", [2, 3])
def test_gram_row_blocks_and_passes(ksb, gram_env, world):
    gram_env(8)
    csr = random_csr(seed=51, n_samples=500, n_colors=3000, s_lo=1, s_hi=50, big=(120, 300, 499), zero_every=17)
    idx = ksb.Index.from_csr(**csr)
    assert idx.info.n_gram_colors > 0
    want = oracle_records(csr)
    parts = [idx.pairwise(rank=r, world=world).records() for r in range(world)]
    assert_records_equal(np.concatenate(parts), want)
    dev = idx.upload(dense_budget_bytes=200_000)
    assert dev.n_passes >= 4
    assert_records_equal(dev.run_all_passes().records(), want)


def test_gram_giant_colours_checksum(ksb, gram_env):
    """BASELINE config 5 shape, scaled: giant dense colours with full-range u32 weights on N = 3000;
    the oracle is too slow here, so: per-sample row sums and the checksum of checksums, and equality
    with the atomic path (KSPIDER_GRAM=off) on the same index."""
    rng = np.random.default_rng(52)
    n, g, s = 3000, 40, 1200
    members = np.concatenate([np.sort(rng.choice(n, size=s, replace=False)) + 1 for _ in range(g)]).astype(np.uint32)
    offsets = (np.arange(g + 1) * s).astype(np.uint64)
    weights = rng.integers(1, 2 ** 32 - 1, size=g, dtype=np.int64).astype(np.uint32)
    ids = np.arange(1, n + 1, dtype=np.uint32)
    totals = rng.integers(1, 10 ** 6, size=n, dtype=np.int64).astype(np.uint32)
    gram_env("auto")
    idx = ksb.Index.from_csr(offsets, members, weights, ids, totals)
    assert idx.info.n_gram_colors == g, "auto routing must pick the giant colours"
    rec = idx.pairwise().records()
    w = weights.astype(object)
    assert sum(int(v) for v in rec["shared_kmers"]) == sum(int(x) * (s * (s - 1) // 2) for x in w)
    gram_env("off")
    ref = ksb.Index.from_csr(offsets, members, weights, ids, totals).pairwise().records()
    assert_records_equal(rec, ref)


def _clade_csr(seed, n_clades, clade, fam, n_core, n_fam, n_noise, w_hi=200, n_samples=None):
    """Clade-structured index (the shape of BASELINE configs 3/4): colours local to a clade, colours
    local to a family of `fam` clades, and a few colours spread over the whole population."""
    rng = np.random.default_rng(seed)
    n = n_samples or n_clades * clade
    cols = []
    for c in range(n_clades):
        lo = c * clade
        for _ in range(n_core):
            k = int(rng.integers(2, clade + 1))
            cols.append(np.sort(rng.choice(clade, size=k, replace=False)) + lo)
    for f in range(0, n_clades, fam):
        lo, width = f * clade, min(fam, n_clades - f) * clade
        for _ in range(n_fam):
            k = int(rng.integers(width // 4, width // 2 + 1))
            cols.append(np.sort(rng.choice(width, size=k, replace=False)) + lo)
    for _ in range(n_noise):
        k = int(rng.integers(2, 40))
        cols.append(np.sort(rng.choice(n, size=k, replace=False)))
    offsets = np.zeros(len(cols) + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum([len(c) for c in cols])
    members = (np.concatenate(cols) + 1).astype(np.uint32)
    weights = rng.integers(0 if seed % 2 else 1, w_hi, size=len(cols), dtype=np.int64).astype(np.uint32)
    ids = np.arange(1, n + 1, dtype=np.uint32)
    totals = rng.integers(1, 10 ** 6, size=n, dtype=np.int64).astype(np.uint32)
    return dict(offsets=offsets, members=members, weights=weights, total_ids=ids, total_vals=totals)


@pytest.fixture(params=["csr", "bitmap"])
def gram_form(request, monkeypatch):
    """Both forms of the Gram section of the blob: member lists that kb_build turns into bitmaps every step, and the
    ready-made bitmaps (IndexDesc::gram_form; the loader takes the smaller one when nothing is forced)."""
    monkeypatch.setenv("KSPIDER_GRAM_FORM", request.param)
    return request.param


@pytest.mark.parametrize("mode", ["auto", 2, 16])
def test_gram_windows_bit_exact_vs_oracle(ksb, gram_env, gram_form, mode):
    """Colours local to clades / families go to small Gram windows (several levels of the window
    grid, windows overlapping by half); the result is the oracle's, bit for bit."""
    gram_env(mode)
    csr = _clade_csr(seed=60, n_clades=12, clade=50, fam=4, n_core=120, n_fam=30, n_noise=200)
    idx = ksb.Index.from_csr(**csr, host_threads=3)
    inf = idx.info
    if mode != "auto":  # auto may decide that an index this small is not worth the path's fixed costs
        assert inf.n_gram_colors > 0 and inf.n_gram_windows > 3
    got = idx.pairwise(device=0).records()
    assert_records_equal(got, oracle_records(csr))


@pytest.mark.parametrize("world", [2, 5])
def test_gram_windows_row_blocks_passes_and_chunks(ksb, gram_env, gram_form, world, monkeypatch):
    gram_env(2)
    monkeypatch.setenv("KSPIDER_GRAM_CHUNK_MB", "1")  # many chunks, windows cut into sub-windows
    csr = _clade_csr(seed=61, n_clades=9, clade=70, fam=3, n_core=3000, n_fam=40, n_noise=50, w_hi=70000)
    idx = ksb.Index.from_csr(**csr)
    assert idx.info.n_gram_windows > 3 and idx.info.gram_planes == 4
    want = oracle_records(csr)
    parts = [idx.pairwise(rank=r, world=world).records() for r in range(world)]
    assert_records_equal(np.concatenate(parts), want)
    dev = idx.upload(dense_budget_bytes=150_000)
    assert dev.n_passes >= 4
    assert_records_equal(dev.run_all_passes().records(), want)
    monkeypatch.setenv("KSPIDER_OVERLAP", "0")
    assert_records_equal(idx.pairwise().records(), want)


def test_gram_windows_beyond_max_width(ksb, gram_env):
    """N > 32768: no whole-population window exists, local colours still use K4, the rest K1/K2."""
    gram_env(16)
    csr = _clade_csr(seed=62, n_clades=8, clade=60, fam=4, n_core=400, n_fam=20, n_noise=300, n_samples=40000)
    idx = ksb.Index.from_csr(**csr)
    inf = idx.info
    assert inf.n_gram_colors > 0 and inf.n_small_colors > 0
    rec = idx.pairwise().records()
    gram_env("off")
    ref = ksb.Index.from_csr(**csr).pairwise().records()
    assert_records_equal(rec, ref)
    sizes = np.diff(csr["offsets"]).astype(object)
    w = csr["weights"].astype(object)
    assert sum(int(v) for v in rec["shared_kmers"]) == sum(int(x) * int(k * (k - 1) // 2) for x, k in zip(w, sizes))


def test_gram_pipelined_upload_pieces(ksb, gram_env, gram_form, monkeypatch):
    """ksb_pairwise copies the Gram members in one piece per K4 chunk (pieces under 4 MB ride with the
    next one) and every chunk's bitmap build waits for its own piece: several chunks and several pieces
    here, result bit-exact; the resident path (upload + run) must agree with it."""
    gram_env(16)
    monkeypatch.setenv("KSPIDER_GRAM_CHUNK_MB", "8")
    rng = np.random.default_rng(71)
    n, s, nc = 1201, 24, 250_000          # n prime: (a + k b) mod n are s distinct members
    a = rng.integers(0, n, size=nc)
    b = rng.integers(1, n, size=nc)
    members = ((a[:, None] + np.arange(s)[None, :] * b[:, None]) % n + 1).astype(np.uint32).ravel()
    small = random_csr(seed=72, n_samples=n, n_colors=20000, s_lo=1, s_hi=12, w_hi=255)
    csr = dict(offsets=np.concatenate([np.arange(nc, dtype=np.uint64) * np.uint64(s), np.uint64(nc * s) + small["offsets"]]),
               members=np.concatenate([members, small["members"]]),
               weights=np.concatenate([rng.integers(1, 256, size=nc).astype(np.uint32), small["weights"]]),
               total_ids=small["total_ids"], total_vals=small["total_vals"])
    idx = ksb.Index.from_csr(**csr, host_threads=4)
    inf = idx.info
    assert inf.n_gram_colors >= nc and inf.n_gram_members * 12 // 8 > (8 << 20)
    want = oracle_records(csr)
    for _ in range(2):   # second call: pooled streams / events are reused
        res = idx.pairwise(device=0)
        assert_records_equal(res.records(), want)
        res.close()
    dev = idx.upload(device=0)
    assert_records_equal(dev.run_all_passes().records(), want)


@pytest.mark.parametrize("mbits,rel", [("16", "0"), ("16", "12"), ("16", "10"), ("32", None), ("32", "12"), ("32", "0"), ("16", "bitmap"), ("32", "bitmap")])
def test_gram_members_relative_to_their_window(ksb, gram_env, monkeypatch, mbits, rel):
    """Gram members shipped as offsets inside their window (IndexDesc::gram_rel_bits; the default for indexes whose ids
    need 32 bits is 16-bit offsets): kb_build, the row-block shards and the spill emitter (k3_emit_csr) decode every
    encoding to the same rows, bit for bit the oracle's."""
    gram_env(16)
    monkeypatch.setenv("KSPIDER_MEMBER_BITS", mbits)
    monkeypatch.setenv("KSPIDER_GRAM_FORM", "bitmap" if rel == "bitmap" else "csr")  # "bitmap": no member lists at all
    if rel in (None, "bitmap"):
        monkeypatch.delenv("KSPIDER_GRAM_REL", raising=False)
    else:
        monkeypatch.setenv("KSPIDER_GRAM_REL", rel)
    csr = _clade_csr(seed=63, n_clades=16, clade=50, fam=4, n_core=150, n_fam=40, n_noise=300, n_samples=5000)
    want = oracle_records(csr)
    idx = ksb.Index.from_csr(**csr, host_threads=3)
    assert idx.info.n_gram_colors > 0 and idx.info.n_gram_windows > 3
    assert_records_equal(idx.pairwise(device=0).records(), want)
    parts = [idx.shard(r, 3).pairwise().records() for r in range(3)]
    assert_records_equal(np.concatenate(parts), want)
    got = idx.pairwise(spill="always", dense_budget_bytes=1).records()
    assert_records_equal(got, want)
