"""CPU-only checks of the product's host side: the C-ABI library loads and exports every symbol the
header declares, the phmap-format loader/writer agree with phmap's own bytes, the text layer
agrees with printf, and compute entry points fail loudly without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle import pairwise_oracle as po
from util import random_csr

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _no_gpu(lib):
    return lib.ksb_cuda_available() != 0


def test_header_symbols_all_exported(built_lib):
    from kspider_b200 import capi
    hdr = open(os.path.join(ROOT, "include", "kspider_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ksb_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 35
    for name in sorted(declared):
        assert hasattr(built_lib, name), f"libkspider_b200.so does not export {name}"
    assert declared == set(capi.exported_names()), declared ^ set(capi.exported_names())
    assert built_lib.ksb_abi_version() == 1


def test_dropin_cxx_symbol_exported():
    """kSpider::pairwise(std::string, int) with the reference's mangled name (include/kSpider.hpp:11)."""
    from kspider_b200 import capi
    L = C.CDLL(capi.LIB_PATH)
    assert hasattr(L, "_ZN7kSpider8pairwiseENSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEi")


@pytest.mark.parametrize("name", ["micro_kat", "small_dense"])
def test_loader_on_phmap_bytes(built_lib, name, tmp_path):
    from kspider_b200 import Index
    ix = Index.load(os.path.join(GOLD, name), host_threads=2)
    o = po.OracleIndex(os.path.join(GOLD, name))
    info = ix.info
    sizes = np.diff(o.offsets).astype(np.int64)
    multi = sizes >= 2
    assert info.n_colors == int(multi.sum())
    assert info.n_members == int(sizes[multi].sum())
    assert info.n_updates == int((sizes[multi] * (sizes[multi] - 1) // 2).sum())
    ids = np.unique(np.concatenate([o.members, o.total_ids]))
    assert np.array_equal(ix.sample_ids, ids)
    want_tot = dict(zip(o.total_ids.tolist(), o.total_vals.tolist()))
    assert ix.totals.tolist() == [want_tot.get(i, 0) for i in ids.tolist()]
    out = tmp_path / "seq.tsv"
    ix.write_seq_to_kmers(str(out))
    assert out.read_text() == open(os.path.join(GOLD, name + ".seq.tsv")).read()


def test_micro_kat_zero_count_singleton_is_not_a_zero_colour(built_lib):
    from kspider_b200 import Index
    info = Index.load(os.path.join(GOLD, "micro_kat")).info
    assert (info.n_samples, info.n_colors, info.n_updates, info.n_zero_colors) == (4, 3, 5, 0)


def test_writer_bytes_equal_phmap(built_lib, tmp_path):
    """ksb_write_index_files must produce what phmap itself dumps for the same insertion order."""
    from kspider_b200 import write_index_files
    for name in ("micro_kat", "small_dense"):
        z = np.load(os.path.join(GOLD, name + ".raw.npz"))
        prefix = str(tmp_path / name)
        write_index_files(prefix, z["color_ids"], z["offsets"], z["members"], z["count_ids"], z["count_vals"],
                          z["total_ids"], z["total_vals"])
        for suf in ("_groupID_to_kmerCount.bin", "_color_count.bin", "_color_to_sources.bin"):
            a = open(prefix + suf, "rb").read()
            b = open(os.path.join(GOLD, name + suf), "rb").read()
            assert len(a) == len(b), suf
            # phmap leaves one ctrl byte per dump uninitialised (the last of the 16 cloned bytes);
            # everything else must match byte for byte
            if "sources" in suf:
                continue  # member sets: slot order depends on phmap's growth history; content checked below
            diff = [i for i in range(len(a)) if a[i] != b[i]]
            assert len(diff) <= 1, (suf, diff[:10])
        mine, ref = po.OracleIndex(prefix), po.OracleIndex(os.path.join(GOLD, name))
        assert np.array_equal(mine.color_ids, ref.color_ids) and np.array_equal(mine.offsets, ref.offsets)
        assert np.array_equal(mine.weights, ref.weights)
        for c in range(mine.color_ids.size):
            lo, hi = int(mine.offsets[c]), int(mine.offsets[c + 1])
            assert sorted(mine.members[lo:hi].tolist()) == sorted(ref.members[lo:hi].tolist())


def test_loader_roundtrip_random(built_lib, tmp_path):
    from kspider_b200 import Index, write_index_files
    csr = random_csr(3, 200, 1500, 1, 30, id_mode="sparse", zero_every=13, big=(70, 199))
    n = csr["weights"].size
    cids = np.arange(10, 10 + n, dtype=np.uint64)
    prefix = str(tmp_path / "rt")
    write_index_files(prefix, cids, csr["offsets"], csr["members"], cids, csr["weights"].astype(np.uint64),
                      csr["total_ids"], csr["total_vals"])
    a, b = Index.load(prefix, 4).info, Index.from_csr(**csr).info
    for f in ("n_samples", "n_colors", "n_members", "n_updates", "n_small_colors", "n_mid_colors", "n_zero_colors",
              "max_color_size", "triangle_cells"):
        assert getattr(a, f) == getattr(b, f), f
    assert a.n_mid_colors == 2 and a.n_zero_colors > 0
    # the oracle reads the same files to the same colours
    o = po.OracleIndex(prefix)
    assert int((np.diff(o.offsets) >= 2).sum()) == a.n_colors


def test_loader_errors(built_lib, tmp_path):
    from kspider_b200 import Index
    from kspider_b200.capi import KsbError, KSB_ERR_IO, KSB_ERR_FORMAT, KSB_ERR_ARG
    with pytest.raises(KsbError) as e:
        Index.load(str(tmp_path / "nope"))
    assert e.value.code == KSB_ERR_IO
    prefix = str(tmp_path / "bad")
    for suf in ("_color_to_sources.bin", "_color_count.bin", "_groupID_to_kmerCount.bin"):
        open(prefix + suf, "wb").write(b"\x07" * 40)
    with pytest.raises(KsbError) as e:
        Index.load(prefix)
    assert e.value.code == KSB_ERR_FORMAT
    with pytest.raises(KsbError) as e:  # duplicate member inside a colour: not a set
        Index.from_csr(np.array([0, 3], np.uint64), np.array([1, 2, 1], np.uint32), np.array([1], np.uint32),
                       np.array([1, 2], np.uint32), np.array([5, 5], np.uint32))
    assert e.value.code == KSB_ERR_ARG


def test_format_float_matches_printf(built_lib):
    from kspider_b200 import format_float
    rng = np.random.default_rng(0)
    vals = np.concatenate([rng.random(2000, dtype=np.float32), (rng.random(500) * 1e-5).astype(np.float32),
                           np.array([0, 1, 0.5, 0.3, 0.466666669, 1e-7, 123456.7, 1234567.0, np.inf, -np.nan, np.nan],
                                    dtype=np.float32)])
    for v in vals:
        assert format_float(float(v)) == po.format_float(float(v)) == ("%g" % float(v) if np.isfinite(v) else po.format_float(float(v)))


def test_fast_format_float_is_exact_on_a_sweep(built_lib):
    """The writer's integer fast path (1e-4 <= v < 1e6) against printf's correctly rounded "%g":
    random values, ratios of small integers (what containments are), exact half-way cases and every
    float next to a power of ten."""
    import ctypes as C
    buf = C.create_string_buffer(64)

    def fmt(x):
        n = built_lib.ksb_format_float(C.c_float(x), buf)
        return buf.raw[:n].decode()
    rng = np.random.default_rng(1)
    a, b = rng.integers(1, 5000, 60000), rng.integers(1, 5000, 60000)
    vals = [rng.random(60000).astype(np.float32),
            (rng.random(60000) * 10.0 ** rng.integers(-6, 7, 60000)).astype(np.float32),
            np.minimum(a, b).astype(np.float32) / np.maximum(a, b).astype(np.float32),
            np.arange(1, 1024, dtype=np.float32) / np.float32(1024),
            np.array([999999.0, 999999.5, 999999.4, 123456.5, 123455.5, 0.9999995, 0.99999994, 99999.95, 12345.65], np.float32)]
    edge = []
    for c in (1e-4, 1e-3, 1e-2, 0.1, 1.0, 10.0, 1e5, 1e6):
        for toward in (0.0, 1e9):
            x = np.float32(c)
            for _ in range(40):
                edge.append(x)
                x = np.nextafter(x, np.float32(toward))
    vals.append(np.array(edge, dtype=np.float32))
    for v in np.concatenate(vals):
        assert fmt(float(v)) == "%g" % float(v), float(v)


def test_tsv_writer_on_host_rows_is_the_reference_text(built_lib, tmp_path):
    """src/pairwise.cpp:242-275 on rows that never saw a GPU: header, tab-separated integers, ostream<<float text, file order
    = row order whatever the thread count and across the writer's batch boundary; appended blocks follow without a header."""
    from kspider_b200.api import RECORD_DTYPE
    rng = np.random.default_rng(7)
    n = (1 << 22) + 12345  # one full batch of the writer and a ragged second one
    rec = np.zeros(n, RECORD_DTYPE)
    rec["source_1"] = rng.integers(0, 2 ** 32, n, dtype=np.uint64)
    rec["source_2"] = rng.integers(0, 2 ** 32, n, dtype=np.uint64)
    rec["shared_kmers"] = rng.integers(0, 2 ** 63, n, dtype=np.uint64) >> rng.integers(0, 63, n, dtype=np.uint64)
    rec["shared_kmers"][:3] = (0, 2 ** 64 - 1, 10 ** 19)
    a, b = rng.integers(1, 70000, n), rng.integers(1, 70000, n)
    rec["min_containment"] = (np.minimum(a, b) / np.maximum(a, b)).astype(np.float32)
    rec["avg_containment"] = (rng.random(n) * 10.0 ** rng.integers(-8, 3, n)).astype(np.float32)
    rec["max_containment"] = rng.random(n, dtype=np.float32)
    rec["max_containment"][:4] = (0.0, 1.0, 1e-7, 0.99999994)

    def text_of(r):
        return (f"{r['source_1']}\t{r['source_2']}\t{r['shared_kmers']}\t{'%g' % float(r['min_containment'])}\t"
                f"{'%g' % float(r['avg_containment'])}\t{'%g' % float(r['max_containment'])}")

    header = "source_1\tsource_2\tshared_kmers\tmin_containment\tavg_containment\tmax_containment"
    paths = []
    for threads in (1, 3, 8):
        p = str(tmp_path / f"t{threads}.tsv")
        assert built_lib.ksb_write_tsv_records(rec.ctypes.data, n, p.encode(), threads, 0) == 0
        paths.append(p)
    blobs = [open(p, "rb").read() for p in paths]
    assert blobs[0] == blobs[1] == blobs[2]
    lines = blobs[0].decode().split("\n")
    assert lines[0] == header and lines[-1] == "" and len(lines) == n + 2
    for k in np.concatenate([np.arange(64), np.arange((1 << 22) - 32, (1 << 22) + 32), np.arange(n - 64, n), rng.integers(0, n, 4000)]):
        assert lines[1 + k] == text_of(rec[k]), k
    # the three numeric columns of every line parse back to the rows (no line lost or torn at a thread / batch seam)
    import pandas as pd
    back = pd.read_csv(paths[0], sep="\t", dtype={"source_1": np.uint32, "source_2": np.uint32, "shared_kmers": np.uint64})
    assert np.array_equal(back["source_1"].to_numpy(), rec["source_1"]) and np.array_equal(back["source_2"].to_numpy(), rec["source_2"])
    assert np.array_equal(back["shared_kmers"].to_numpy(), rec["shared_kmers"])

    # two blocks, the second appended == one call; an empty block writes the header alone; a missing directory is an IO error
    p2 = str(tmp_path / "two.tsv")
    cut = 1000
    assert built_lib.ksb_write_tsv_records(rec.ctypes.data, cut, p2.encode(), 4, 0) == 0
    assert built_lib.ksb_write_tsv_records(rec[cut:].ctypes.data, 5000 - cut, p2.encode(), 4, 1) == 0
    assert open(p2, "rb").read() == ("\n".join(lines[:5001]) + "\n").encode()
    p3 = str(tmp_path / "none.tsv")
    assert built_lib.ksb_write_tsv_records(None, 0, p3.encode(), 4, 0) == 0
    assert open(p3).read() == header + "\n"
    from kspider_b200.capi import KSB_ERR_IO, KSB_ERR_ARG
    assert built_lib.ksb_write_tsv_records(rec.ctypes.data, 10, str(tmp_path / "no" / "dir.tsv").encode(), 1, 0) == KSB_ERR_IO
    assert built_lib.ksb_write_tsv_records(None, 10, p3.encode(), 1, 0) == KSB_ERR_ARG


def test_threshold_float_is_the_exact_boundary(built_lib):
    """c* = smallest float32 whose printed text passes `float(text)*100 >= cutoff*100` (ks_clustering.py:97-105)."""
    from kspider_b200 import threshold_float

    def kept(c, cutoff):
        return not (float("%g" % float(np.float32(c))) * 100 < float(cutoff) * 100)

    for cutoff in (0.0, 0.05, 0.1, 0.3, 0.5, 0.7, 0.95, 0.999999, 1.0, 0.123456789, 1e-7):
        cs = np.float32(threshold_float(cutoff))
        assert kept(cs, cutoff)
        if cs > 0:
            below = np.nextafter(cs, np.float32(-1), dtype=np.float32)
            assert not kept(below, cutoff), (cutoff, cs, below)
    rng = np.random.default_rng(1)
    for cutoff in rng.random(50):
        cs = np.float32(threshold_float(float(cutoff)))
        for c in np.concatenate([rng.random(40, dtype=np.float32), [cs]]):
            assert kept(c, cutoff) == bool(np.float32(c) >= cs)


def test_compute_fails_loudly_without_device(built_lib):
    """No CPU fallback: on a box without an sm_100 device every compute entry point errors."""
    if not _no_gpu(built_lib):
        pytest.skip("a GPU is present")
    from kspider_b200 import Index
    from kspider_b200.capi import KsbError, KSB_ERR_CUDA
    ix = Index.from_csr(**random_csr(1, 10, 20))
    with pytest.raises(KsbError) as e:
        ix.pairwise()
    assert e.value.code == KSB_ERR_CUDA
    with pytest.raises(KsbError):
        ix.upload()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "kspider_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cpp", ".cu", ".cuh", ".hpp", ".h")):
                src = open(os.path.join(dirpath, fn), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), fn
                assert "oracle/" not in src or fn == "README", fn


def test_row_bands_are_tight_and_cover_every_pair(built_lib, oracle_lib):
    """Loader: band_lo/band_hi of a row = exactly the min / max+1 column it is paired with in the
    oracle's rows (zero-weight colours included); untouched rows have an empty band."""
    import kspider_b200 as ksb
    from util import oracle_records, random_csr
    csr = random_csr(seed=71, n_samples=90, n_colors=150, s_lo=1, s_hi=5, id_mode="sparse", zero_every=9)
    idx = ksb.Index.from_csr(**csr, host_threads=3)
    lo, hi, cells = idx.bands()
    ids = idx.sample_ids
    want = oracle_records(csr)
    i = np.searchsorted(ids, want["source_1"])
    j = np.searchsorted(ids, want["source_2"])
    n = ids.size
    wlo = np.full(n, 2 ** 32 - 1, dtype=np.int64)
    whi = np.zeros(n, dtype=np.int64)
    np.minimum.at(wlo, i, j)
    np.maximum.at(whi, i, j + 1)
    wlo[whi == 0] = 0
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)
    assert cells == int((whi - wlo).sum()) and cells < n * (n - 1) // 2


@pytest.mark.parametrize("world,bits", [(2, None), (5, None), (2, "12"), (2, "16"), (2, "32")])
def test_row_block_shards_hold_exactly_the_needed_colours(built_lib, world, bits, monkeypatch):
    """ksb_index_shard: rank r's shard = colours with a row in [cut[r], cut[r+1]), cut down to their
    members >= cut[r]; the owned rows keep their bands, and the shards' updates add up to U.
    Runs in every member packing of the small / Gram sections (10-bit by default at N = 300; 12-bit,
    u16, u32 forced): the shard is rebuilt from the PACKED sections, so a wrong pack or unpack shows."""
    from kspider_b200 import Index
    from util import random_csr
    if bits:
        monkeypatch.setenv("KSPIDER_MEMBER_BITS", bits)
    csr = random_csr(seed=77, n_samples=300, n_colors=4000, s_lo=1, s_hi=40, big=(70, 150, 290), zero_every=31)
    idx = Index.from_csr(**csr, host_threads=2)
    cuts = idx.row_cuts(world).astype(np.int64)
    lo, hi, _ = idx.bands()
    ids = np.unique(np.concatenate([csr["total_ids"], csr["members"]]))
    off = csr["offsets"].astype(np.int64)
    total = 0
    for r in range(world):
        sh = idx.shard(r, world, host_threads=2)
        inf = sh.info
        assert inf.n_samples == idx.info.n_samples and np.array_equal(sh.sample_ids, idx.sample_ids)
        assert np.array_equal(sh.totals, idx.totals)
        want_members = want_colors = want_upd = 0
        for c in range(off.size - 1):
            m = np.sort(np.searchsorted(ids, csr["members"][off[c]:off[c + 1]]))
            t = m[m >= cuts[r]]
            if t.size >= 2 and t[0] < cuts[r + 1]:
                want_colors += 1
                want_members += t.size
                if csr["weights"][c] != 0:
                    p = np.nonzero(t < cuts[r + 1])[0]
                    want_upd += int((t.size - 1 - p).sum())
        assert (inf.n_colors, inf.n_members, inf.n_updates) == (want_colors, want_members, want_upd)
        slo, shi, _ = sh.bands()
        a, b = int(cuts[r]), int(cuts[r + 1])
        assert np.array_equal(slo[a:b], lo[a:b]) and np.array_equal(shi[a:b], hi[a:b])
        total += inf.n_updates
        with pytest.raises(Exception):
            sh.shard(0, 2)
        # the (descriptor, payload) pair rebuilds the same host index in another process (ksb_index_from_blob)
        cp = Index.from_blob(sh.descriptor(), sh.payload())
        assert cp.descriptor() == sh.descriptor() and np.array_equal(cp.payload(), sh.payload())
        assert np.array_equal(cp.sample_ids, idx.sample_ids) and cp.info.n_updates == inf.n_updates
        with pytest.raises(Exception):
            Index.from_blob(sh.descriptor()[:-8], sh.payload())
        with pytest.raises(Exception):
            Index.from_blob(sh.descriptor(), sh.payload()[:-1])
        with pytest.raises(Exception):
            Index.from_blob(b"\0" * len(sh.descriptor()), sh.payload())
    assert total == idx.info.n_updates


def test_gram_windows_shrink_to_the_touched_samples(built_lib, monkeypatch):
    """Loader: the grid window of a Gram-routed colour only groups colours; the window K4 works on is the range
    of samples they touch.  Same routed colours either way, never more MACs, and far fewer when a clade-local
    family straddles the grid (700 samples inside a 2048-wide grid window: 6 x 6 tiles instead of 16 x 16)."""
    from kspider_b200 import Index
    rng = np.random.default_rng(5)
    n, fam0, fam = 6000, 900, 700  # samples 900..1599 fit no 1024-wide grid window (stride 512): the 2048 level
    ids = np.arange(1, n + 1, dtype=np.uint32)
    members, offsets = [], [0]
    for _ in range(400):
        m = np.sort(rng.choice(fam, size=350, replace=False)) + fam0
        members.append(ids[m])
        offsets.append(offsets[-1] + m.size)
    csr = dict(offsets=np.array(offsets, dtype=np.uint64), members=np.concatenate(members).astype(np.uint32),
               weights=rng.integers(1, 200, size=400).astype(np.uint32), total_ids=ids, total_vals=np.full(n, 10 ** 6, dtype=np.uint32))
    monkeypatch.setenv("KSPIDER_GRAM", "16")
    info = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("KSPIDER_GRAM_SHRINK", mode)
        idx = Index.from_csr(**csr, host_threads=2)
        info[mode] = (idx.info.n_gram_colors, idx.info.n_gram_macs, idx.info.n_updates)
        idx.close()
    assert info["0"][0] == info["1"][0] == 400 and info["0"][2] == info["1"][2]
    assert info["1"][1] == 400 * (768 * 767 // 2)           # 700 samples -> six tiles of 128
    assert info["0"][1] == 400 * (2048 * 2047 // 2)         # the grid window itself


def test_from_blob_rejects_a_blob_that_does_not_match_its_descriptor(built_lib):
    """a (descriptor, payload) pair received over NCCL / gloo is checked before any offset in it is trusted"""
    import numpy as np
    from kspider_b200 import Index
    from kspider_b200.capi import KsbError
    from util import random_csr
    idx = Index.from_csr(**random_csr(seed=5, n_samples=200, n_colors=2000, s_lo=1, s_hi=30, big=(70, 150), zero_every=13))
    desc, payload = idx.descriptor(), idx.payload().copy()
    ok = Index.from_blob(desc, payload)
    assert ok.info.n_updates == idx.info.n_updates and np.array_equal(ok.payload(), payload)
    words = np.frombuffer(desc, dtype=np.uint64).copy()
    rng = np.random.default_rng(1)
    rejected = 0
    for _ in range(200):   # any single corrupted 8-byte field of the descriptor: rejected, or harmless (a counter the layout does not use)
        w = words.copy()
        k = int(rng.integers(1, w.size))
        w[k] = w[k] + np.uint64(int(rng.integers(1, 1 << 20))) if rng.random() < 0.5 else np.uint64(int(rng.integers(0, 1 << 62)))
        try:
            Index.from_blob(w.tobytes(), payload)
        except KsbError:
            rejected += 1
    assert rejected >= 60
    with pytest.raises(KsbError):
        Index.from_blob(desc, payload[:-256])          # truncated payload
    bad = payload.copy()
    gram_cols = int(idx.info.n_gram_colors)
    if gram_cols:
        # the Gram offsets are a monotone u64 array ending at n_gram_members: find it and break the monotony
        arr = np.frombuffer(payload.tobytes(), dtype=np.uint64)
        want_last = int(idx.info.n_gram_members)
        hits = [i for i in np.flatnonzero(arr == want_last) if i >= gram_cols and arr[i - gram_cols] == 0 and np.all(np.diff(arr[i - gram_cols:i + 1].astype(np.int64)) >= 0)]
        if hits:
            i = hits[0]
            bad8 = np.frombuffer(bad, dtype=np.uint64)
            bad8 = bad8.copy()
            bad8[i - 1] = np.uint64(want_last + 12345)
            with pytest.raises(KsbError):
                Index.from_blob(desc, np.frombuffer(bad8.tobytes(), dtype=np.uint8))


def test_gram_members_stored_relative_to_their_window(built_lib, monkeypatch):
    """Loader: Gram-routed colours may ship their members as offsets inside their window (IndexDesc::gram_rel_bits:
    2-byte members for an index whose compact ids need 4 bytes).  The encoding never changes what the index says:
    shards rebuilt from the PACKED Gram section of every encoding are byte-identical, and the blob shrinks."""
    from kspider_b200 import Index
    rng = np.random.default_rng(15)
    n = 6000
    ids = np.arange(1, n + 1, dtype=np.uint32)
    members, offsets = [], [0]
    for fam0, fam, s, cnt in ((900, 700, 350, 300), (3000, 100, 60, 500), (5100, 800, 90, 200)):
        for _ in range(cnt):
            m = np.sort(rng.choice(fam, size=s, replace=False)) + fam0
            members.append(ids[m])
            offsets.append(offsets[-1] + m.size)
    nc = len(offsets) - 1
    csr = dict(offsets=np.array(offsets, dtype=np.uint64), members=np.concatenate(members).astype(np.uint32),
               weights=rng.integers(1, 200, size=nc).astype(np.uint32), total_ids=ids, total_vals=np.full(n, 10 ** 6, dtype=np.uint32))
    monkeypatch.setenv("KSPIDER_GRAM", "16")
    shards, sizes = {}, {}
    for mbits, rel in (("16", "0"), ("16", "12"), ("16", "10"), ("32", None), ("32", "0"), ("32", "12"), ("16", "bitmap"), ("32", "auto")):
        monkeypatch.setenv("KSPIDER_MEMBER_BITS", mbits)
        # "bitmap" / "auto": the Gram colours travel as the bitmaps K4 reads (IndexDesc::gram_form 1) -- what the loader
        # picks by itself for colours this dense; the other cases pin the CSR form and vary its member encoding
        monkeypatch.delenv("KSPIDER_GRAM_FORM", raising=False)
        if rel != "auto":
            monkeypatch.setenv("KSPIDER_GRAM_FORM", "bitmap" if rel == "bitmap" else "csr")
        if rel in (None, "bitmap", "auto"):
            monkeypatch.delenv("KSPIDER_GRAM_REL", raising=False)
        else:
            monkeypatch.setenv("KSPIDER_GRAM_REL", rel)
        idx = Index.from_csr(**csr, host_threads=2)
        assert idx.info.n_gram_colors == nc
        sizes[(mbits, rel)] = int(idx.info.payload_bytes)
        cp = Index.from_blob(idx.descriptor(), idx.payload())   # the validator accepts every encoding it wrote
        assert cp.info.n_updates == idx.info.n_updates
        # the shards are built under ONE setting, from the parents' differently encoded sections
        monkeypatch.setenv("KSPIDER_MEMBER_BITS", "16")
        monkeypatch.setenv("KSPIDER_GRAM_REL", "0")
        monkeypatch.setenv("KSPIDER_GRAM_FORM", "csr")
        shards[(mbits, rel)] = []
        for r in range(2):
            sh = idx.shard(r, 2, host_threads=2)   # kept alive while its payload view is copied
            shards[(mbits, rel)].append(bytes(sh.payload()))
            sh.close()
        idx.close()
    base = shards[("16", "0")]
    for key, got in shards.items():
        assert got == base, key
    gm = int(csr["members"].size)
    assert sizes[("16", "0")] - sizes[("16", "12")] >= gm // 2 - 4096          # 16 -> 12 bits per Gram member
    assert sizes[("16", "12")] - sizes[("16", "10")] >= gm // 4 - 4096          # every window here is at most 1024 samples wide
    assert sizes[("32", "0")] - sizes[("32", None)] >= 2 * gm - 4096          # default: 32 -> 16 bits when ids need 32
    assert sizes[("32", None)] - sizes[("32", "12")] >= gm // 2 - 4096
    # bitmaps: about 300 colours x 768 samples + 500 x 128 + 200 x 896 bits in K-blocks of 128 colours (~80 kB) against
    # 153 000 members and 1 000 offsets: smaller than the narrowest CSR, and what the loader picks by itself
    assert sizes[("16", "10")] - sizes[("16", "bitmap")] >= 100_000
    assert sizes[("32", "auto")] < sizes[("32", "12")] and sizes[("32", "auto")] - sizes[("16", "bitmap")] < 16384


@pytest.mark.parametrize("shape", ["close_ids", "wide_ids", "top_of_u32", "mixed"])
def test_loader_sorts_member_lists_like_from_csr(built_lib, tmp_path, shape):
    """The file loader puts a colour's members (hash-slot order in the file) into ascending order either through a bitmap
    over [min, max] (ids close together) or with std::sort (ids far apart): both must give exactly the index that
    ksb_index_from_csr builds from the same lists -- byte-identical blobs -- also for ids at the top of the u32 range."""
    from kspider_b200 import Index, write_index_files
    rng = np.random.default_rng({"close_ids": 1, "wide_ids": 2, "top_of_u32": 3, "mixed": 4}[shape])
    if shape == "close_ids":
        pool = np.arange(1000, 1400, dtype=np.uint64)
    elif shape == "wide_ids":
        pool = np.sort(rng.choice(np.arange(1, 2 ** 32 - 2, 104729, dtype=np.uint64), size=400, replace=False))
    elif shape == "top_of_u32":
        pool = np.arange(2 ** 32 - 401, 2 ** 32 - 1, dtype=np.uint64)
    else:
        pool = np.sort(np.concatenate([np.arange(5, 205, dtype=np.uint64), rng.choice(np.arange(10 ** 6, 4 * 10 ** 9, 7919, dtype=np.uint64), size=200, replace=False)]))
    cols, offsets = [], [0]
    for s in list(rng.integers(2, 90, size=600)) + [300, 399]:
        cols.append(rng.permutation(rng.choice(pool, size=int(s), replace=False)))  # unsorted on purpose
        offsets.append(offsets[-1] + int(s))
    members = np.concatenate(cols).astype(np.uint32)
    csr = dict(offsets=np.array(offsets, dtype=np.uint64), members=members,
               weights=rng.integers(1, 5000, size=len(cols)).astype(np.uint32),
               total_ids=pool.astype(np.uint32), total_vals=rng.integers(1, 10 ** 6, size=pool.size).astype(np.uint32))
    n = len(cols)
    cids = np.arange(7, 7 + n, dtype=np.uint64)
    prefix = str(tmp_path / shape)
    write_index_files(prefix, cids, csr["offsets"], csr["members"], cids, csr["weights"].astype(np.uint64), csr["total_ids"], csr["total_vals"])
    a = Index.load(prefix, 3)
    b = Index.from_csr(**csr, host_threads=3)
    assert a.descriptor() == b.descriptor()
    assert np.array_equal(a.payload(), b.payload())
    o = po.OracleIndex(prefix)
    assert int((np.diff(o.offsets) >= 2).sum()) == a.info.n_colors
    a.close(); b.close()


def test_from_blob_checks_the_bitmap_form_and_the_row_tables(built_lib, monkeypatch):
    """A blob whose Gram section holds bitmaps (IndexDesc::gram_form 1) carries the first bitmap row of every window and
    the accumulator's row tables (bandstart, rowbase): a descriptor / payload pair whose tables do not add up is refused."""
    from kspider_b200 import Index
    from kspider_b200.capi import KsbError
    rng = np.random.default_rng(21)
    n = 3000
    ids = np.arange(1, n + 1, dtype=np.uint32)
    cols, offsets = [], [0]
    for lo in (100, 1200, 2500):
        for _ in range(300):
            m = np.sort(rng.choice(200, size=90, replace=False)) + lo
            cols.append(ids[m]); offsets.append(offsets[-1] + m.size)
    csr = dict(offsets=np.array(offsets, dtype=np.uint64), members=np.concatenate(cols).astype(np.uint32),
               weights=rng.integers(1, 200, size=len(cols)).astype(np.uint32), total_ids=ids, total_vals=np.full(n, 10 ** 6, dtype=np.uint32))
    monkeypatch.setenv("KSPIDER_GRAM", "16")
    monkeypatch.setenv("KSPIDER_GRAM_FORM", "bitmap")
    idx = Index.from_csr(**csr)
    assert idx.info.gram_form & 1 and idx.info.n_gram_windows == 3
    desc, payload = idx.descriptor(), idx.payload().copy()
    assert np.array_equal(Index.from_blob(desc, payload).payload(), payload)
    p64 = np.frombuffer(payload.tobytes(), dtype=np.uint64).copy()
    lo, hi, cells = idx.bands()
    start = np.concatenate([[0], np.cumsum(hi.astype(np.int64) - lo)]).astype(np.uint64)
    # bandstart is the only place in the blob where this exact prefix-sum array stands: find it and break one entry
    hits = [i for i in np.flatnonzero(p64 == start[-1]) if i >= n and np.array_equal(p64[i - n:i + 1], start)]
    assert hits, "bandstart section not found"
    bad = p64.copy()
    bad[hits[0] - n // 2] += np.uint64(1)
    with pytest.raises(KsbError):
        Index.from_blob(desc, np.frombuffer(bad.tobytes(), dtype=np.uint8))
    # the window row table: [0, rows of window 0, ...]; 3 windows of 3 K-blocks x 200-sample extent
    rows = 3 * 200
    want = np.array([0, rows, 2 * rows, 3 * rows], dtype=np.uint64)
    whits = [i for i in range(p64.size - 3) if np.array_equal(p64[i:i + 4], want)]
    assert whits, "window row table not found"
    bad = p64.copy()
    bad[whits[0] + 1] += np.uint64(8)
    with pytest.raises(KsbError):
        Index.from_blob(desc, np.frombuffer(bad.tobytes(), dtype=np.uint8))


def _blob_fields(desc: bytes):
    """The u64 fields of ksb::IndexDesc behind its size-class table (ksb_common.hpp), by name."""
    names = ("off_totals off_sample_ids off_row_updates off_band_lo off_band_hi band_cells off_bandstart off_rowbase "
             "off_small_members off_small_weights off_mid_offsets off_mid_sizes off_mid_members off_mid_weights off_mid_items "
             "n_mid_items n_mid_members_padded off_zero_offsets off_zero_members n_gram_colors n_gram_members n_gram_updates "
             "gram_planes gram_min_size off_gram_offsets off_gram_members off_gram_weights off_gram_wins n_gram_wins "
             "n_small_updates n_gram_macs n_gram_cells member_bits packed_bits weight_bits gram_rel_bits gram_form "
             "gram_bitmap_rows shard_world shard_rank shard_rb shard_re payload_bytes").split()
    head, cls = 14 * 8, 65 * 40
    assert len(desc) == head + cls + 8 * len(names), "IndexDesc changed: update this table"
    tail = np.frombuffer(desc[head + cls:], dtype=np.uint64)
    f = {k: int(v) for k, v in zip(names, tail)}
    # size classes: {s, pairs, magic, pad} u32, {n, member_off, weight_off} u64
    f["cls"] = [dict(zip(("s", "pairs", "magic", "pad"), np.frombuffer(desc[head + 40 * s:head + 40 * s + 16], np.uint32).tolist()),
                     **dict(zip(("n", "member_off", "weight_off"), np.frombuffer(desc[head + 40 * s + 16:head + 40 * s + 40], np.uint64).tolist())))
                for s in range(65)]
    f["cls_at"] = head
    return f


def test_from_blob_scans_the_member_values(built_lib, monkeypatch):
    """The kernels check a row against the owned block and add a column to the row's base as it comes, and the shard
    builder indexes the sample table with what it reads: ksb_index_from_blob refuses a blob with a member list that is not
    strictly ascending below N, a pair outside its row's band, a mid colour or K2 item that does not fit its offsets, sample
    ids out of order, a size class with the wrong pair count, a Gram bit of a colour the window does not have."""
    from kspider_b200 import Index
    from kspider_b200.capi import KsbError
    rng = np.random.default_rng(33)
    n = 5000
    ids = np.arange(1, n + 1, dtype=np.uint32)

    def refused(desc, payload, what):
        with pytest.raises(KsbError) as e:
            Index.from_blob(bytes(desc), payload)
        assert what in str(e.value), str(e.value)

    # ---- (A) small, mid and zero-weight colours, no Gram path; the last sample is in no colour
    cols, w = [], []
    for _ in range(500):
        cols.append(np.sort(rng.choice(n - 1, size=5, replace=False))); w.append(int(rng.integers(1, 50)))
    for _ in range(3):
        cols.append(np.sort(rng.choice(n - 1, size=300, replace=False))); w.append(7)
    for _ in range(20):
        cols.append(np.sort(rng.choice(n - 1, size=4, replace=False))); w.append(0)
    offsets = np.concatenate([[0], np.cumsum([c.size for c in cols])]).astype(np.uint64)
    monkeypatch.setenv("KSPIDER_GRAM", "off")
    idx = Index.from_csr(offsets=offsets, members=ids[np.concatenate(cols)], weights=np.array(w, np.uint32), total_ids=ids,
                         total_vals=np.full(n, 10 ** 6, dtype=np.uint32))
    info = idx.info
    assert info.n_mid_colors == 3 and info.n_zero_colors == 20 and info.n_small_colors == 500 and info.n_gram_colors == 0
    desc, payload = idx.descriptor(), idx.payload().copy()
    f = _blob_fields(desc)
    assert f["packed_bits"] == 16 and f["member_bits"] == 16
    Index.from_blob(desc, payload)

    def u16(p, off):  # view of a u16 section from byte `off` on
        return p[off:].view(np.uint16) if (p.size - off) % 2 == 0 else p[off:-1].view(np.uint16)

    c5 = f["cls"][5]
    assert c5["n"] == 500
    p = payload.copy(); m = u16(p, f["off_small_members"])[c5["member_off"]:]
    m[0], m[1] = m[1], m[0]
    refused(desc, p, "ascending")
    p = payload.copy(); m = u16(p, f["off_small_members"])[c5["member_off"]:]
    m[5 + 2] = m[5 + 1]                          # a sample twice in one colour
    refused(desc, p, "ascending")
    p = payload.copy(); m = u16(p, f["off_small_members"])[c5["member_off"]:]
    m[4] = n - 1                                 # ascending, below N, but no row's band reaches the last sample
    refused(desc, p, "band")
    p = payload.copy(); m = u16(p, f["off_small_members"])[c5["member_off"]:]
    m[4] = n                                     # not a sample
    refused(desc, p, "ascending")
    p = payload.copy(); z = u16(p, f["off_zero_members"])
    z[3] = n - 1
    refused(desc, p, "band")
    # mid colours: 300 members in 304 slots
    p = payload.copy(); p[f["off_mid_sizes"]:].view(np.uint32)[1] = 400
    refused(desc, p, "mid colour")
    p = payload.copy(); p[f["off_mid_sizes"]:].view(np.uint32)[1] = 302      # fits the slots, reads the padding as members
    refused(desc, p, "")
    p = payload.copy(); p[f["off_mid_sizes"]:].view(np.uint32)[2] = 1
    refused(desc, p, "mid colour")
    p = payload.copy(); mm = u16(p, f["off_mid_members"]); mm[10], mm[11] = mm[11], mm[10]
    refused(desc, p, "ascending")
    assert f["n_mid_items"] == 3 * 2             # columns [0, 256) x rows [0, 255), columns [256, 300) x rows [0, 299)
    for word, val in ((2, 512), (2, 8), (1, 512), (1, 128), (0, 3)):   # {color, b0, a0, pad}
        p = payload.copy(); p[f["off_mid_items"]:].view(np.uint32)[4 * 1 + word] = val
        refused(desc, p, "mid item")
    p = payload.copy(); s_ = p[f["off_sample_ids"]:].view(np.uint32); s_[10], s_[11] = s_[11], s_[10]
    refused(desc, p, "sample ids")
    d = bytearray(desc); at = f["cls_at"] + 40 * 5 + 4          # cls[5].pairs
    d[at:at + 4] = (11).to_bytes(4, "little")
    refused(d, payload, "pair count")
    # what the scan does not look at can change: a weight, a total
    p = payload.copy(); p[f["off_totals"]:].view(np.uint32)[7] = 123
    assert Index.from_blob(desc, p).info.n_samples == n

    # ---- (B) Gram bitmaps: one window of 130 colours (two K-blocks, the second with two colours)
    # (compact samples; every colour spans sample 128 and stays inside [64, 192): one window of the 128-wide grid level)
    cols = [np.arange(100, 150)]                                   # colour A: its rows pair with nothing past 170
    for _ in range(128):
        cols.append(np.sort(np.concatenate([[100, 170], rng.choice(np.arange(101, 170), size=38, replace=False)])))
    cols.append(np.concatenate([np.arange(64, 73), np.arange(180, 192)]))   # the only colour that reaches the window's last samples
    offsets = np.concatenate([[0], np.cumsum([c.size for c in cols])]).astype(np.uint64)
    monkeypatch.setenv("KSPIDER_GRAM", "16")
    monkeypatch.setenv("KSPIDER_GRAM_FORM", "bitmap")
    idx = Index.from_csr(offsets=offsets, members=ids[np.concatenate(cols)], weights=np.full(len(cols), 3, np.uint32), total_ids=ids,
                         total_vals=np.full(n, 10 ** 6, dtype=np.uint32))
    assert idx.info.gram_form & 1 and idx.info.n_gram_windows == 1 and idx.info.n_gram_colors == 130
    desc, payload = idx.descriptor(), idx.payload().copy()
    f = _blob_fields(desc)
    Index.from_blob(desc, payload)
    win = payload[f["off_gram_wins"]:f["off_gram_wins"] + 24]
    x0, width, extent = (int(v) for v in win[12:24].view(np.uint32))
    assert (x0, extent) == (64, 128) and int(win[8:12].view(np.uint32)[0]) == 130
    bits = payload[f["off_gram_members"]:f["off_gram_members"] + 16 * 2 * extent].view(np.uint32).reshape(2, extent, 4)
    # the K-block and bit of colour A: the only column with exactly the rows of samples 100..149
    col_a = [(kb, k) for kb in range(2) for k in range(128)
             if np.array_equal(np.flatnonzero((bits[kb, :, k // 32] >> np.uint32(k % 32)) & 1), np.arange(36, 86))]
    assert len(col_a) == 1
    kb, k = col_a[0]
    p = payload.copy()
    b = p[f["off_gram_members"]:f["off_gram_members"] + 16 * 2 * extent].view(np.uint32).reshape(2, extent, 4)
    b[kb, 127, k // 32] |= np.uint32(1 << (k % 32))                 # sample 191 joins colour A: (100, 191) is outside row 100's band
    refused(desc, p, "band")
    p = payload.copy()
    b = p[f["off_gram_members"]:f["off_gram_members"] + 16 * 2 * extent].view(np.uint32).reshape(2, extent, 4)
    b[1, 5, 1] |= np.uint32(1 << 9)                                  # K-block 1 holds two colours: bit 41 is nobody's
    refused(desc, p, "window")


def test_blob_row_cuts_refuses_a_descriptor_that_points_outside_the_payload(built_lib):
    """ksb_blob_row_cuts reads the per-row update counts straight out of a (descriptor, payload) pair: a sample count or
    a section offset that would leave the payload -- also one large enough to wrap the sum -- is an argument error."""
    from kspider_b200 import Index
    from kspider_b200.capi import KSB_ERR_ARG
    csr = random_csr(5, 60, 400, 1, 9)
    idx = Index.from_csr(**csr)
    desc, payload = idx.descriptor(), idx.payload().copy()
    f = _blob_fields(desc)
    cuts = np.zeros(4, np.uint64)
    assert built_lib.ksb_blob_row_cuts(desc, len(desc), payload.ctypes.data, payload.size, 3, cuts.ctypes.data) == 0
    assert cuts[0] == 0 and cuts[-1] == idx.info.n_samples and np.all(np.diff(cuts.astype(np.int64)) >= 0)
    names_at = 14 * 8 + 65 * 40
    for field_at, val in ((8, 2 ** 61 + 3), (8, payload.size), (names_at + 8 * 2, 2 ** 64 - 8), (names_at + 8 * 2, payload.size),
                          (names_at + 8 * 2, f["off_row_updates"] + 4)):      # n_samples; off_row_updates
        d = bytearray(desc)
        d[field_at:field_at + 8] = int(val).to_bytes(8, "little")
        assert built_lib.ksb_blob_row_cuts(bytes(d), len(d), payload.ctypes.data, payload.size, 3, cuts.ctypes.data) == KSB_ERR_ARG
