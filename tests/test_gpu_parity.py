"""GPU parity tests proper (-m gpu): the CUDA path, called through the C-ABI, against the oracle
on the same seeded inputs, against the committed golden vectors of the unmodified reference, and
— at sizes the oracle cannot reach — through size-independent properties.
Bar: bit-exact (integers equal, float32 containments equal by bit pattern, TSV text equal)."""
import glob
import gzip
import os

import numpy as np
import pytest

from util import assert_records_equal, canonical_partition, oracle_records, random_csr, sort_records

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
CASES = sorted(os.path.basename(p)[:-len(".raw.npz")] for p in glob.glob(os.path.join(GOLD, "*.raw.npz")))


@pytest.fixture(scope="module")
def ksb(built_lib):
    from kspider_b200.capi import check
    check(built_lib.ksb_cuda_available())
    import kspider_b200
    return kspider_b200


def golden_rows(name):
    p = os.path.join(GOLD, name + ".pairwise.tsv")
    f = gzip.open(p + ".gz", "rt") if os.path.exists(p + ".gz") else open(p)
    with f:
        header = f.readline()
        return header, f.read().splitlines()


def write_golden_files(ksb, name, prefix):
    z = np.load(os.path.join(GOLD, name + ".raw.npz"))
    ksb.write_index_files(prefix, z["color_ids"], z["offsets"], z["members"], z["count_ids"], z["count_vals"],
                          z["total_ids"], z["total_vals"])


def copy_golden_bins(name, prefix):
    """the three index files exactly as the reference's own vendored phmap dumped them"""
    import shutil
    for suf in ("_color_to_sources.bin", "_color_count.bin", "_groupID_to_kmerCount.bin"):
        shutil.copy(os.path.join(GOLD, name + suf), prefix + suf)


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("writer", ["phmap", "ours"])
def test_dropin_tsv_equals_reference_golden(ksb, name, writer, tmp_path, capfd):
    """_kSpider_internal.pairwise(prefix, threads) on the golden inputs: TSV text identical to the
    unmodified reference's (as sorted rows: the reference's row order is hash-map order).
    writer=phmap: the index files are the bytes phmap itself dumped; writer=ours: the same content
    written by ksb_write_index_files (slot order may differ, which only permutes seqToKmersNo rows)."""
    from kspider_b200 import kSpider_internal
    prefix = str(tmp_path / name)
    if writer == "phmap":
        copy_golden_bins(name, prefix)
    else:
        write_golden_files(ksb, name, prefix)
    kSpider_internal.pairwise(index_prefix=prefix, user_threads=3)
    out = capfd.readouterr().out
    for line in ("mapping colors to groups: ", "parsing index colors: ", "kmer counting: ",
                 "pairwise hashmap construction: ", "writing pairwise matrix to " + prefix + "_kSpider_pairwise.tsv"):
        assert line in out
    header, want = golden_rows(name)
    with open(prefix + "_kSpider_pairwise.tsv") as f:
        assert f.readline() == header
        got = f.read().splitlines()
    assert got == want  # our rows come out sorted by (source_1, source_2)
    got_seq = open(prefix + "_kSpider_seqToKmersNo.tsv").read()
    want_seq = open(os.path.join(GOLD, name + ".seq.tsv")).read()
    if writer == "phmap":
        assert got_seq == want_seq  # rows in the file's slot order, like the reference's map iteration
    else:
        strip = lambda t: sorted(tuple(l.split("\t")[1:]) for l in t.splitlines()[1:])
        assert got_seq.splitlines()[0] == want_seq.splitlines()[0]
        assert strip(got_seq) == strip(want_seq)


@pytest.mark.parametrize("name", CASES)
def test_loads_phmap_written_files(ksb, name):
    idx = ksb.Index.load(os.path.join(GOLD, name), 2)
    rec = idx.pairwise().records()
    _, want = golden_rows(name)
    got = [f"{r['source_1']}\t{r['source_2']}\t{r['shared_kmers']}\t{ksb.format_float(float(r['min_containment']))}\t"
           f"{ksb.format_float(float(r['avg_containment']))}\t{ksb.format_float(float(r['max_containment']))}" for r in rec]
    assert got == want


SEEDED = [
    ("tiny", dict(seed=1, n_samples=5, n_colors=8, s_lo=1, s_hi=4)),
    ("pairs_only", dict(seed=2, n_samples=50, n_colors=3000, s_lo=2, s_hi=2)),
    ("small_all_sizes", dict(seed=3, n_samples=200, n_colors=6000, s_lo=1, s_hi=64)),
    ("boundary_32_33_64_65", dict(seed=4, n_samples=300, n_colors=10, s_lo=2, s_hi=3, big=(31, 32, 33, 63, 64, 65, 66))),
    ("mid", dict(seed=5, n_samples=1500, n_colors=200, s_lo=2, s_hi=20, big=(255, 256, 257, 511, 512, 513, 769, 1025, 1499))),
    ("sparse_ids", dict(seed=6, n_samples=400, n_colors=3000, s_lo=1, s_hi=40, id_mode="sparse", big=(100, 399))),
    ("zero_weights", dict(seed=7, n_samples=120, n_colors=2000, s_lo=1, s_hi=30, zero_every=5, big=(90,))),
    ("big_weights", dict(seed=8, n_samples=64, n_colors=5000, s_lo=2, s_hi=64, w_hi=2 ** 32 - 1)),
    ("whole_population_colour", dict(seed=9, n_samples=700, n_colors=50, s_lo=2, s_hi=10, big=(700, 700))),
]


@pytest.mark.parametrize("name,kw", SEEDED, ids=[s[0] for s in SEEDED])
def test_pairwise_bit_exact_vs_oracle(ksb, name, kw):
    csr = random_csr(**kw)
    idx = ksb.Index.from_csr(**csr, host_threads=3)
    res = idx.pairwise(device=0)
    got = res.records()
    want = oracle_records(csr)
    assert_records_equal(got, want)
    assert np.array_equal(got, sort_records(got)), "records must come out sorted by (source_1, source_2)"
    inf = res.info
    assert inf.n_rows == want.size and inf.n_updates == idx.info.n_updates and inf.gpu_launches >= 3


def test_empty_and_degenerate(ksb):
    # no multi-member colour at all -> no rows
    csr = random_csr(10, 20, 30, 1, 1)
    res = ksb.Index.from_csr(**csr).pairwise()
    assert res.records().size == 0 and res.info.n_rows == 0
    # totals missing for some members -> 0 -> inf / nan exactly like operator[] default (pairwise.cpp:257-258)
    csr = random_csr(11, 30, 200, 2, 6)
    csr["total_ids"], csr["total_vals"] = csr["total_ids"][:10], csr["total_vals"][:10]
    got = ksb.Index.from_csr(**csr).pairwise().records()
    want = oracle_records(csr)
    assert_records_equal(got, want)
    assert np.isinf(want["max_containment"]).any()
    # zero totals AND zero weights -> 0/0 = -nan
    csr = random_csr(12, 12, 60, 2, 4, zero_every=2)
    csr["total_vals"][:] = 0
    got = ksb.Index.from_csr(**csr).pairwise().records()
    want = oracle_records(csr)
    assert np.isnan(want["min_containment"]).any()
    assert_records_equal(got, want)
    # single sample
    one = dict(offsets=np.array([0, 1], np.uint64), members=np.array([9], np.uint32), weights=np.array([3], np.uint32),
               total_ids=np.array([9], np.uint32), total_vals=np.array([4], np.uint32))
    assert ksb.Index.from_csr(**one).pairwise().records().size == 0


@pytest.mark.parametrize("world", [2, 3, 8])
def test_row_blocks_are_disjoint_and_complete(ksb, world):
    """Multi-GPU decomposition run rank by rank on one device: blocks disjoint, union == oracle."""
    csr = random_csr(21, 500, 4000, 1, 50, big=(120, 300, 499), zero_every=17)
    idx = ksb.Index.from_csr(**csr)
    parts, upd, rows = [], 0, []
    for r in range(world):
        res = idx.pairwise(rank=r, world=world)
        inf = res.info
        rows.append((inf.row_begin, inf.row_end))
        upd += inf.n_updates
        parts.append(res.records())
    assert rows[0][0] == 0 and rows[-1][1] == idx.info.n_samples
    assert all(rows[i][1] == rows[i + 1][0] for i in range(world - 1))
    assert upd == idx.info.n_updates
    got = np.concatenate(parts)
    assert_records_equal(got, oracle_records(csr))
    assert np.array_equal(got, sort_records(got))  # blocks concatenate in order


def test_dense_budget_forces_multiple_passes(ksb):
    csr = random_csr(22, 600, 3000, 1, 40, big=(200, 599), zero_every=19)
    idx = ksb.Index.from_csr(**csr)
    dev = idx.upload(dense_budget_bytes=200_000)  # 25k cells per pass of 179700
    assert dev.n_passes >= 7
    res = dev.run_all_passes()
    assert_records_equal(res.records(), oracle_records(csr))
    # and combined with row blocks
    parts = [idx.pairwise(rank=r, world=2, dense_budget_bytes=300_000).records() for r in range(2)]
    assert_records_equal(np.concatenate(parts), oracle_records(csr))


def test_accumulate_is_idempotent_and_triangle_matches(ksb):
    """accumulate() clears first: running it twice gives the same block; the dense block read back
    equals the oracle's shared counts cell by cell."""
    csr = random_csr(23, 150, 2500, 2, 33, big=(80,))
    idx = ksb.Index.from_csr(**csr)
    dev = idx.upload()
    dev.accumulate()
    dev.accumulate()
    ptr, cells, rb, re = dev.triangle()
    n = int(idx.info.n_samples)
    lo, hi, band_cells = idx.bands()
    assert (cells, rb, re) == (band_cells, 0, n) and band_cells <= n * (n - 1) // 2
    assert np.all(lo[hi > lo] > np.nonzero(hi > lo)[0]), "bands lie strictly above the diagonal"
    start = np.concatenate([[0], np.cumsum(hi.astype(np.int64) - lo)])
    host = dev.read_triangle()
    want = oracle_records(csr)
    ids = idx.sample_ids
    pos = {int(v): i for i, v in enumerate(ids)}
    dense = np.zeros(cells, dtype=np.uint64)
    i = np.array([pos[int(v)] for v in want["source_1"]], dtype=np.int64)
    j = np.array([pos[int(v)] for v in want["source_2"]], dtype=np.int64)
    assert np.all((j >= lo[i]) & (j < hi[i])), "every touched pair lies inside its row band"
    dense[start[i] + (j - lo[i])] = want["shared_kmers"]
    assert np.array_equal(host, dense)


@pytest.mark.parametrize("dist,cutoff", [("max_cont", 0.0), ("max_cont", 0.3), ("min_cont", 0.1), ("avg_cont", 0.2),
                                         ("max_cont", 1.0)])
def test_cluster_partition_vs_oracle(ksb, dist, cutoff, tmp_path):
    """Threshold on the PRINTED text + connected components (ks_clustering.py:97-120) as a partition."""
    from oracle import cluster_oracle
    rng = np.random.default_rng(31)
    csr = random_csr(31, 300, 500, 2, 5)
    # totals comparable to the shared counts so containments spread over (0, 1]
    csr["total_vals"] = rng.integers(500, 4000, size=300).astype(np.uint32)
    idx = ksb.Index.from_csr(**csr)
    res = idx.pairwise()
    prefix = str(tmp_path / "c")
    res.write_tsv(prefix + "_kSpider_pairwise.tsv", 2)
    with open(prefix + ".namesMap", "w") as f:
        f.write("300\n" + "".join(f"{i} g{i}\n" for i in range(1, 301)))
    labels, n_edges = res.cluster(300, dist, cutoff)
    col = cluster_oracle.DISTANCE_TO_COL[dist]
    edges = cluster_oracle.edges_from_tsv(prefix + "_kSpider_pairwise.tsv", col, cutoff)
    assert n_edges == len(edges)
    assert canonical_partition(labels) == canonical_partition(cluster_oracle.components(300, edges))
    assert all(labels[g[0]] == g[0] for g in canonical_partition(labels)), "label = smallest member"
    # the mirrored command writes the reference's file name and the same partition
    from kspider_b200 import clustering
    idx.write_seq_to_kmers(prefix + "_kSpider_seqToKmersNo.tsv")
    want = cluster_oracle.cluster(prefix, cutoff, dist)
    out = clustering.cluster(prefix, cutoff, dist)  # file-driven, like the reference: parses the TSV on disk
    assert out == cluster_oracle.output_name(prefix, cutoff)
    assert cluster_oracle.read_clusters_file(out) == want
    live = str(tmp_path / "live.tsv")                # and from the live result, nothing read back from disk
    clustering.cluster_result(idx, res, cluster_oracle.load_names_map(prefix + ".namesMap"), live, cutoff, dist)
    assert cluster_oracle.read_clusters_file(live) == want


def test_cluster_merge_of_row_blocks(ksb):
    csr = random_csr(32, 400, 700, 2, 4)
    csr["total_vals"] = np.random.default_rng(32).integers(300, 3000, size=400).astype(np.uint32)
    idx = ksb.Index.from_csr(**csr)
    full, _ = idx.pairwise().cluster(400, "max_cont", 0.2)
    parts = [idx.pairwise(rank=r, world=3).cluster(400, "max_cont", 0.2)[0] for r in range(3)]
    merged = ksb.cluster_merge(ksb.cluster_merge(parts[0], parts[1]), parts[2])
    assert np.array_equal(merged, full)


def test_full_size_properties_cfg2_scaled(ksb):
    """BASELINE config 2 shape (Zipf sizes <= 64, N = 1000) at 1 M colours — too many updates for a
    quick oracle run, so checked through invariants: checksum of checksums and per-sample row sums."""
    from kspider_b200 import synth
    s = synth.zipf_index(1000, 1_000_000)
    idx = ksb.Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=4)
    assert idx.info.n_updates == s.n_updates
    rec = idx.pairwise().records()
    sizes = np.diff(s.offsets).astype(np.int64)
    w = s.weights.astype(np.int64)
    assert int(rec["shared_kmers"].sum()) == int((w * sizes * (sizes - 1) // 2).sum())
    # sum over pairs containing sample i = sum over colours containing i of w*(s-1)
    per = np.zeros(1001, dtype=np.int64)
    np.add.at(per, rec["source_1"], rec["shared_kmers"].astype(np.int64))
    np.add.at(per, rec["source_2"], rec["shared_kmers"].astype(np.int64))
    want = np.bincount(s.members, weights=np.repeat((w * (sizes - 1)).astype(np.float64), sizes), minlength=1001)
    assert np.array_equal(per, want.astype(np.int64))
    assert rec.size == 499500 and np.array_equal(rec, sort_records(rec))


def test_banded_accumulator_on_clade_structured_index(ksb):
    """Colours confined to clades / families (BASELINE configs 3 and 4 in miniature): the accumulator
    stores narrow row bands (band cells << N(N-1)/2, rows without any pair have no cells); K5 must
    decode (i, j) from the banded layout, also per row block and over several dense passes."""
    from kspider_b200 import synth
    s = synth.fracminhash_index(n_clades=12, clade_size=20, sketch=120, core=60, family_clades=3, family_pool=30, seed=7)
    csr = dict(offsets=s.offsets, members=s.members, weights=s.weights, total_ids=s.total_ids, total_vals=s.total_vals)
    idx = ksb.Index.from_csr(**csr)
    n = int(idx.info.n_samples)
    lo, hi, cells = idx.bands()
    assert cells < n * (n - 1) // 4 and (hi == lo).any()
    want = oracle_records(csr)
    got = idx.pairwise().records()
    assert_records_equal(got, want)
    assert np.array_equal(got, sort_records(got))
    parts = [idx.pairwise(rank=r, world=3).records() for r in range(3)]
    assert_records_equal(np.concatenate(parts), want)
    dev = idx.upload(dense_budget_bytes=8 * 3000, spill="never")
    assert dev.n_passes >= 3
    assert_records_equal(dev.run_all_passes().records(), want)


@pytest.mark.parametrize("cells", ["32", "64"])
@pytest.mark.parametrize("bits", ["10", "12", "16", "32"])
def test_k1_kernels_and_member_packings_agree(ksb, cells, bits, monkeypatch):
    """K1 has two kernels (k1_small_c32: 32-bit cell arithmetic, the default below 2^32 owned cells;
    k1_small: 64-bit row addresses, also the spill sink's) and the small / Gram member sections four
    packings (10-bit <= 1024 samples, 12-bit <= 4096, u16, u32).  Every combination, on the whole
    triangle and on row blocks (where the owned-row runs and the ballot masks matter), against the
    oracle.  No Gram routing, so that all colours <= 64 go through K1."""
    monkeypatch.setenv("KSPIDER_K1_CELLS", cells)
    monkeypatch.setenv("KSPIDER_GRAM", "off")
    if bits != "10":
        monkeypatch.setenv("KSPIDER_MEMBER_BITS", bits)
    csr = random_csr(seed=123, n_samples=700, n_colors=9000, s_lo=1, s_hi=64, big=(31, 32, 33, 63, 64), zero_every=37)
    want = oracle_records(csr)
    idx = ksb.Index.from_csr(**csr, host_threads=2)
    assert_records_equal(idx.pairwise().records(), want)
    parts = [idx.pairwise(rank=r, world=3).records() for r in range(3)]
    assert_records_equal(np.concatenate(parts), want)
    parts = [idx.shard(r, 5).pairwise().records() for r in range(5)]
    assert_records_equal(np.concatenate(parts), want)


@pytest.mark.parametrize("world", [2, 3, 7])
def test_row_block_shards_equal_row_blocks(ksb, world):
    """A device made from rank r's shard produces the same records as the whole index run with
    rank=r/world, and all shards together are the oracle's result."""
    csr = random_csr(seed=91, n_samples=400, n_colors=6000, s_lo=1, s_hi=50, big=(80, 200, 399), zero_every=23)
    idx = ksb.Index.from_csr(**csr, host_threads=2)
    parts = []
    for r in range(world):
        sh = idx.shard(r, world)
        got = sh.pairwise().records()
        assert_records_equal(got, idx.pairwise(rank=r, world=world).records())
        parts.append(got)
    assert_records_equal(np.concatenate(parts), oracle_records(csr))


def test_cluster_labels_left_on_device_and_merged_there(ksb):
    """ksb_cluster_pct_device + ksb_cluster_merge_device (what the multi-GPU step uses between the
    NCCL all-gather): same labels as the host-returning calls."""
    import ctypes as C
    import torch
    from kspider_b200.capi import check, lib
    csr = random_csr(seed=93, n_samples=300, n_colors=3000, s_lo=1, s_hi=12, big=(40,))
    idx = ksb.Index.from_csr(**csr)
    n = int(idx.info.n_samples)
    whole, ne = idx.pairwise().cluster(n, "max_cont", 0.05)
    parts, dparts, edges = [], [], 0
    for r in range(3):
        res = idx.shard(r, 3).pairwise()
        lab, e1 = res.cluster(n, "max_cont", 0.05)
        t = torch.empty(n, dtype=torch.int32, device="cuda:0")
        e2 = res.cluster_into(t.data_ptr(), "max_cont", 0.05)
        torch.cuda.synchronize()
        assert e1 == e2 and np.array_equal(t.cpu().numpy().view(np.uint32), lab)
        parts.append(lab)
        dparts.append(t)
        edges += e1
    assert edges == ne
    allv = torch.cat(dparts)
    out = torch.empty(n, dtype=torch.int32, device="cuda:0")
    check(lib().ksb_cluster_merge_device(n, 3, C.c_void_p(allv.data_ptr()), C.c_void_p(out.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert canonical_partition(out.cpu().numpy().view(np.uint32)) == canonical_partition(whole)


def test_l2_blocking_of_atomic_kernels(ksb, monkeypatch):
    """K1/K2 launched once per row sub-block (KSPIDER_L2_BLOCK_MB; on by default only for blocks far
    larger than L2): same result, also on a row-block shard and with the pipelined upload."""
    monkeypatch.setenv("KSPIDER_GRAM", "off")
    csr = random_csr(seed=95, n_samples=900, n_colors=20000, s_lo=2, s_hi=40, big=(100, 300, 899), zero_every=29)
    want = ksb.Index.from_csr(**csr).pairwise().records()
    monkeypatch.setenv("KSPIDER_L2_BLOCK_MB", "1")
    idx = ksb.Index.from_csr(**csr)
    dev = idx.upload()
    got = dev.run_all_passes().records()
    assert dev.info.gpu_launches >= 6  # several K1 + K2 launches
    assert_records_equal(got, want)
    assert_records_equal(idx.pairwise().records(), want)
    parts = [idx.shard(r, 2).pairwise().records() for r in range(2)]
    assert_records_equal(np.concatenate(parts), want)


@pytest.mark.parametrize("name", ["small_dense", "mid_colors"])
def test_dropin_on_two_gpus_writes_the_same_tsv(ksb, name, tmp_path, monkeypatch, capfd):
    """KSPIDER_GPUS=2: each device gets the row-block shard of its rows; the concatenated TSV is the
    golden one.  Needs two devices (skipped on a single-GPU box)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from kspider_b200 import kSpider_internal
    prefix = str(tmp_path / name)
    copy_golden_bins(name, prefix)
    monkeypatch.setenv("KSPIDER_GPUS", "2")
    kSpider_internal.pairwise(index_prefix=prefix, user_threads=4)
    capfd.readouterr()
    header, want = golden_rows(name)
    with open(prefix + "_kSpider_pairwise.tsv") as f:
        assert f.readline() == header
        got = f.read().splitlines()
    assert got == want


@pytest.mark.parametrize("name", ["cfg3", "cfg4"])
def test_full_size_properties_clade_structured(ksb, name):
    """BASELINE configs 3 and 4 at FULL size (10 k / 100 k samples; windowed K4, banded accumulator,
    u16 resp. u32 members): too many updates for the oracle, so size-independent properties —
    checksum of checksums, per-sample row sums, sortedness — and, for cfg3, equality with the atomic
    path (KSPIDER_GRAM=off) on the same index."""
    from kspider_b200 import synth
    s = getattr(synth, name)()
    idx = ksb.Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=8)
    assert idx.info.n_updates == s.n_updates and idx.info.n_gram_windows > 100
    rec = idx.pairwise().records()
    sizes = np.diff(s.offsets).astype(np.int64)
    w = s.weights.astype(np.int64)
    assert int(rec["shared_kmers"].astype(np.int64).sum()) == int((w * sizes * (sizes - 1) // 2).sum())
    n = int(s.total_ids.max()) + 1
    per = np.bincount(rec["source_1"], weights=rec["shared_kmers"].astype(np.float64), minlength=n) + \
        np.bincount(rec["source_2"], weights=rec["shared_kmers"].astype(np.float64), minlength=n)
    want = np.bincount(s.members, weights=np.repeat((w * (sizes - 1)).astype(np.float64), sizes), minlength=n)
    assert np.array_equal(per.astype(np.int64), want.astype(np.int64))   # sums < 2^53: exact in float64
    key = rec["source_1"].astype(np.int64) * (1 << 32) + rec["source_2"].astype(np.int64)
    assert np.all(np.diff(key) > 0)
    if name == "cfg3":
        old = os.environ.get("KSPIDER_GRAM")
        os.environ["KSPIDER_GRAM"] = "off"
        try:
            ref = ksb.Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=8).pairwise().records()
        finally:
            if old is None:
                os.environ.pop("KSPIDER_GRAM", None)
            else:
                os.environ["KSPIDER_GRAM"] = old
        assert_records_equal(rec, ref)


def test_host_only_records_release_the_device_copies(ksb):
    """ksb_opts.want_records = 2 (what the TSV-writing drop-in uses): the same rows on the host, pass by pass, and the
    device copies are gone -- clustering such a result is refused loudly (44 B per row stayed resident)"""
    from kspider_b200.capi import KsbError
    csr = random_csr(seed=61, n_samples=400, n_colors=3000, s_lo=2, s_hi=6)
    idx = ksb.Index.from_csr(**csr)
    want = oracle_records(csr)
    res = idx.pairwise(want_records="host", dense_budget_bytes=200_000)   # several passes
    assert_records_equal(res.records(), want)
    with pytest.raises(KsbError, match="host records only"):
        res.cluster(400, "max_cont", 0.1)
    res2 = idx.pairwise(want_records=True)   # default budget: cells and rows of a pass both fit
    assert_records_equal(res2.records(), want)
    assert res2.cluster(400, "max_cont", 0.1)[1] >= 0
