"""Host side of the file-driven consumers (csrc/table_reader.cpp): the multithreaded TSV parser against Python's
own int() / float() on the same text (ks_clustering.py:97-101, ks_pairwise.py:68-73), and the binary side-car.
CPU only -- the arithmetic on these columns is GPU work (tests/test_gpu_cluster_golden.py)."""
import ctypes as C
import os

import numpy as np
import pytest

from kspider_b200.api import RECORD_DTYPE
from kspider_b200.capi import KsbError, check
from util import GOLD, stage_golden_prefix

HEADER = "source_1\tsource_2\tshared_kmers\tmin_containment\tavg_containment\tmax_containment\n"


def read_table(lib, tsv, col, col2=0, ani=None, threads=3):
    h = C.c_void_p()
    check(lib.ksb_table_read_tsv(tsv.encode(), col, col2, ani.encode() if ani else None, threads, C.byref(h)))
    return h


def columns(lib, h):
    n = C.c_uint64()
    s1, s2, d, c = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
    check(lib.ksb_table_get(h, C.byref(n), C.byref(s1), C.byref(s2), C.byref(d), C.byref(c)))
    def view(p, ct, dt):
        return np.frombuffer((ct * n.value).from_address(p.value), dtype=dt).copy() if p.value and n.value else None
    return n.value, view(s1, C.c_uint32, np.uint32), view(s2, C.c_uint32, np.uint32), view(d, C.c_double, np.float64), view(c, C.c_float, np.float32)


@pytest.mark.parametrize("name", ["micro_kat", "clades", "mid_colors", "sparse_ids_zero_w"])
@pytest.mark.parametrize("threads", [1, 5])
def test_tsv_parse_equals_python(built_lib, name, threads, tmp_path):
    if name == "sparse_ids_zero_w":
        prefix = str(tmp_path / name)
        import shutil
        shutil.copy(os.path.join(GOLD, name + ".pairwise.tsv"), prefix + "_kSpider_pairwise.tsv")
    else:
        prefix = stage_golden_prefix(name, tmp_path)
    tsv = prefix + "_kSpider_pairwise.tsv"
    rows = [l.strip().split("\t") for l in open(tsv).read().splitlines()[1:]]
    for col in (3, 4, 5):
        h = read_table(built_lib, tsv, col, threads=threads)
        n, s1, s2, d, c = columns(built_lib, h)
        built_lib.ksb_table_free(h)
        assert n == len(rows) and c is None
        assert s1.tolist() == [int(r[0]) for r in rows] and s2.tolist() == [int(r[1]) for r in rows]
        want = np.array([float(r[col]) for r in rows])
        assert np.array_equal(d.view(np.uint64), want.view(np.uint64))  # bit for bit, NaN included


def test_tsv_parse_special_values_and_ani_column(built_lib, tmp_path):
    tsv = str(tmp_path / "t.tsv")
    ani = str(tmp_path / "a.tsv")
    vals = ["inf", "-nan", "nan", "0", "1", "2.31972e-05", "1e-320", "0.466667", "1e999", " 0.5 ", "+0.25"]
    with open(tsv, "w") as f:
        f.write(HEADER)
        for k, v in enumerate(vals):
            f.write(f"{k + 1}\t{k + 2}\t7\t{v}\t0.5\t{v}\n")
    anis = ["0.9507929037223541", "1.0", "0.0", "nan", "5e-324", "0.1", "0.2", "0.3", "0.4", "0.5", "0.6", "0.7"]  # one spare line
    open(ani, "w").write("avg_ani\n" + "\n".join(anis) + "\n")
    h = read_table(built_lib, tsv, 5, 3, threads=4)
    n, s1, s2, d, _ = columns(built_lib, h)
    built_lib.ksb_table_free(h)
    want = np.array([float(v) for v in vals])
    assert n == len(vals) and np.array_equal(d.view(np.uint64), want.view(np.uint64))
    h = read_table(built_lib, tsv, 6, 0, ani, threads=2)
    n, _, _, d, _ = columns(built_lib, h)
    built_lib.ksb_table_free(h)
    assert np.array_equal(d.view(np.uint64), np.array([float(v) for v in anis[:n]]).view(np.uint64))
    open(ani, "w").write("avg_ani\n0.5\n")  # fewer values than rows: the reference's next() raises StopIteration
    with pytest.raises(KsbError):
        read_table(built_lib, tsv, 6, 0, ani)
    open(tsv, "a").write("x\ty\t1\t1\t1\t1\n")
    with pytest.raises(KsbError):
        read_table(built_lib, tsv, 5)
    with pytest.raises(KsbError):
        read_table(built_lib, str(tmp_path / "missing.tsv"), 5)


def test_sidecar_round_trip_and_invalidation(built_lib, tmp_path):
    tsv = str(tmp_path / "p_kSpider_pairwise.tsv")
    side = str(tmp_path / "p_kSpider_pairwise.ksb")
    rec = np.zeros(1000, dtype=RECORD_DTYPE)
    rng = np.random.default_rng(3)
    rec["source_1"] = rng.integers(1, 500, 1000)
    rec["source_2"] = rec["source_1"] + rng.integers(1, 500, 1000).astype(np.uint32)
    rec["shared_kmers"] = rng.integers(0, 1 << 40, 1000)
    for f in ("min_containment", "avg_containment", "max_containment"):
        rec[f] = rng.random(1000, dtype=np.float32)
    open(tsv, "w").write(HEADER + "1\t2\t3\t0.1\t0.2\t0.3\n")
    st = os.stat(tsv)
    check(built_lib.ksb_write_sidecar_records(rec[:600].ctypes.data, 600, side.encode(), st.st_size, st.st_mtime_ns, 0))
    check(built_lib.ksb_write_sidecar_records(rec[600:].ctypes.data, 400, side.encode(), st.st_size, st.st_mtime_ns, 1))
    assert built_lib.ksb_sidecar_matches(side.encode(), tsv.encode()) == 1
    h = C.c_void_p()
    check(built_lib.ksb_table_read_sidecar(side.encode(), 4, 0, 3, C.byref(h)))
    n, s1, s2, d, c = columns(built_lib, h)
    built_lib.ksb_table_free(h)
    assert n == 1000 and d is None
    assert np.array_equal(s1, rec["source_1"]) and np.array_equal(s2, rec["source_2"])
    assert np.array_equal(c.view(np.uint32), rec["avg_containment"].view(np.uint32))
    # the TSV rewritten by anybody (same size, later mtime) or of another size: the side-car no longer belongs to it
    os.utime(tsv, ns=(st.st_atime_ns, st.st_mtime_ns + 1_000_000))
    assert built_lib.ksb_sidecar_matches(side.encode(), tsv.encode()) == 0
    os.utime(tsv, ns=(st.st_atime_ns, st.st_mtime_ns))
    assert built_lib.ksb_sidecar_matches(side.encode(), tsv.encode()) == 1
    open(side, "ab").write(b"x")  # truncated / padded file
    assert built_lib.ksb_sidecar_matches(side.encode(), tsv.encode()) == 0
    with pytest.raises(KsbError):
        check(built_lib.ksb_table_read_sidecar(side.encode(), 4, 0, 1, C.byref(h)))
