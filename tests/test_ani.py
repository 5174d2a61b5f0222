"""ANI column (`kSpider pairwise --estimate-ani`, ks_pairwise.py:29-82).

CPU: the repr(float) formatter against Python's own, the column writer, the oracle's restatement on
hand-computed values.  GPU (-m gpu): K8 through the C-ABI against oracle/ani_oracle.py —
  * ksize = 1: pow(c, 1.0) is exact, so the column must be BIT-IDENTICAL; this pins the text round
    trip float32 -> "%g" -> double (text6) on every row;
  * ksize = 21 / 31: the device's double-double pow (csrc/ddpow.hpp, correctly rounded in practice) against
    CPython's (glibc, < 1 ulp but not always correctly rounded) on values in (0, 1]: tolerance 2 ulp of 1.0 =
    2 * 2**-53 absolute, and at least 99 % of the rows bit-identical, stated here as the bar for this
    floating-point column (a CPU check of the same header against glibc: 0.09 % of 2e6 values differ, by 1 ulp).
"""
import math
import os
import random
import struct

import numpy as np
import pytest

from util import random_csr

TOL = 2 * 2.0 ** -53


def consistent_totals(csr):
    """totals[i] = 1000 + sum of the weights of the colours holding i, so every containment is <= 1"""
    ids = csr["total_ids"]
    tot = np.full(ids.size, 1000, dtype=np.uint64)
    off = csr["offsets"].astype(np.int64)
    for c in range(off.size - 1):
        m = np.unique(csr["members"][off[c]:off[c + 1]])
        tot[np.searchsorted(ids, m)] += np.uint64(csr["weights"][c])
    out = dict(csr)
    out["total_vals"] = np.minimum(tot, 2 ** 32 - 1).astype(np.uint32)
    return out


def test_repr_formatter_matches_python(built_lib):
    from kspider_b200.api import format_double_repr as f
    rng = random.Random(7)
    vals = [0.0, -0.0, 1.0, 0.5, 1e-4, 9.999e-5, 1e-5, 1e16, 9.999999999999998e15, 1e15, 1e22, 1e-7, 0.1,
            0.30000000000000004, 1 / 3, 5e-324, 1.7976931348623157e308, float("inf"), float("-inf"), 100.0,
            0.9876543210987654, 1e100, 1.5e-100, 123.0, 1e21, 1e-310, 0.9999999999999999, 0.99]
    # both ends of repr's fixed range from either side, integer-valued doubles up to 2^53 and past it, negatives
    for e in range(-5, 18):
        for toward in (0.0, 1e300):
            x = 10.0 ** e
            for _ in range(4):
                vals += [x, -x]
                x = math.nextafter(x, toward)
    vals += [float(rng.randint(0, 10 ** 16)) for _ in range(2000)] + [rng.randint(1, 2 ** 53) / 2.0 ** rng.randint(0, 60) for _ in range(2000)]
    vals += [4503599627370496.5, 9007199254740993.0, 9999999999999998.0, 123456789012345678.0, -0.25, -123456.789]
    for _ in range(100000):
        r = rng.random()
        if r < 0.5:
            vals.append((rng.random() + rng.random()) / 2)  # what the column holds: means of values in [0, 1]
        elif r < 0.8:
            vals.append(struct.unpack("<d", struct.pack("<Q", rng.getrandbits(64)))[0])
        else:
            vals.append(rng.random() * 10.0 ** rng.randint(-20, 20))
    for v in vals:
        if math.isnan(v):
            assert f(v) == "nan"
        else:
            assert f(v) == repr(v), (repr(v), f(v))


def test_ani_column_writer(built_lib, tmp_path):
    from kspider_b200.api import write_ani_column
    from oracle.ani_oracle import ani_column_text
    rng = np.random.default_rng(3)
    vals = np.concatenate([rng.random(5000), [0.0, 1.0, 0.5, 1e-5]])
    for threads in (1, 3):
        p = tmp_path / f"col{threads}.tsv"
        write_ani_column(vals, str(p), threads)
        assert p.read_text() == ani_column_text(vals.tolist())


def test_oracle_restates_the_sourmash_formula():
    from oracle.ani_oracle import avg_ani_of_text, containment_to_ani
    assert containment_to_ani(0, 31) == 0.0 and containment_to_ani(1, 31) == 1.0
    assert containment_to_ani(0.5, 1) == 1 - (1.0 - 0.5)
    assert containment_to_ani(0.25, 2) == 1 - (1.0 - 0.25 ** 0.5) == 0.5
    assert avg_ani_of_text("0.466667", "1", 21) == ((1 - (1.0 - 0.466667 ** (1.0 / 21))) + 1.0) / 2.0
    with pytest.raises(ValueError):
        containment_to_ani(1.5, 21)
    with pytest.raises(ValueError):
        containment_to_ani(float("inf"), 21)


def test_estimate_ani_preconditions_need_no_device(built_lib, tmp_path):
    """ks_pairwise.py:35-36,44-45: the TSV must exist and a scale must be given -- checked before any GPU work"""
    from kspider_b200.ani import estimate_ani
    prefix = str(tmp_path / "nothing")
    with pytest.raises(FileNotFoundError):
        estimate_ani(prefix, 1000)
    open(prefix + "_kSpider_pairwise.tsv", "w").write("source_1\tsource_2\tshared_kmers\tmin_containment\tavg_containment\tmax_containment\n")
    with pytest.raises(ValueError):
        estimate_ani(prefix, 0)


DDPOW_CHECK = r"""
#include "ddpow.hpp"
#include <cstdio>
#include <cstdlib>
#include <random>
int main() {
    std::mt19937_64 rng(7);
    long bad = 0, n = 0, worse = 0;
    const int ks[5] = {21, 31, 51, 2, 7};
    for (int ki = 0; ki < 5; ki++) {
        const double e = 1.0 / ks[ki];
        for (int i = 0; i < 60000; i++) {
            char buf[64];  // six-digit decimals like the TSV text, over the range of containments
            snprintf(buf, sizeof buf, "%.0fe%d", (double)(100000 + rng() % 900000), -6 - (int)(rng() % 10));
            const double x = strtod(buf, nullptr);
            if (!(x > 0 && x < 1)) continue;
            const double a = ksb::dd_pow(x, e), b = pow(x, e);
            n++;
            if (a != b) { bad++; if (a != nextafter(b, 0.0) && a != nextafter(b, 2.0)) worse++; }
        }
    }
    printf("%ld %ld %ld\n", n, bad, worse);
    return 0;
}
"""


def test_ddpow_host_build_agrees_with_glibc(tmp_path):
    """K8's power function (csrc/ddpow.hpp, every step an explicit IEEE operation) compiled for the host with
    contraction off: against glibc's pow -- what CPython's float ** float calls -- on 3e5 six-digit inputs it may
    differ on a fraction of a per cent (glibc is < 1 ulp, not correctly rounded), and never by more than one ulp."""
    import shutil
    import subprocess
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    src = tmp_path / "ddpow_check.cpp"
    src.write_text(DDPOW_CHECK)
    exe = tmp_path / "ddpow_check"
    inc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "kspider_b200", "csrc")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-I", inc, str(src), "-o", str(exe)], check=True)
    n, bad, worse = (int(v) for v in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split())
    assert n > 250000 and worse == 0 and bad / n < 0.003, (n, bad, worse)


@pytest.fixture(scope="module")
def ksb(built_lib):
    from kspider_b200.capi import check
    check(built_lib.ksb_cuda_available())
    import kspider_b200
    return kspider_b200


@pytest.mark.gpu
@pytest.mark.parametrize("ksize", [1, 21, 31])
def test_ani_column_against_oracle(ksb, ksize):
    from oracle.ani_oracle import ani_column_of_records
    csr = consistent_totals(random_csr(seed=91, n_samples=400, n_colors=30000, s_lo=1, s_hi=9, big=(80, 200), w_hi=50))
    idx = ksb.Index.from_csr(**csr, host_threads=2)
    res = idx.pairwise(want_records=True)
    rec = res.records()
    got, bad = res.ani_column(ksize)
    want = np.array(ani_column_of_records(rec, ksize))
    assert bad == 0 and got.size == rec.size > 50000
    assert len(np.unique(rec["min_containment"])) > 10000  # a real spread of containments, tiny ones included
    if ksize == 1:
        assert np.array_equal(got.view(np.uint64), want.view(np.uint64))
    else:
        assert np.max(np.abs(got - want)) <= TOL
        assert np.mean(got == want) > 0.99  # the rest differ in the last place
    res.close()
    idx.close()


@pytest.mark.gpu
def test_ani_counts_rows_the_reference_would_reject(ksb):
    """totals that make a containment exceed 1 (or inf: total 0): sourmash raises ValueError; K8 counts them"""
    from oracle.ani_oracle import avg_ani_of_text
    csr = random_csr(seed=5, n_samples=60, n_colors=900, s_lo=2, s_hi=6, w_hi=1000)
    csr["total_vals"][:] = 5000
    csr["total_vals"][::7] = 0
    idx = ksb.Index.from_csr(**csr, host_threads=1)
    res = idx.pairwise(want_records=True)
    rec = res.records()
    got, bad = res.ani_column(21)
    n_bad = 0
    for k, r in enumerate(rec):
        try:
            want = avg_ani_of_text("%g" % float(r["min_containment"]), "%g" % float(r["max_containment"]), 21)
            assert abs(got[k] - want) <= TOL
        except ValueError:
            n_bad += 1
    assert bad == n_bad > 0
    res.close()
    idx.close()


@pytest.mark.gpu
def test_estimate_ani_follows_the_row_order_of_the_tsv_on_disk(ksb, tmp_path):
    """The mirror of `kSpider pairwise --estimate-ani`: the TSV is written first (drop-in pairwise), its
    rows shuffled as the reference's hash-map order would, and the column must line up with THAT file.
    ksize = 1 makes the expected text exact."""
    from kspider_b200 import kSpider_internal
    from kspider_b200.ani import estimate_ani
    from oracle.ani_oracle import ani_column_of_tsv, ani_column_text
    csr = consistent_totals(random_csr(seed=17, n_samples=120, n_colors=5000, s_lo=1, s_hi=8, w_hi=30))
    prefix = str(tmp_path / "idx")
    off = csr["offsets"]
    cids = np.arange(1, off.size, dtype=np.uint64)
    ksb.write_index_files(prefix, cids, off, csr["members"], cids, csr["weights"].astype(np.uint64), csr["total_ids"], csr["total_vals"])
    with pytest.raises(FileNotFoundError):
        estimate_ani(prefix, 1000)
    kSpider_internal.pairwise(prefix, 2)
    tsv = prefix + "_kSpider_pairwise.tsv"
    lines = open(tsv).read().splitlines()
    rows = lines[1:]
    random.Random(4).shuffle(rows)
    open(tsv, "w").write("\n".join([lines[0]] + rows) + "\n")
    with pytest.raises(ValueError):
        estimate_ani(prefix, 0)
    for ksize in (1, 31):
        open(prefix + ".extra", "w").write(f"{ksize}\n1\n1\n-1\n")
        out = estimate_ani(prefix, 1000, user_threads=2)
        assert out == prefix + "_kSpider_pairwise.ani_col.tsv"
        want = ani_column_of_tsv(tsv, ksize)
        if ksize == 1:
            assert open(out).read() == ani_column_text(want)
        else:
            got = open(out).read().splitlines()
            assert got[0] == "avg_ani" and len(got) == len(want) + 1
            assert max(abs(float(a) - b) for a, b in zip(got[1:], want)) <= TOL
