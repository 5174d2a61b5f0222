"""the kProcessor `colored_kDataFrame` loader (csrc/colored_loader.cu, ksb_index_load_colored)
against fixtures written AND read back by the reference's own code (tests/golden/make_golden_ckf.py:
lib/kProcessor/src/colorTable.cpp unmodified + the vendored phmap), and against the unmodified src/pairwise.cpp run
on the .bin state the same index corresponds to."""
import os
import shutil

import numpy as np
import pytest

from util import GOLD

pytestmark = pytest.mark.gpu
FILES = [".phmap", "colors.intvectors", "colorTableConfig.json", ".namesMap", ".extra"]


@pytest.fixture(scope="module")
def ksb(built_lib):
    from kspider_b200.capi import check
    check(built_lib.ksb_cuda_available())
    import kspider_b200
    return kspider_b200


def reference_dump(name):
    colors, totals = {}, {}
    for line in open(os.path.join(GOLD, name + ".dump.txt")):
        t = line.split()
        if t[0] == "C":
            colors[int(t[1])] = (int(t[2]), sorted(int(x) for x in t[4:]))
        elif t[0] == "T":
            totals[int(t[1])] = int(t[2])
    return colors, totals


def test_loader_reads_what_the_reference_reads(ksb):
    idx = ksb.Index.load_colored(os.path.join(GOLD, "ckf_small"), host_threads=3)
    colors, totals = reference_dump("ckf_small")
    inf = idx.info
    multi = {c: v for c, v in colors.items() if len(v[1]) >= 2}
    assert inf.n_colors == len(multi) and inf.n_colors_file == len(colors)
    assert inf.n_members == sum(len(v[1]) for v in multi.values())
    assert inf.n_zero_colors == sum(1 for v in multi.values() if v[0] == 0)   # colours whose k-mers are all gone
    ids, tot = idx.sample_ids, idx.totals
    names = [int(l.split()[0]) for l in open(os.path.join(GOLD, "ckf_small.namesMap")).read().splitlines()[1:]]
    assert ids.tolist() == sorted(names)                                       # every name is a sample
    assert {int(i): int(t) for i, t in zip(ids, tot)} == {i: totals.get(i, 0) for i in names}


def test_pairwise_from_colored_index_equals_unmodified_reference(ksb, tmp_path):
    prefix = str(tmp_path / "ckf_small")
    for suf in FILES:
        shutil.copy(os.path.join(GOLD, "ckf_small" + suf), prefix + suf)
    idx = ksb.Index.load_colored(prefix, host_threads=2)
    res = idx.pairwise()
    res.write_tsv(prefix + "_kSpider_pairwise.tsv", 2)
    assert open(prefix + "_kSpider_pairwise.tsv").read() == open(os.path.join(GOLD, "ckf_small.pairwise.tsv")).read()


def test_malformed_inputs_fail_loudly(ksb, tmp_path):
    from kspider_b200.capi import KsbError
    prefix = str(tmp_path / "bad")
    for suf in FILES:
        shutil.copy(os.path.join(GOLD, "ckf_small" + suf), prefix + suf)
    lines = open(prefix + "colors.intvectors").read().splitlines()
    # a colour that k-mers carry is missing from the table
    victim = next(l for l in lines[1:] if l.split()[0] == "1001")
    open(prefix + "colors.intvectors", "w").write("\n".join([str(len(lines) - 2)] + [l for l in lines[1:] if l is not victim]) + "\n")
    with pytest.raises(KsbError, match="does not list"):
        ksb.Index.load_colored(prefix)
    # fewer lines than declared, a size that does not match its line
    open(prefix + "colors.intvectors", "w").write("\n".join([str(len(lines) + 5)] + lines[1:]) + "\n")
    with pytest.raises(KsbError, match="declared"):
        ksb.Index.load_colored(prefix)
    open(prefix + "colors.intvectors", "w").write("\n".join(lines[:1] + [lines[1] + " 77"] + lines[2:]) + "\n")
    with pytest.raises(KsbError, match="exactly"):
        ksb.Index.load_colored(prefix)
    open(prefix + "colors.intvectors", "w").write("\n".join(lines) + "\n")
    blob = open(prefix + ".phmap", "rb").read()
    open(prefix + ".phmap", "wb").write(blob[: len(blob) // 2])
    with pytest.raises(KsbError):
        ksb.Index.load_colored(prefix)
    os.remove(prefix + ".phmap")
    with pytest.raises(KsbError):
        ksb.Index.load_colored(prefix)


def test_duplicate_colour_line_last_one_wins(ksb, tmp_path):
    """colorTable.cpp:237 `colors[color] = ...`: a repeated id overwrites"""
    prefix = str(tmp_path / "dup")
    for suf in FILES:
        shutil.copy(os.path.join(GOLD, "ckf_small" + suf), prefix + suf)
    lines = open(prefix + "colors.intvectors").read().splitlines()
    first = lines[1].split()
    lines.insert(1, f"{first[0]} 2 1 2")          # same id, other members, BEFORE the real line
    lines[0] = str(len(lines) - 1)
    open(prefix + "colors.intvectors", "w").write("\n".join(lines) + "\n")
    a = ksb.Index.load_colored(prefix).pairwise().records()
    b = ksb.Index.load_colored(os.path.join(GOLD, "ckf_small")).pairwise().records()
    assert np.array_equal(a, b)
