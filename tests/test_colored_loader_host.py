"""Host-side behaviour of ksb_index_load_colored without a device (CPU suite): malformed files are rejected before
any GPU work, a well-formed set fails LOUDLY with KSB_ERR_CUDA (the histogram is device work; no CPU fallback)."""
import ctypes as C
import os
import shutil

import pytest

from kspider_b200 import capi
from util import GOLD

FILES = [".phmap", "colors.intvectors", "colorTableConfig.json", ".namesMap", ".extra"]


def load(lib, prefix):
    h = C.c_void_p()
    return lib.ksb_index_load_colored(prefix.encode(), 2, -1, C.byref(h)), h


def test_fixture_is_the_references_own_output():
    """first line = colour count, one `id size members` line per colour, as intVectorsTable::save writes (colorTable.cpp:271-281)"""
    lines = open(os.path.join(GOLD, "ckf_smallcolors.intvectors")).read().splitlines()
    assert int(lines[0]) == len(lines) - 1
    assert all(int(l.split()[1]) == len(l.split()) - 2 for l in lines[1:])
    dump = [l.split() for l in open(os.path.join(GOLD, "ckf_small.dump.txt"))]
    assert dump[0][0] == "K" and int(dump[0][2]) == 0                      # no orphan k-mers
    assert sum(int(t[2]) for t in dump if t[0] == "C") == int(dump[0][1])  # every k-mer is counted once


def test_errors_before_the_device(built_lib, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("device present: tests/test_gpu_colored_loader.py covers the whole path")
    rc, _ = load(built_lib, str(tmp_path / "nothing"))
    assert rc == capi.KSB_ERR_IO
    prefix = str(tmp_path / "x")
    for suf in FILES:
        shutil.copy(os.path.join(GOLD, "ckf_small" + suf), prefix + suf)
    open(prefix + "colors.intvectors", "w").write("3\n1 2 5 6\n")
    rc, _ = load(built_lib, prefix)
    assert rc == capi.KSB_ERR_FORMAT and b"declared" in built_lib.ksb_last_error()
    shutil.copy(os.path.join(GOLD, "ckf_smallcolors.intvectors"), prefix + "colors.intvectors")
    rc, _ = load(built_lib, prefix)
    assert rc == capi.KSB_ERR_CUDA and b"no CPU fallback" in built_lib.ksb_last_error()
