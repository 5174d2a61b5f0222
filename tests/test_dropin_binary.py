"""The two native callers of the drop-in symbol kSpider::pairwise(std::string, int) (include/kSpider.hpp:11):
  * kspider_b200/bin/pairwise -- csrc/pairwise_main.cpp, the reference's `pairwise <prefix> <threads>` binary
    (/root/reference/pairwise.cpp:3-5, CMakeLists.txt:49-50), built by csrc/Makefile;
  * a caller compiled HERE from three lines against csrc/kspider_dropin.hpp and linked to libkspider_b200.so:
    the C++ symbol is linked and CALLED, not looked up.
CPU: both link, and both fail loudly without a device.  GPU: both produce the reference's golden TSV."""
import gzip
import os
import shutil
import subprocess

import pytest

from util import GOLD

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "kspider_b200")
BIN = os.path.join(PKG, "bin", "pairwise")
CALLER = r'''
#include "kspider_dropin.hpp"
int main(int argc, char** argv) { kSpider::pairwise(argv[1], 2); return argc > 1 ? 0 : 1; }
'''


def build_caller(tmp_path) -> str:
    src = tmp_path / "caller.cpp"
    src.write_text(CALLER)
    exe = str(tmp_path / "caller")
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.run([cxx, "-std=c++17", "-O1", "-I", os.path.join(PKG, "csrc"), str(src), "-o", exe,
                    "-L", PKG, "-lkspider_b200", f"-Wl,-rpath,{PKG}"], check=True)
    return exe


def golden_rows(name):
    p = os.path.join(GOLD, name + ".pairwise.tsv")
    with (gzip.open(p + ".gz", "rt") if os.path.exists(p + ".gz") else open(p)) as f:
        return f.read()


def stage(name, tmp_path) -> str:
    prefix = str(tmp_path / name)
    for suf in ("_color_to_sources.bin", "_color_count.bin", "_groupID_to_kmerCount.bin"):
        shutil.copy(os.path.join(GOLD, name + suf), prefix + suf)
    return prefix


def test_symbol_is_exported_with_the_reference_mangling(built_lib):
    out = subprocess.run(["nm", "-D", "--defined-only", os.path.join(PKG, "libkspider_b200.so")], check=True,
                         stdout=subprocess.PIPE, text=True).stdout
    # kSpider::pairwise(std::__cxx11::basic_string<char, ...>, int) -- what pairwise.cpp:4 and a SWIG wrapper reference
    assert "_ZN7kSpider8pairwiseENSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEi" in out


def test_binary_and_caller_link_and_refuse_to_run_without_a_device(built_lib, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present: covered by the GPU test below")
    assert os.access(BIN, os.X_OK), "kspider_b200/bin/pairwise is built by csrc/Makefile (__graft_entry__.build)"
    prefix = stage("micro_kat", tmp_path)
    for exe in (BIN, build_caller(tmp_path)):
        p = subprocess.run([exe, prefix, "2"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert p.returncode != 0                      # no CPU fallback: no TSV appears
        assert not os.path.exists(prefix + "_kSpider_pairwise.tsv")
    p = subprocess.run([BIN], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert p.returncode == 2 and "usage" in p.stderr  # the reference dereferences argv[2] == NULL here


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["micro_kat", "clades", "mid_colors"])
def test_binary_and_caller_write_the_reference_tsv(built_lib, name, tmp_path):
    want = golden_rows(name)
    for k, exe in enumerate((BIN, build_caller(tmp_path))):
        d = tmp_path / f"run{k}"
        d.mkdir()
        prefix = stage(name, d)
        p = subprocess.run([exe, prefix, "3"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert p.returncode == 0, p.stderr
        for line in ("mapping colors to groups: ", "parsing index colors: ", "kmer counting: ",
                     "pairwise hashmap construction: ", "writing pairwise matrix to " + prefix + "_kSpider_pairwise.tsv"):
            assert line in p.stdout
        assert open(prefix + "_kSpider_pairwise.tsv").read() == want  # our rows are sorted like the stored goldens
        assert open(prefix + "_kSpider_seqToKmersNo.tsv").read() == open(os.path.join(GOLD, name + ".seq.tsv")).read()
