"""TEST INFRASTRUCTURE ONLY — runs the UNMODIFIED reference Python package
(/root/reference/pykSpider/kSpider2: click CLI, ks_clustering.py, ks_pairwise.py) in a subprocess, with the
three modules it needs and this image lacks supplied by oracle/py_shim (rustworkx, sourmash,
_kSpider_internal -- see their headers for what each does and does not pin).

Only available where /root/reference exists (the build container).  It is used to (1) generate the
committed fixtures tests/golden/*.clusters.json / *.ani_k*.tsv (tests/golden/make_golden_py.py) and (2) pin
oracle/cluster_oracle.py / oracle/ani_oracle.py live in the CPU test-suite.  Nothing on the GPU box calls it.
"""
from __future__ import annotations

import os
import subprocess
import sys
from typing import Sequence

HERE = os.path.dirname(os.path.abspath(__file__))
SHIM = os.path.join(HERE, "py_shim")
REF_PY = "/root/reference/pykSpider"


def have_reference_python() -> bool:
    return os.path.isfile(os.path.join(REF_PY, "kSpider2", "ks_clustering.py"))


def run_cli(args: Sequence[str], timeout: float = 600.0, check: bool = True) -> subprocess.CompletedProcess:
    """`kSpider <args>` = kSpider2.kSpider_main.cli (setup.py entry point), reference sources untouched."""
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([SHIM, REF_PY] + ([env["PYTHONPATH"]] if env.get("PYTHONPATH") else []))
    env.pop("BRANCH_NAME", None)  # kSpider_version.get_version would go to the network
    code = "import sys; from kSpider2.kSpider_main import cli; cli(prog_name='kSpider')"
    return subprocess.run([sys.executable, "-c", code, *[str(a) for a in args]], env=env, timeout=timeout,
                          stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, check=check)


def cluster(index_prefix: str, cutoff01: float, dist_type: str = "max_cont") -> str:
    """`kSpider cluster -i P -c cutoff -d dist` (ks_clustering.py:150-163); returns the output path."""
    run_cli(["cluster", "-i", index_prefix, "-c", repr(float(cutoff01)), "-d", dist_type])
    return index_prefix + f"_kSpider_clusters_{float(cutoff01) * 100}%.tsv"


def estimate_ani(index_prefix: str, scale: int = 1000, check: bool = True) -> subprocess.CompletedProcess:
    """`kSpider pairwise -i P --estimate-ani -s scale` (ks_pairwise.py:29-82) -> P_kSpider_pairwise.ani_col.tsv"""
    return run_cli(["pairwise", "-i", index_prefix, "--estimate-ani", "-s", str(int(scale))], check=check)


def pairwise(index_prefix: str, threads: int = 1) -> subprocess.CompletedProcess:
    """`kSpider pairwise -i P -t T`: the CLI calling _kSpider_internal.pairwise = oracle/_ref/pairwise_ref."""
    return run_cli(["pairwise", "-i", index_prefix, "-t", str(int(threads))])
