"""Generates the cluster / ANI golden fixtures by running the UNMODIFIED reference Python package
(/root/reference/pykSpider/kSpider2: `kSpider cluster`, `kSpider pairwise --estimate-ani`) through
oracle/refpy.py (shims for rustworkx / sourmash / _kSpider_internal only: oracle/py_shim/*), on the committed
pairwise goldens the unmodified src/pairwise.cpp produced (make_golden.py).

Run in the build container (needs /root/reference):
    python tests/golden/make_golden_py.py
For every case with sample ids 1..N (the reference's node = id - 1 needs them, ks_clustering.py:75-76,99-100):
    <name>.clusters.json    {"<dist_type>|<cutoff>": partition as sorted lists of sample ids}, dist types
                            min_cont / avg_cont / max_cont (+ ani where the ANI column exists), cutoffs CUTOFFS
    <name>.ani_k<K>.tsv     the reference's P_kSpider_pairwise.ani_col.tsv, verbatim, for the rows of the
                            committed (sorted) TSV; or <name>.ani_k<K>.err when the reference raises
                            (a printed containment above 1: sourmash's ANIResult rejects the distance)
"""
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import refpy  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import stage_golden_prefix as stage  # noqa: E402

CASES = ["micro_kat", "small_dense", "mid_colors", "wide_counts", "clades"]
CUTOFFS = [0.0, 0.004, 0.02, 0.3, 0.5, 0.7, 0.8, 0.95, 1.0]
KSIZES = {"micro_kat": [1, 21], "clades": [1, 21, 31], "small_dense": [21]}
ANI_CLUSTER_K = 21  # the ani_col file present when dist_type "ani" is clustered


def read_partition(path: str):
    with open(path) as f:
        return sorted(sorted(int(n[1:]) for n in line.rstrip("\n").split(",")) for line in f if line.strip())


def main():
    assert refpy.have_reference_python(), "needs /root/reference"
    for name in CASES:
        out = {}
        with tempfile.TemporaryDirectory() as td:
            prefix = stage(name, td, ANI_CLUSTER_K)
            have_ani = False
            for k in KSIZES.get(name, []):
                with open(prefix + ".extra", "w") as f:
                    f.write(f"{k}\n1\n1\n\n")
                r = refpy.estimate_ani(prefix, 1000, check=False)
                dst = os.path.join(HERE, f"{name}.ani_k{k}")
                for suf in (".tsv", ".err"):
                    if os.path.exists(dst + suf):
                        os.remove(dst + suf)
                if r.returncode == 0:
                    shutil.copy(prefix + "_kSpider_pairwise.ani_col.tsv", dst + ".tsv")
                    if k == ANI_CLUSTER_K:
                        have_ani = True
                        shutil.copy(prefix + "_kSpider_pairwise.ani_col.tsv", prefix + ".keep_ani")
                else:
                    last = [ln for ln in r.stderr.strip().splitlines() if ln.strip()][-1]
                    with open(dst + ".err", "w") as f:
                        f.write(last + "\n")
            if have_ani:
                shutil.copy(prefix + ".keep_ani", prefix + "_kSpider_pairwise.ani_col.tsv")
            for dist in ["min_cont", "avg_cont", "max_cont"] + (["ani"] if have_ani else []):
                for c in CUTOFFS:
                    path = refpy.cluster(prefix, c, dist)
                    out[f"{dist}|{c!r}"] = read_partition(path)
        with open(os.path.join(HERE, name + ".clusters.json"), "w") as f:
            json.dump(out, f, separators=(",", ":"), sort_keys=True)
            f.write("\n")
        sizes = {k: len(v) for k, v in out.items() if k.startswith("max_cont")}
        print(name, "clusters per max_cont cutoff:", sizes)


if __name__ == "__main__":
    main()
