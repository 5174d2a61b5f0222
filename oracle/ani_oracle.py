"""TEST INFRASTRUCTURE ONLY — CPU restatement of `kSpider pairwise --estimate-ani`
(pykSpider/kSpider2/ks_pairwise.py:29-82).

Parity status: the text handling (which columns, float() of the printed values, mean, f-string repr,
header, row order) is PINNED: tests/test_cluster_pinning.py compares this restatement verbatim with the
ani_col files the UNMODIFIED ks_pairwise.py wrote through oracle/refpy.py (tests/golden/*.ani_k*.tsv).
The formula itself stays "parity unpinned": the arithmetic lives in a third-party dependency that is absent
from /root/reference and from this image: `sourmash.distance_utils.containment_to_distance`
(imported at ks_pairwise.py:50; sourmash is not listed in setup.py:203-210 nor requirements.txt, so
no version is pinned).  Its published algorithm (sourmash >= 4.4, distance_utils.py) is restated
here for the call the reference makes — containment_to_distance(containment, ksize, scaled,
n_unique_kmers=...) with the defaults confidence=0.95, estimate_ci=False, prob_threshold=1e-3 —
and for the one attribute it reads, `.ani`:

    containment == 0  -> point_estimate = 1.0
    containment == 1  -> point_estimate = 0.0
    otherwise         -> point_estimate = 1.0 - containment ** (1.0 / ksize)
    ANIResult.__post_init__: not (0 <= dist <= 1) -> ValueError
    ANIResult.ani = 1 - dist      (size_is_inaccurate is False on this path)

`scaled` and `n_unique_kmers` only enter p_nothing_in_common, which `.ani` does not depend on.  None
of the reference's tests touch this option.  The text handling around it (ks_pairwise.py:64-80:
float() of the printed columns, mean of the two values, f-string repr) is plain Python and is
followed exactly.
"""
from __future__ import annotations

from typing import Iterable, List

import numpy as np


def containment_to_ani(containment: float, ksize: int) -> float:
    """sourmash containment_to_distance(...).ani (see the module docstring)."""
    if containment == 0:
        dist = 1.0
    elif containment == 1:
        dist = 0.0
    else:
        dist = 1.0 - containment ** (1.0 / ksize)
    if not (0 <= dist <= 1):
        raise ValueError(f"Error: distance value {dist :.4f} and p_threshold value 0.0010 must both be between 0 and 1.")
    return 1 - dist


def avg_ani_of_text(min_text: str, max_text: str, ksize: int) -> float:
    """ks_pairwise.py:70-78 for one TSV row (columns 3 and 5 as printed)."""
    ani_1_in_2 = containment_to_ani(float(min_text), ksize)
    ani_2_in_1 = containment_to_ani(float(max_text), ksize)
    return (ani_1_in_2 + ani_2_in_1) / 2.0


def ani_column_of_tsv(pairwise_tsv: str, ksize: int) -> List[float]:
    """ks_pairwise.py:64-80: one avg_ani per row of the pairwise TSV, in file order."""
    out: List[float] = []
    with open(pairwise_tsv) as f:
        next(f)
        for line in f:
            col = line.strip().split("\t")
            out.append(avg_ani_of_text(col[3], col[5], ksize))
    return out


def ani_column_of_records(records: np.ndarray, ksize: int) -> List[float]:
    """Same, from float32 containments: the text the reference writes is '%g' of the float
    (src/pairwise.cpp:266-273, ostream<<float)."""
    return [avg_ani_of_text("%g" % float(a), "%g" % float(b), ksize)
            for a, b in zip(records["min_containment"], records["max_containment"])]


def ani_column_text(values: Iterable[float]) -> str:
    """ks_pairwise.py:66,79-80: the column file."""
    return "avg_ani\n" + "".join(f"{v}\n" for v in values)
