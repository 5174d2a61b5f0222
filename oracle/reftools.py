"""TEST INFRASTRUCTURE ONLY — drivers for the prebuilt reference binaries under oracle/_ref/.

oracle/_ref/pairwise_ref     = the UNMODIFIED /root/reference/src/pairwise.cpp (see oracle/Makefile)
oracle/_ref/index_writer_ref = writes the 3 index files with the reference's vendored phmap

Nothing here is imported by the product (kspider_b200/); only tests/, bench.py's reference /
cpu_baseline legs and __graft_entry__ may use it.  Nothing here reads /root/reference at run time.
"""
from __future__ import annotations

import os
import re
import struct
import subprocess
from typing import Dict, Iterable, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
PAIRWISE_REF = os.path.join(REF_DIR, "pairwise_ref")
INDEX_WRITER_REF = os.path.join(REF_DIR, "index_writer_ref")
RAW_MAGIC = 0x313057415242534B  # "KSBRAW01"


def have_ref() -> bool:
    return os.access(PAIRWISE_REF, os.X_OK) and os.access(INDEX_WRITER_REF, os.X_OK)


def write_raw(path: str,
              totals: Dict[int, int],
              counts: Dict[int, int],
              colors: Iterable[Tuple[int, Sequence[int]]]) -> None:
    """Serialise an index description in the KSBRAW01 layout (see ref_tools/index_writer_ref.cpp)."""
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", RAW_MAGIC))
        f.write(struct.pack("<Q", len(totals)))
        for k, v in totals.items():
            f.write(struct.pack("<II", k & 0xFFFFFFFF, v & 0xFFFFFFFF))
        f.write(struct.pack("<Q", len(counts)))
        for k, v in counts.items():
            f.write(struct.pack("<QQ", k, v))
        colors = list(colors)
        f.write(struct.pack("<Q", len(colors)))
        for cid, members in colors:
            m = np.asarray(members, dtype="<u4")
            f.write(struct.pack("<QI", cid, m.size))
            f.write(m.tobytes())


def write_raw_arrays(path: str, total_ids: np.ndarray, total_vals: np.ndarray,
                     count_ids: np.ndarray, count_vals: np.ndarray,
                     color_ids: np.ndarray, offsets: np.ndarray, members: np.ndarray) -> None:
    """Vectorised KSBRAW01 writer for large synthetic indexes (CSR form)."""
    n_col = int(color_ids.size)
    offsets = np.asarray(offsets).astype(np.int64)
    sizes = np.diff(offsets).astype("<u4")
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", RAW_MAGIC))
        f.write(struct.pack("<Q", int(total_ids.size)))
        tv = np.empty((total_ids.size, 2), dtype="<u4")
        tv[:, 0] = total_ids
        tv[:, 1] = total_vals
        f.write(tv.tobytes())
        f.write(struct.pack("<Q", int(count_ids.size)))
        cv = np.empty((count_ids.size, 2), dtype="<u8")
        cv[:, 0] = count_ids
        cv[:, 1] = count_vals
        f.write(cv.tobytes())
        f.write(struct.pack("<Q", n_col))
        # record = u64 id | u32 s | s x u32  -> build as one u32 stream
        rec_words = 3 + sizes.astype(np.int64)
        starts = np.zeros(n_col + 1, dtype=np.int64)
        np.cumsum(rec_words, out=starts[1:])
        out = np.empty(int(starts[-1]), dtype="<u4")
        cid = color_ids.astype("<u8")
        out[starts[:-1]] = (cid & 0xFFFFFFFF).astype("<u4")
        out[starts[:-1] + 1] = (cid >> 32).astype("<u4")
        out[starts[:-1] + 2] = sizes
        # member positions
        mpos = np.arange(members.size, dtype=np.int64) + np.repeat(starts[:-1] + 3 - offsets[:-1], sizes)
        out[mpos] = members.astype("<u4")
        f.write(out.tobytes())


def run_index_writer_ref(raw_path: str, prefix: str) -> None:
    subprocess.run([INDEX_WRITER_REF, raw_path, prefix], check=True)


_T_RE = re.compile(r"pairwise hashmap construction: ([0-9.eE+-]+) secs")


def run_pairwise_ref(prefix: str, threads: int = 1, timeout: float | None = None) -> dict:
    """Run the reference pass; returns its stdout and its own accumulation timer
    (src/pairwise.cpp:239 'pairwise hashmap construction')."""
    env = dict(os.environ)
    p = subprocess.run([PAIRWISE_REF, prefix, str(int(threads))], check=True, timeout=timeout,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env)
    m = _T_RE.search(p.stdout)
    return {"stdout": p.stdout, "construct_secs": float(m.group(1)) if m else None}


def read_pairwise_tsv_sorted(path: str) -> Tuple[str, list]:
    """Header + data lines sorted by (source_1, source_2) numerically — the reference's row
    order is hash-map order and unspecified."""
    with open(path, "r") as f:
        header = f.readline()
        rows = f.read().splitlines()
    rows.sort(key=lambda r: tuple(int(x) for x in r.split("\t")[:2]))
    return header, rows
