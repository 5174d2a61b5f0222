"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's incremental colouring loop,
src/sourmash_indexing.cpp:141-277 (the same algorithm as src/index.cpp:189-331).

PARITY UNPINNED: the reference indexer cannot be built in this image (it needs kProcessor's
kDataFrame), so this restatement is checked only against the definition the reference's
own test uses (test/generate_golden_files.py:40-49: shared k-mers of a pair = |mins_a & mins_b|) and,
through the files it leads to, against the unmodified pairwise binary (oracle/_ref/pairwise_ref).

`incremental(sketches)` follows the loop statement by statement (frame = hash -> colour, legend,
tagsMap keyed by the ';'-joined sorted member string, colorsCount, freeColors min-heap, per-signature
convertMap); `end_state(...)` reduces its result to what the pairwise pass can observe.
"""
from __future__ import annotations

import heapq
from typing import Dict, List, Sequence, Tuple


def incremental(sketches: Sequence[Sequence[int]]):
    """Returns (legend: colour -> sorted members, colors_count: colour -> hashes, totals: sample -> mins.size())."""
    n = len(sketches)
    tags: Dict[str, int] = {}
    legend: Dict[int, List[int]] = {}
    colors_count: Dict[int, int] = {}
    group_counter: Dict[int, int] = {}
    group_id = 1
    for g in range(1, n + 1):                      # :100-116 one colour per sample, id = groupID
        tags[str(g)] = g
        legend[g] = [g]
        colors_count[g] = 0
        group_counter[g] = 1
        group_id += 1
    inv_group = set(range(1, n + 1))               # inv_groupNameMap (:126-128)
    free: List[int] = []
    frame: Dict[int, int] = {}                     # hash -> colour; getCount of an absent hash is 0
    totals: Dict[int, int] = {}
    for g, mins in enumerate(sketches, start=1):
        read_tag = g
        convert = {0: read_tag, read_tag: read_tag}   # :198-200
        totals[g] = len(mins)                         # :205
        for h in mins:
            cur = frame.get(h, 0)
            if cur not in convert:                    # :211-243
                colors = list(legend[cur])
                if read_tag not in colors:
                    colors.append(read_tag)
                    colors.sort()
                key = ";".join(str(c) for c in colors)
                if key not in tags:
                    if not free:
                        new = group_id
                        group_id += 1
                    else:
                        new = heapq.heappop(free)
                    tags[key] = new
                    legend[new] = colors
                    colors_count[new] = 0
                convert[cur] = tags[key]
            tgt = convert[cur]
            if tgt != cur:                            # :245-266
                colors_count[cur] = colors_count.get(cur, 0) - 1
                if colors_count[cur] == 0 and cur != 0:
                    if cur not in inv_group:
                        heapq.heappush(free, cur)
                        del tags[";".join(str(c) for c in legend[cur])]
                        del legend[cur]
                        convert.pop(cur, None)
                colors_count[tgt] = colors_count.get(tgt, 0) + 1
            frame[h] = tgt
        group_counter[g] -= 1                         # :275-282
        if colors_count[read_tag] == 0 and group_counter[g] == 0:
            heapq.heappush(free, read_tag)
            legend.pop(read_tag, None)
    return legend, colors_count, totals


def end_state(legend, colors_count) -> Dict[Tuple[int, ...], int]:
    """member set -> number of hashes, over the colours the pairwise pass can see (the legend), merged
    by member set; entries with count 0 are dropped (they add nothing to any pair)."""
    out: Dict[Tuple[int, ...], int] = {}
    for color, members in legend.items():
        c = colors_count.get(color, 0)
        if c:
            key = tuple(sorted(members))
            out[key] = out.get(key, 0) + c
    return out


def by_definition(sketches: Sequence[Sequence[int]]) -> Dict[Tuple[int, ...], int]:
    """The same end state from its definition: colour of a hash = the set of samples containing it."""
    where: Dict[int, List[int]] = {}
    for g, mins in enumerate(sketches, start=1):
        for h in set(mins):
            where.setdefault(h, []).append(g)
    out: Dict[Tuple[int, ...], int] = {}
    for ms in where.values():
        key = tuple(sorted(ms))
        out[key] = out.get(key, 0) + 1
    return out
