"""N > 1 host logic on CPU: world_size 2 over gloo (no GPU).  Covers what the multi-GPU path does
around the kernels: the blob broadcast, row-block ownership derived by every rank
from the blob it received, disjointness/completeness of the row blocks, and the label all-gather
+ merge giving the global partition.  The per-block "kernel" work is done by the oracle here."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _union_find_merge(parts: np.ndarray) -> np.ndarray:
    """host merge of [world, N] label arrays (label = smallest member), the CPU stand-in of k7_merge"""
    n = parts.shape[1]
    parent = np.arange(n)

    def find(x):
        while parent[x] != x:
            parent[x] = parent[parent[x]]
            x = parent[x]
        return x
    for row in parts:
        for v, l in enumerate(row.tolist()):
            a, b = find(v), find(l)
            if a != b:
                parent[max(a, b)] = min(a, b)
    return np.array([find(v) for v in range(n)], dtype=np.uint32)


def _worker(rank: int, world: int, port: int, q) -> None:
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        import ctypes as C
        from kspider_b200 import Index
        from kspider_b200 import distributed as kd
        from kspider_b200.capi import check, lib
        from oracle import cluster_oracle
        from oracle.pairwise_oracle import format_float
        from util import canonical_partition, oracle_records, random_csr

        csr = random_csr(seed=21, n_samples=120, n_colors=1500, s_lo=1, s_hi=14, big=(70, 100), id_mode="sparse")
        idx = Index.from_csr(**csr) if rank == 0 else None
        desc, payload = kd.broadcast_blob(idx, torch.device("cpu"), 0)
        assert payload.dtype == torch.uint8 and not payload.is_cuda
        if rank == 0:
            assert desc == idx.descriptor() and np.array_equal(payload.numpy(), idx.payload())
        # every rank derives the same cuts from the blob it received
        cuts = np.zeros(world + 1, dtype=np.uint64)
        pay = np.ascontiguousarray(payload.numpy())
        dbuf = C.create_string_buffer(desc, len(desc))
        check(lib().ksb_blob_row_cuts(dbuf, len(desc), pay.ctypes.data, pay.size, world, cuts.ctypes.data))
        if rank == 0:
            assert np.array_equal(cuts, idx.row_cuts(world))
        n = int(np.unique(np.concatenate([csr["total_ids"], csr["members"]])).size)
        assert cuts[0] == 0 and cuts[-1] == n and np.all(np.diff(cuts.astype(np.int64)) >= 0)
        # this rank's row block = rows whose smaller sample (compact index) lies in [cuts[r], cuts[r+1])
        ids = np.unique(np.concatenate([csr["total_ids"], csr["members"]]))
        want = oracle_records(csr)
        row = np.searchsorted(ids, want["source_1"])
        mine = want[(row >= cuts[rank]) & (row < cuts[rank + 1])]
        # local connected components over the block's kept edges, then all-gather + merge
        cutoff = 0.02
        pos = {int(v): i for i, v in enumerate(ids)}
        keep = lambda r: not (float(format_float(float(r["max_containment"]))) * 100 < cutoff * 100)
        edges = [(pos[int(r["source_1"])], pos[int(r["source_2"])]) for r in mine if keep(r)]
        local = np.asarray(cluster_oracle.components(n, edges), dtype=np.uint32)
        gathered = kd.merge_labels(local, torch.device("cpu"))
        assert gathered.shape == (world, n)
        merged = _union_find_merge(gathered)
        all_edges = [(pos[int(r["source_1"])], pos[int(r["source_2"])]) for r in want if keep(r)]
        ref = cluster_oracle.components(n, all_edges)
        assert canonical_partition(merged) == canonical_partition(np.asarray(ref))
        # blocks are disjoint and complete, and balanced by weighted pair-updates
        cnt = torch.tensor([mine.size, int(mine["shared_kmers"].sum())], dtype=torch.int64)
        dist.all_reduce(cnt)
        assert int(cnt[0]) == want.size and int(cnt[1]) == int(want["shared_kmers"].sum())
        sizes = np.diff(csr["offsets"].astype(np.int64))
        total_upd = int((sizes * (sizes - 1) // 2).sum())
        off = csr["offsets"].astype(np.int64)
        my_upd = 0
        for c in range(sizes.size):
            m = np.sort(np.searchsorted(ids, csr["members"][off[c]:off[c + 1]]))
            p = np.nonzero((m >= cuts[rank]) & (m < cuts[rank + 1]))[0]
            my_upd += int((m.size - 1 - p).sum())
        upd = torch.tensor([my_upd], dtype=torch.int64)
        lst = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(lst, upd)
        assert sum(int(t) for t in lst) == total_upd
        assert max(int(t) for t in lst) <= 0.6 * total_upd  # 2 ranks: each near one half
        # row-block shards: rank 0 cuts the index, every rank receives exactly its own shard
        shards = kd.make_shards(idx, world, host_threads=2) if rank == 0 else None
        sdesc, spay = kd.scatter_shards(shards, torch.device("cpu"), 0)
        sums = torch.zeros(world, dtype=torch.int64)
        if rank == 0:
            for r in range(world):
                sums[r] = int(shards[r].payload().astype(np.int64).sum()) + len(shards[r].descriptor())
            assert sdesc == shards[0].descriptor()
        dist.broadcast(sums, 0)
        assert int(spay.numpy().astype(np.int64).sum()) + len(sdesc) == int(sums[rank])
        scuts = np.zeros(2, dtype=np.uint64)  # a shard's blob still yields row cuts (world 1 = everything)
        sp = np.ascontiguousarray(spay.numpy())
        sbuf = C.create_string_buffer(sdesc, len(sdesc))
        check(lib().ksb_blob_row_cuts(sbuf, len(sdesc), sp.ctypes.data, sp.size, 1, scuts.ctypes.data))
        assert scuts[0] == 0 and scuts[1] == n
        # ... and keeps it as a host index of its own (ksb_index_from_blob): same blob, same shard rows
        mine = kd.host_shard(shards, torch.device("cpu"), 0)
        assert mine.descriptor() == sdesc and np.array_equal(mine.payload(), sp)
        assert mine.info.n_samples == n
        with pytest.raises(Exception):
            mine.shard(0, 2)
        mine.close()
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # surface the failure in the parent
        import traceback
        q.put((rank, traceback.format_exc() + repr(e)))


@pytest.mark.timeout(300)
def test_two_ranks_gloo(built_lib, oracle_lib):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}: {msg}"
