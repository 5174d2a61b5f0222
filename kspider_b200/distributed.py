"""One process per GPU (torch.distributed): rank 0 holds the parsed index, the binned CSR blob is
broadcast ONCE (NCCL over NVLink on GPUs; gloo in the CPU tests), every rank adopts it and owns a
disjoint row block of the triangle — so the accumulation needs no reduction.  The only other
exchange is the all-gather of the N cluster labels followed by the K7 merge kernel.
SURVEY.md §8(e)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from .api import Device, Index
from .capi import check, lib


def broadcast_blob(index: Optional[Index], device: torch.device, src: int = 0) -> Tuple[bytes, torch.Tensor]:
    """Broadcasts (descriptor, payload) from `src`.  Returns the descriptor bytes and the payload as a
    uint8 tensor on `device` (device memory when `device` is CUDA: the blob goes host->GPU once on
    `src`, then GPU->GPU)."""
    rank = dist.get_rank()
    if rank == src:
        assert index is not None
        desc = index.descriptor()
        payload = torch.from_numpy(index.payload())  # view of the pinned blob
        head = torch.tensor([len(desc), payload.numel()], dtype=torch.int64)
    else:
        desc, payload = b"", None
        head = torch.zeros(2, dtype=torch.int64)
    head = head.to(device)
    dist.broadcast(head, src)
    n_desc, n_payload = (int(v) for v in head.cpu())
    if rank == src:
        d_desc = torch.frombuffer(bytearray(desc), dtype=torch.uint8).to(device)
        d_payload = payload.to(device, non_blocking=False)
    else:
        d_desc = torch.empty(n_desc, dtype=torch.uint8, device=device)
        d_payload = torch.empty(n_payload, dtype=torch.uint8, device=device)
    dist.broadcast(d_desc, src)
    dist.broadcast(d_payload, src)
    return bytes(d_desc.cpu().numpy().tobytes()), d_payload


def make_shards(index: Index, world: int, host_threads: int = 4):
    """Row-block shards of `index`, one per rank (built on the host, part of loading the index)."""
    return [index.shard(r, world, host_threads) for r in range(world)]


def scatter_shards(shards, device: torch.device, src: int = 0) -> Tuple[bytes, torch.Tensor]:
    """Rank `src` holds the row-block shards (make_shards); every rank receives ITS shard's
    (descriptor, payload): host -> GPU of `src` -> peer GPU over NVLink (NCCL send/recv; gloo in the
    CPU tests).  Unlike broadcast_blob every rank gets only the colours of its own rows, so a step's
    work is proportional to the rank's share of the pair-updates."""
    rank, world = dist.get_rank(), dist.get_world_size()
    head = torch.zeros(2 * world, dtype=torch.int64)
    if rank == src:
        assert shards is not None and len(shards) == world
        descs = [s.descriptor() for s in shards]
        pays = [torch.from_numpy(s.payload()) for s in shards]
        for r in range(world):
            head[2 * r], head[2 * r + 1] = len(descs[r]), pays[r].numel()
    head = head.to(device)
    dist.broadcast(head, src)
    head = head.cpu()
    n_desc, n_payload = int(head[2 * rank]), int(head[2 * rank + 1])
    if rank == src:
        mine = None
        for r in range(world):
            d_desc = torch.frombuffer(bytearray(descs[r]), dtype=torch.uint8).to(device)
            d_pay = pays[r].to(device, non_blocking=False)
            if r == src:
                mine = (descs[r], d_pay)
            else:
                dist.send(d_desc, r)
                dist.send(d_pay, r)
        return mine
    d_desc = torch.empty(n_desc, dtype=torch.uint8, device=device)
    d_payload = torch.empty(n_payload, dtype=torch.uint8, device=device)
    dist.recv(d_desc, src)
    dist.recv(d_payload, src)
    return bytes(d_desc.cpu().numpy().tobytes()), d_payload


def bind_to_gpu_numa(device_index: int):
    """Pins the calling process to the CPUs next to GPU `device_index` (NVML's ideal affinity), so that
    the pinned host buffers it allocates afterwards (first touch) sit on that GPU's NUMA node and the
    per-rank H2D / D2H copies of a multi-GPU run do not cross the socket interconnect.  Returns the
    CPU list, or None when NVML or the topology is unavailable (nothing is changed then)."""
    try:
        import os
        import pynvml
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(device_index).uuid)
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return sorted(os.sched_getaffinity(0))
    except Exception:
        return None


def host_shard(shards, device: torch.device, src: int = 0) -> Index:
    """scatter_shards, then the received shard is kept as an Index in THIS rank's pinned host memory
    (ksb_index_from_blob).  From there every rank runs the host-facing calls (Index.upload /
    Index.pairwise) on its own shard, i.e. N uploads over N PCIe links at once — the multi-GPU
    shape of the reference-facing call, where the one-off distribution is part of loading the index."""
    desc, payload = scatter_shards(shards, device, src)
    return Index.from_blob(desc, payload.cpu().numpy())


def adopt(desc: bytes, d_payload: torch.Tensor, **opts) -> Device:
    """Device handle of this rank's row block over an already-resident blob (CUDA tensor)."""
    assert d_payload.is_cuda and d_payload.dtype == torch.uint8
    return Device.adopt(desc, d_payload.data_ptr(), d_payload.numel(), keepalive=d_payload,
                        device=d_payload.device.index, rank=dist.get_rank(), world=dist.get_world_size(), **opts)


def cluster_merged(result, n_samples: int, dist_type: str, cutoff01: float, device: torch.device, to_host: bool = False):
    """K6 + K7 on this rank's records with the labels left on the GPU, all-gather of every rank's
    labels over NVLink, K7 merge kernel: (global labels as a CUDA int32 tensor -- with to_host a numpy view of a
    pinned buffer that the next call overwrites --, this rank's kept edges).
    No host round trip between the three."""
    world = dist.get_world_size()
    key = (n_samples, world, device.index)
    buf = _LABEL_BUFFERS.get(key)
    if buf is None:  # allocated once per shape: a step must not pay for allocator calls
        buf = (torch.empty(n_samples, dtype=torch.int32, device=device),
               torch.empty(world * n_samples, dtype=torch.int32, device=device),
               torch.empty(n_samples, dtype=torch.int32, device=device),
               torch.empty(n_samples, dtype=torch.int32, pin_memory=True))
        _LABEL_BUFFERS.clear()
        _LABEL_BUFFERS[key] = buf
    mine, allv, out, host = buf
    n_edges = result.cluster_into(mine.data_ptr(), dist_type, cutoff01)
    dist.all_gather_into_tensor(allv, mine)
    stream = torch.cuda.current_stream(device).cuda_stream
    check(lib().ksb_cluster_merge_device(n_samples, world, C.c_void_p(allv.data_ptr()), C.c_void_p(out.data_ptr()),
                                         C.c_void_p(stream)))
    if to_host:
        host.copy_(out, non_blocking=True)
        torch.cuda.current_stream(device).synchronize()
        return host.numpy().view(np.uint32), n_edges
    return out, n_edges


_LABEL_BUFFERS = {}
_STEP_STREAMS = {}


def _label_buffers(n_samples: int, world: int, device: torch.device):
    key = (n_samples, world, device.index)
    buf = _LABEL_BUFFERS.get(key)
    if buf is None:  # allocated once per shape: a step must not pay for allocator calls
        buf = (torch.empty(n_samples, dtype=torch.int32, device=device),
               torch.empty(world * n_samples, dtype=torch.int32, device=device),
               torch.empty(n_samples, dtype=torch.int32, device=device),
               torch.empty(n_samples, dtype=torch.int32, pin_memory=True))
        _LABEL_BUFFERS.clear()
        _LABEL_BUFFERS[key] = buf
    return buf


def pairwise_cluster_step(shard: Index, n_samples: int, dist_type: str, cutoff01: float, device: torch.device, **opts):
    """The multi-GPU END-TO-END cluster-only step: every rank uploads its own row-block shard from its pinned host
    memory (ksb_pairwise_cluster: pipelined H2D, accumulate, fused scan, unions; labels left on the GPU), then the
    label all-gather, the K7 merge kernel and one pinned D2H.  Returns (labels, kept edges, touched pairs, info)."""
    world = dist.get_world_size()
    mine, allv, out, host = _label_buffers(n_samples, world, device)
    _, n_edges, n_rows, info = shard.pairwise_cluster(dist_type, cutoff01, d_labels_ptr=mine.data_ptr(), to_host=False,
                                                      device=device.index, **opts)
    dist.all_gather_into_tensor(allv, mine)
    stream = torch.cuda.current_stream(device).cuda_stream
    check(lib().ksb_cluster_merge_device(n_samples, world, C.c_void_p(allv.data_ptr()), C.c_void_p(out.data_ptr()),
                                         C.c_void_p(stream)))
    host.copy_(out, non_blocking=True)
    torch.cuda.current_stream(device).synchronize()
    return host.numpy().view(np.uint32), n_edges, n_rows, info


def cluster_step(dev: Device, n_samples: int, dist_type: str, cutoff01: float, device: torch.device):
    """The multi-GPU cluster-only step, ONE host synchronisation: ksb_device_cluster_enqueue on this rank's row block
    (one CUDA graph launch: accumulate + fused scan + unions, labels left on the GPU), all-gather of every rank's
    labels over NVLink, K7 merge kernel, one pinned D2H, stream synchronise.  Returns (global labels as a numpy view
    of a pinned buffer the next call overwrites, this rank's kept edges, this rank's touched pairs)."""
    from .capi import KSB_ERR_BUDGET, KsbError
    world = dist.get_world_size()
    mine, allv, out, host = _label_buffers(n_samples, world, device)
    # Everything of the step is ordered on ONE non-default stream: torch's default stream has handle 0, which the C-ABI
    # reads as "the library's own stream" -- the all-gather would then not wait for the labels.
    side = _STEP_STREAMS.get(device.index)
    if side is None:
        side = _STEP_STREAMS[device.index] = torch.cuda.Stream(device)
    stream = side.cuda_stream
    while True:
        with torch.cuda.stream(side):
            dev.cluster_enqueue(dist_type, cutoff01, stream=stream, d_labels_ptr=mine.data_ptr())
            dist.all_gather_into_tensor(allv, mine)
            check(lib().ksb_cluster_merge_device(n_samples, world, C.c_void_p(allv.data_ptr()), C.c_void_p(out.data_ptr()),
                                                 C.c_void_p(stream)))
            host.copy_(out, non_blocking=True)
        side.synchronize()
        # every rank must take the same branch (the all-gather is collective): a rank whose edge list was too short
        # grows it inside collect; all ranks then repeat the step
        short = False
        try:
            _, n_edges, n_rows = dev.cluster_collect()
        except KsbError as e:
            if e.code != KSB_ERR_BUDGET:
                raise
            short = True
            n_edges = n_rows = 0
        if dev._first_steps_left > 0:   # overflow can only happen while the list is still at its initial size
            again = torch.tensor([1 if short else 0], dtype=torch.int32, device=device)
            dist.all_reduce(again)
            dev._first_steps_left -= 1
            if int(again.item()):
                dev._first_steps_left = 2
                continue
        elif short:
            raise RuntimeError("kept-edge list overflow after the warm-up steps (the cut-off changed?): call Device.cluster once first")
        return host.numpy().view(np.uint32), n_edges, n_rows


def merge_labels(labels: np.ndarray, device: torch.device) -> np.ndarray:
    """all-gather of every rank's label array + K7 merge kernel (on CUDA); on a CPU `device` the
    gathered arrays are returned unmerged-by-kernel, merged with a host union-find used only by the
    gloo tests."""
    world = dist.get_world_size()
    mine = torch.from_numpy(labels.astype(np.int32, copy=False).view(np.int32)).to(device)
    allv = torch.empty(world * mine.numel(), dtype=torch.int32, device=device)
    dist.all_gather_into_tensor(allv, mine) if device.type == "cuda" else \
        dist.all_gather(list(allv.view(world, -1).unbind(0)), mine)
    if device.type == "cuda":
        out = torch.empty_like(mine)
        stream = torch.cuda.current_stream(device).cuda_stream
        check(lib().ksb_cluster_merge_device(mine.numel(), world, C.c_void_p(allv.data_ptr()),
                                             C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
        return out.cpu().numpy().view(np.uint32)
    return allv.view(world, -1).numpy().view(np.uint32)
