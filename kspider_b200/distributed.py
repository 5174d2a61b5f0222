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


def adopt(desc: bytes, d_payload: torch.Tensor, **opts) -> Device:
    """Device handle of this rank's row block over an already-resident blob (CUDA tensor)."""
    assert d_payload.is_cuda and d_payload.dtype == torch.uint8
    return Device.adopt(desc, d_payload.data_ptr(), d_payload.numel(), keepalive=d_payload,
                        device=d_payload.device.index, rank=dist.get_rank(), world=dist.get_world_size(), **opts)


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
