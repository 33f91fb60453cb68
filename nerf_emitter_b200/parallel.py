"""Multi-GPU plumbing of the emitter path: one process per GPU, rays sharded, weights replicated.

Forward / forward-mode need no exchange (rays are independent).  The backward has exactly one collective: a
sum all-reduce of the parameter-gradient buffer (what DDP does in the reference,
NS/pipelines/base_pipeline.py:355-357 with the no_sync accumulation of NS/pipelines/mitsuba_sdf.py:1179-1190).
``NerfEmitter.bind_grads()`` lays all gradient tensors out as views of ONE flat fp32 buffer, so this is a
single NCCL call over NVLink (19.4 M floats = 77.6 MB).

The single-producer mode of the reference (rank 0 owns all rays and scatters / gathers them,
DS/emitters/nerf_op.py:92-166) is kept as ``scatter_rays`` / ``gather_rows`` for drop-in use; it sends exact
shard sizes with batched point-to-point ops instead of zero-padding every shard to ceil(N/W).
Works with the ``nccl`` backend on GPUs and ``gloo`` on CPU tensors (tests).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_sizes(total: int, world_size: int) -> List[int]:
    """get_target_sizes (DS/emitters/nerf_op.py:92-95): divmod-balanced contiguous shards."""
    cnt, rem = divmod(int(total), int(world_size))
    return [cnt + 1 if i < rem else cnt for i in range(world_size)]


def shard_range(total: int, rank: int, world_size: int) -> Tuple[int, int]:
    sizes = shard_sizes(total, world_size)
    start = sum(sizes[:rank])
    return start, start + sizes[rank]


def allreduce_grads(flat_grads: torch.Tensor, group=None, scale: Optional[float] = None) -> torch.Tensor:
    """Sum a gradient buffer over ranks with torch.distributed (any backend).  The GPU path does NOT use this:
    ``NerfEmitter.allreduce_grads`` goes through the library (nemit_allreduce_grads: NCCL with the scale folded into
    the collective).  This torch-level twin serves host-side tests on the gloo backend and CPU tensors."""
    _, w = world()
    if w > 1:
        dist.all_reduce(flat_grads, op=dist.ReduceOp.SUM, group=group)
    if scale is not None and scale != 1.0:
        flat_grads.mul_(scale)
    return flat_grads


def scatter_rays(payload: Optional[torch.Tensor], total: int, width: int, device, src: int = 0,
                 dtype=torch.float32) -> torch.Tensor:
    """pad_scatter (DS/emitters/nerf_op.py:124-153) without the padding: rank ``src`` holds ``payload``
    (total, width); every rank returns its own (shard, width) rows."""
    rank, w = world()
    if w == 1:
        return payload
    sizes = shard_sizes(total, w)
    recv = torch.empty(sizes[rank], width, device=device, dtype=dtype)
    ops = []
    if rank == src:
        chunks = list(torch.split(payload, sizes, dim=0))
        recv.copy_(chunks[src])
        for r in range(w):
            if r != src and sizes[r] > 0:
                ops.append(dist.P2POp(dist.isend, chunks[r].contiguous(), r))
    elif sizes[rank] > 0:
        ops.append(dist.P2POp(dist.irecv, recv, src))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return recv


def gather_rows(local: torch.Tensor, total: int, dst: int = 0) -> Optional[torch.Tensor]:
    """pad_gather (DS/emitters/nerf_op.py:98-121) without the padding: concatenates the ranks' rows on ``dst``
    (returns None elsewhere)."""
    rank, w = world()
    if w == 1:
        return local
    sizes = shard_sizes(total, w)
    out = None
    ops = []
    if rank == dst:
        out = torch.empty(total, *local.shape[1:], device=local.device, dtype=local.dtype)
        start = 0
        for r in range(w):
            view = out[start:start + sizes[r]]
            if r == dst:
                view.copy_(local)
            elif sizes[r] > 0:
                ops.append(dist.P2POp(dist.irecv, view, r))
            start += sizes[r]
    elif sizes[rank] > 0:
        ops.append(dist.P2POp(dist.isend, local.contiguous(), dst))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return out
