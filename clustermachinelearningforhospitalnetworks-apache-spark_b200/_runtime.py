"""Device / process-group plumbing shared by the ml classes (torch is only the allocator,
stream and torch.distributed host here)."""
from __future__ import annotations

import torch

from ._native import HkmError


def device() -> torch.device:
    if not torch.cuda.is_available():
        raise HkmError("no CUDA device is visible: the KMeans path has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def dist_info():
    """(rank, world_size, initialised) of the default process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size(), True
    return 0, 1, False


def all_reduce_sum(t: torch.Tensor) -> None:
    """In-place sum over ranks on the current stream (NCCL over NVLink); no-op for one rank."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)


def all_gather_object(obj):
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        out = [None] * dist.get_world_size()
        dist.all_gather_object(out, obj)
        return out
    return [obj]
