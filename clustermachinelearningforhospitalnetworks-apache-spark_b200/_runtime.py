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


class PeerExchange:
    """Per-process-group NVLink exchange buffers for hkm_lloyd_step_peer.

    torch's symmetric memory does the allocation and the peer mapping (plumbing); the data movement itself is
    plain peer stores / loads issued by the Lloyd kernel's last CTA."""

    _cached = None

    def __init__(self, dev: torch.device):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        from . import _native as N
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        nbytes = int(N.load().hkm_peer_exchange_bytes(self.world))
        self.buf = symm.empty((nbytes + 7) // 8, dtype=torch.int64, device=dev)
        self.buf.zero_()
        self.handle = symm.rendezvous(self.buf, dist.group.WORLD)
        self.ptrs_dev = int(self.handle.buffer_ptrs_dev)
        self.err = torch.zeros(1, dtype=torch.int32, device=dev)
        self.epoch = 0
        torch.cuda.synchronize(dev)
        dist.barrier()          # every rank has zeroed its buffer before anyone writes into it

    @classmethod
    def get(cls, dev: torch.device):
        """The shared instance, or None when peer memory is unavailable (then the NCCL path is used)."""
        import os
        if os.environ.get("HKM_NO_PEER_EXCHANGE") == "1":
            return None
        if cls._cached is None:
            try:
                cls._cached = cls(dev)
            except Exception as e:  # noqa: BLE001  (no symmetric memory on this platform / build)
                cls._cached = False
                import warnings
                warnings.warn(f"NVLink peer exchange unavailable ({type(e).__name__}: {e}); using NCCL all-reduce")
            # every rank must take the same path: one rank on NCCL while the others spin on peer flags would hang
            import torch.distributed as dist
            ok = torch.tensor([1 if cls._cached else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok.item()) == 0:
                cls._cached = False
        return cls._cached or None

    def fits(self, k: int, d: int) -> bool:
        """Whether the k*d + k + 1 partial sums fit one exchange slot (else the NCCL all-reduce path is used)."""
        from . import _native as N
        return k * d + k + 1 <= int(N.load().hkm_peer_exchange_slot_doubles())

    def check(self):
        """Raises on EVERY rank if any rank timed out in an exchange (the flag is summed over the ranks first, so no rank
        goes on into a later collective alone), and clears the flag so the next fit starts clean."""
        import torch.distributed as dist
        flag = self.err.clone()
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        if int(flag.item()) != 0:
            self.err.zero_()
            raise HkmError("hkm_lloyd_step_peer: a peer rank never arrived at the exchange (time-out)")
