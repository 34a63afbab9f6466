"""One-process-per-GPU context: rank / world, the control-plane process group, and CUDA-IPC
exchange of shard base pointers so every rank can address every peer's HBM directly.

Control plane = ``torch.distributed`` (NCCL on GPUs for barriers and the all-to-all baseline;
``all_gather_object`` for the 64-byte IPC handles).  Data plane = plain loads / ``red.add`` on
peer-mapped pointers from inside the fused kernels (no collective on the hot path).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional


@dataclass
class DeviceContext:
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    device: object = None
    group: object = None

    @property
    def distributed(self) -> bool:
        return self.world_size > 1

    def barrier(self) -> None:
        if self.distributed:
            import torch.distributed as dist
            dist.barrier(group=self.group)

    def all_gather_object(self, obj) -> list:
        if not self.distributed:
            return [obj]
        import torch.distributed as dist
        out = [None] * self.world_size
        dist.all_gather_object(out, obj, group=self.group)
        return out


_ctx: Optional[DeviceContext] = None


def init_context(backend: Optional[str] = None) -> DeviceContext:
    """Reads RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* (torchrun) and initialises the process
    group when WORLD_SIZE > 1.  Idempotent."""
    global _ctx
    if _ctx is not None:
        return _ctx
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    device = None
    if torch.cuda.is_available():
        torch.cuda.set_device(local % torch.cuda.device_count())
        device = torch.device("cuda", local % torch.cuda.device_count())
    if world > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            be = backend or ("nccl" if device is not None else "gloo")
            kw = {"device_id": device} if be == "nccl" else {}
            dist.init_process_group(be, rank=rank, world_size=world, **kw)
    _ctx = DeviceContext(rank, world, local, device, None)
    return _ctx


def get_context() -> DeviceContext:
    return _ctx if _ctx is not None else init_context()


def reset_context() -> None:
    global _ctx
    _ctx = None
