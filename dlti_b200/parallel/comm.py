"""Process-group bootstrap and the baseline collectives.

One process per GPU; rendezvous from the torchrun/deepspeed-launcher environment
(RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT — the reference scripts read the same
variables: training/train_deepspeed_zero1.py:120-121). Backend: NCCL for CUDA tensors plus gloo
for host-side objects; gloo alone on CPU (the multi-process CPU test tier).

These torch.distributed collectives are the *baseline* path ("nccl" backend). The product path
(`parallel.fused`) does the same data movement from inside sm_100a kernels over peer memory.
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class DistInfo:
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    device: torch.device = torch.device("cpu")

    @property
    def is_main(self) -> bool:
        return self.rank == 0

    @property
    def distributed(self) -> bool:
        return self.world_size > 1


def init_distributed(device: str = "cuda", local_rank: int = -1, timeout_s: int = 1800) -> DistInfo:
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lr = int(os.environ.get("LOCAL_RANK", str(local_rank if local_rank >= 0 else 0)))
    use_cuda = device.startswith("cuda") and torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(lr % max(1, torch.cuda.device_count()))
        dev = torch.device("cuda", torch.cuda.current_device())
    else:
        dev = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = "cpu:gloo,cuda:nccl" if use_cuda else "gloo"
        kw = {"device_id": dev} if use_cuda else {}
        dist.init_process_group(backend=backend, rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=timeout_s), **kw)
    return DistInfo(rank, world, lr, dev)


def shutdown() -> None:
    if dist.is_initialized():
        dist.destroy_process_group()


# ---- collectives with single-process short-circuits and gloo fallbacks -------------------------
def all_reduce_sum_(t: torch.Tensor, info: DistInfo) -> torch.Tensor:
    if info.distributed:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def all_reduce_max_(t: torch.Tensor, info: DistInfo) -> torch.Tensor:
    if info.distributed:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t


def reduce_scatter_sum(inp: torch.Tensor, out: torch.Tensor, info: DistInfo) -> None:
    """out[part] = sum over ranks of inp[rank*part:(rank+1)*part]. inp is [world*part] contiguous."""
    if not info.distributed:
        out.copy_(inp)
        return
    if inp.is_cuda:
        dist.reduce_scatter_tensor(out, inp, op=dist.ReduceOp.SUM)
    else:  # gloo has no reduce_scatter: all-reduce then slice (CPU test tier only)
        tmp = inp.clone()
        dist.all_reduce(tmp, op=dist.ReduceOp.SUM)
        part = out.numel()
        out.copy_(tmp[info.rank * part:(info.rank + 1) * part])


def all_gather_into(out: torch.Tensor, part: torch.Tensor, info: DistInfo) -> None:
    """out[world*part] = concat over ranks of part. `part` may alias its own slot of `out`."""
    if not info.distributed:
        if out.data_ptr() != part.data_ptr():
            out.copy_(part)
        return
    if out.is_cuda:
        dist.all_gather_into_tensor(out, part)
    else:
        chunks = list(out.view(info.world_size, -1).unbind(0))
        dist.all_gather(chunks, part.clone())


def broadcast_(t: torch.Tensor, src: int, info: DistInfo) -> torch.Tensor:
    if info.distributed:
        dist.broadcast(t, src=src)
    return t


def broadcast_object(obj, src: int, info: DistInfo):
    if not info.distributed:
        return obj
    box = [obj]
    dist.broadcast_object_list(box, src=src)
    return box[0]


def all_reduce_host_scalar(v: float, info: DistInfo, op: str = "sum") -> float:
    """Tiny host-side all-reduce (gloo): token counts, flags. No device sync."""
    if not info.distributed:
        return v
    t = torch.tensor([v], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX)
    return float(t.item())


def barrier(info: DistInfo) -> None:
    if info.distributed:
        dist.barrier()
