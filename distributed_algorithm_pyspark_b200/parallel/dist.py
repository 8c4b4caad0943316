"""Process-group plumbing: one process per GPU under ``torchrun`` (NCCL), ``gloo`` on CPU.

Replaces the reference's SparkSession / driver-executor split (``optimization/ssgd.py:78-81``):
there is no driver -- every rank holds the replicated state (``w``, centroids, ranks vector) and
applies identical updates after each collective, which removes Spark's per-iteration
``sc.broadcast`` (``ssgd.py:95``) entirely.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


@dataclass
class Dist:
    rank: int = 0
    world: int = 1
    local_rank: int = 0
    device: torch.device = torch.device("cpu")
    initialized_here: bool = False

    @property
    def is_main(self) -> bool:
        return self.rank == 0

    # -- collectives (no-ops at world == 1) ------------------------------------------------
    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def all_reduce_max(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t

    def all_gather(self, t: torch.Tensor) -> List[torch.Tensor]:
        if self.world == 1:
            return [t]
        out = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(out, t.contiguous())
        return out

    def all_gather_into(self, out: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            out.copy_(t.view_as(out))
        elif self.device.type != "cuda":        # gloo: list form
            parts = [torch.empty_like(t) for _ in range(self.world)]
            dist.all_gather(parts, t.contiguous())
            out.copy_(torch.cat(parts).view_as(out))
        else:
            dist.all_gather_into_tensor(out, t.contiguous())
        return out

    def all_gather_object(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        dist.all_gather_object(out, obj)
        return out

    def all_to_all(self, send: Sequence[torch.Tensor]) -> List[torch.Tensor]:
        """Variable-size all-to-all of 1-D/2-D tensors (sizes exchanged first)."""
        if self.world == 1:
            return [send[0]]
        if self.device.type != "cuda":          # gloo has no all_to_all: exchange through object gather
            objs = self.all_gather_object([t.cpu().numpy() for t in send])
            return [torch.from_numpy(objs[r][self.rank]).to(send[0].device) for r in range(self.world)]
        sizes = torch.tensor([s.shape[0] for s in send], dtype=torch.int64, device=send[0].device)
        rsizes = torch.empty_like(sizes)
        dist.all_to_all_single(rsizes, sizes)
        tail = tuple(send[0].shape[1:])
        recv = [torch.empty((int(n),) + tail, dtype=send[0].dtype, device=send[0].device)
                for n in rsizes.tolist()]
        dist.all_to_all(recv, [s.contiguous() for s in send])
        return recv

    def broadcast(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        if self.world > 1:
            dist.broadcast(t, src)
        return t

    def barrier(self):
        if self.world > 1:
            if self.device.type == "cuda":
                dist.barrier(device_ids=[self.device.index])
            else:
                dist.barrier()

    def max_over_ranks(self, value: float) -> float:
        """Every multi-GPU time is reported as the max over ranks (BASELINE.json north star)."""
        if self.world == 1:
            return float(value)
        t = torch.tensor([value], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, value: float) -> float:
        if self.world == 1:
            return float(value)
        t = torch.tensor([value], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())


_CURRENT: Optional[Dist] = None


def init_distributed(device: Optional[str] = None) -> Dist:
    """Join the torchrun-provided group if WORLD_SIZE > 1, else run single-process.
    Rendezvous always goes through 127.0.0.1 unless MASTER_ADDR says otherwise."""
    global _CURRENT
    if _CURRENT is not None:
        return _CURRENT
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    use_cuda = torch.cuda.is_available() and device != "cpu"
    if use_cuda:
        dev = torch.device("cuda", local_rank % max(torch.cuda.device_count(), 1))
        torch.cuda.set_device(dev)
    else:
        dev = torch.device("cpu")
    here = False
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        kw = {}
        if use_cuda:
            kw["device_id"] = dev
        dist.init_process_group(backend="nccl" if use_cuda else "gloo", rank=rank,
                                world_size=world, **kw)
        here = True
    _CURRENT = Dist(rank=rank, world=world, local_rank=local_rank, device=dev, initialized_here=here)
    return _CURRENT


def current() -> Dist:
    return _CURRENT if _CURRENT is not None else init_distributed()


def shutdown():
    global _CURRENT
    if _CURRENT is not None and _CURRENT.initialized_here and dist.is_initialized():
        dist.destroy_process_group()
    _CURRENT = None
