"""Multi-GPU plumbing: one process per GPU, batch sharded on dim 0, NO collective on the forward path.

Replaces torch.nn.DataParallel as the reference uses it (ReSIDE/train.py:93-98, Benchmark/reside_model.py:47-49),
which re-broadcasts every weight each forward.  Here every rank holds a resident replica; the only exchange
is ONE all-reduce(SUM) of the 16 fp64 loss/metric partial sums (128 bytes) over NCCL/NVLink (gloo on CPU).
"""
import os
from typing import Tuple

import torch
import torch.distributed as dist


def init_from_env(backend: str = None) -> Tuple[int, int, int]:
    """(rank, world_size, local_rank) from the torchrun environment; initialises the process group when
    WORLD_SIZE > 1 (NCCL when CUDA is available, else gloo)."""
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, world, local


def shard_bounds(global_batch: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous split of the batch axis; the first (global_batch % world) ranks get one extra image."""
    base, extra = divmod(global_batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(x: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    lo, hi = shard_bounds(x.shape[0], rank, world)
    return x[lo:hi]


def all_reduce_sums(sums: torch.Tensor) -> torch.Tensor:
    """Sum the 16-double partial-sum vector over all ranks (no-op for a single process)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    return sums


def reduce_max(value: float, device=None) -> float:
    """Max over ranks of a host scalar (timings are reported as the slowest rank)."""
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier():
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
