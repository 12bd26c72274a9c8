"""Multi-GPU plumbing for the per-ray path.  One process per GPU; rays shard naturally,
parameters are replicated, and the only collective of a training step is one all-reduce(sum) of the flat
gradient buffer the backward kernels wrote into.  Test-view rendering splits the image by interleaved
rows and needs one all-gather of the finished tiles.  The reference has no multi-GPU code at all.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def init_from_env(backend: str | None = None) -> tuple[int, int]:
    """Join the process group torchrun described; returns (rank, world).  No-op for a single process."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1 and not dist.is_initialized():
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        kw = {}
        if backend == "nccl":
            os.environ.setdefault("NCCL_MIN_CTAS", "64")      # more CTAs for the one large all-reduce of a step
            local = int(os.environ.get("LOCAL_RANK", "0"))
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, **kw)
    return rank, world


def shard_indices(n: int, rank: int, world: int) -> torch.Tensor:
    """Rank-strided slice of a global batch of n rays (every rank gets floor/ceil(n/world) rays)."""
    return torch.arange(rank, n, world)


def grad_scale(n_local: int, n_global: int) -> float:
    """The loss is a mean over the GLOBAL batch (train.py:338): a rank that back-propagates the mean over its
    own n_local rays must scale by n_local / n_global before the sum-reduce."""
    return float(n_local) / float(n_global)


def all_reduce_gradients(flat_grad: torch.Tensor, group=None) -> torch.Tensor:
    """One all-reduce(sum) over the flat fp32 gradient buffer (TensorBase.grad_buffer()[0])."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    return flat_grad


def row_tiles(height: int, rank: int, world: int, block: int = 8) -> torch.Tensor:
    """Interleaved row blocks of an image for tile-split rendering: rank r renders blocks r, r+world, ...
    (tiles through the object carry more surviving samples, so contiguous tiles would be imbalanced)."""
    rows = torch.arange(height)
    return rows[(rows // block) % world == rank]


def gather_rows(local: torch.Tensor, height: int, rank: int, world: int, block: int = 8, group=None) -> torch.Tensor:
    """Reassemble [H, W, C] from every rank's `row_tiles` part ([rows_r, W, C])."""
    if world == 1:
        return local
    counts = [int(row_tiles(height, r, world, block).numel()) for r in range(world)]
    pad = max(counts)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    out = torch.empty((height,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    for r in range(world):
        out[row_tiles(height, r, world, block).to(local.device)] = parts[r][: counts[r]]
    return out
