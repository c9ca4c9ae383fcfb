"""Point-sharded multi-GPU layout (new design; the reference is single-GPU -- SURVEY 8e).

One process per GPU.  Each rank holds a contiguous range of the mesh points (all batch items, all
channels); weights are replicated.  Per layer the only data-path collective is an all-gather of the
packed rank-partial slice state ``[B*heads*G, C+2]`` (<= 557 KB per rank at 8 heads x 512 slices x 32
channels), after which every rank runs the same rank-ordered log-sum-exp merge, the same (replicated,
deterministic) ball-tree build and Erwin pass, and deslices its own points.  LayerNorm / MLP of the
surrounding block are row-local, so a whole HAETransolver stays sharded end to end.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn


def shard_range(n_points: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced point range of ``rank``: sizes differ by at most one, every rank gets >= 1 point
    when n_points >= world."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    lo = (n_points * rank) // world
    hi = (n_points * (rank + 1)) // world
    return lo, hi


def shard_points(x: torch.Tensor, rank: int, world: int, dim: int = 1) -> torch.Tensor:
    lo, hi = shard_range(x.shape[dim], rank, world)
    return x.narrow(dim, lo, hi - lo).contiguous()


def gather_partials(partial: torch.Tensor, group=None) -> Tuple[torch.Tensor, int]:
    """All-gather the packed rank-partials, rank-major (what haet_slice_merge expects)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    if world == 1:
        return partial, 1
    out = torch.empty(world * partial.numel(), dtype=partial.dtype, device=partial.device)
    dist.all_gather_into_tensor(out, partial.contiguous(), group=group)
    return out, world


def set_point_group(model: nn.Module, group) -> nn.Module:
    """Mark every haet_b200 PhysicsAttention inside ``model`` as point-sharded over ``group``."""
    from .physics_attention import _PhysicsAttentionBase
    for mod in model.modules():
        if isinstance(mod, _PhysicsAttentionBase):
            mod.point_group = group
    return model


def init_from_env(backend: Optional[str] = None):
    """torchrun plumbing: RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* from the environment."""
    import os
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    elif torch.cuda.is_available():
        torch.cuda.set_device(local)
    return rank, local, world
