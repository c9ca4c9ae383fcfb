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


class PeerExchange:
    """The all-gather of the packed slice state as ONE kernel over NVLink peer memory (csrc/peer_exchange.cu).

    Every rank owns a symmetric buffer mapped by all peers (torch.distributed._symmetric_memory); the kernel copies this
    rank's partial into it, publishes a device-side epoch flag, waits for the peers' flags and pulls their partials with
    P2P loads.  No host-side counter, so the layer stays CUDA-graph capturable (a captured NCCL all-gather hung on replay).
    ``PeerExchange.create`` returns None when symmetric memory cannot be set up (CPU / gloo group, no P2P): the caller
    then uses the NCCL all-gather.
    """

    def __init__(self, shape, group, device):
        import ctypes as C
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        l = _lib.lib()
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.n = int(l.haet_partial_floats(C.byref(shape)))
        nbytes = int(l.haet_peer_buffer_bytes(C.byref(shape)))
        self.buf = symm.empty(nbytes // 4, dtype=torch.float32, device=device)
        self.buf.zero_()
        self.handle = symm.rendezvous(self.buf, group.group_name if hasattr(group, "group_name") else group)
        ptrs = list(self.handle.buffer_ptrs)
        if len(ptrs) != self.world:
            raise RuntimeError("symmetric memory returned %d peer pointers for %d ranks" % (len(ptrs), self.world))
        self.peer_ptrs = (C.c_void_p * self.world)(*ptrs)
        self.state = torch.zeros(4, dtype=torch.int32, device=device)
        # all exchanges of this process run on ONE side stream, in launch order: the epoch protocol needs every rank to
        # execute them in the same order, which concurrent caller streams (HostStream) would not guarantee
        self.stream = torch.cuda.Stream(device)
        torch.cuda.synchronize(device)
        dist.barrier(group)   # every rank's buffer is zeroed and mapped before the first exchange

    @classmethod
    def create(cls, shape, group, device) -> Optional["PeerExchange"]:
        try:
            import torch.distributed as dist
            if torch.device(device).type != "cuda" or dist.get_backend(group) != "nccl":
                return None
            return cls(shape, group, device)
        except Exception as ex:   # no P2P / symmetric memory on this system: NCCL all-gather instead
            import warnings
            warnings.warn(f"haet_b200: peer-memory exchange unavailable ({type(ex).__name__}: {ex}); using NCCL all-gather")
            return None

    def __call__(self, shape, partial: torch.Tensor) -> Tuple[torch.Tensor, int]:
        import ctypes as C
        from . import _lib
        cur = torch.cuda.current_stream(partial.device)
        gathered = torch.empty(self.world * self.n, dtype=torch.float32, device=partial.device)
        if torch.cuda.is_current_stream_capturing():
            # inside a CUDA graph the replays are ordered on the replaying stream already
            _lib.check(_lib.lib().haet_peer_exchange(C.byref(shape), self.peer_ptrs, self.rank, self.world,
                                                     partial.data_ptr(), gathered.data_ptr(), self.state.data_ptr(),
                                                     cur.cuda_stream))
            return gathered, self.world
        self.stream.wait_stream(cur)
        partial.record_stream(self.stream)
        gathered.record_stream(self.stream)
        _lib.check(_lib.lib().haet_peer_exchange(C.byref(shape), self.peer_ptrs, self.rank, self.world, partial.data_ptr(),
                                                 gathered.data_ptr(), self.state.data_ptr(), self.stream.cuda_stream))
        cur.wait_stream(self.stream)
        return gathered, self.world

    def error(self) -> bool:
        """True when a wait for a peer timed out (host sync)."""
        return bool(self.state[3].item())


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
