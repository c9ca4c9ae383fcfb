"""Point-sharded multi-GPU layout (new design; the reference is single-GPU -- SURVEY 8e).

One process per GPU.  Each rank holds a contiguous range of the mesh points (all batch items, all
channels); weights are replicated.  Per layer the only data-path collective is an all-gather of the
packed rank-partial slice state ``[B*heads*G, C+2]`` (<= 557 KB per rank at 8 heads x 512 slices x 32
channels), after which every rank runs the same rank-ordered log-sum-exp merge, the same (replicated,
deterministic) ball-tree build and Erwin pass, and deslices its own points.  LayerNorm / MLP of the
surrounding block are row-local, so a whole HAETransolver stays sharded end to end.

Structured meshes (Conv2d / Conv3d front-ends) shard by SLABS of grid planes: the same flow, preceded by one halo exchange
of kernel // 2 planes with each neighbour (exchange_halo -> haet_slice_reduce_halo).
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


def shard_planes(n_planes: int, rank: int, world: int) -> Tuple[int, int]:
    """Slab sharding of a structured mesh: rank owns the planes [lo, hi) of the first grid axis (balanced like shard_range)."""
    return shard_range(n_planes, rank, world)


def shard_grid(x: torch.Tensor, grid, rank: int, world: int) -> torch.Tensor:
    """The rows of ``x [B, H*W(*D), ...]`` that belong to ``rank``'s slab of planes (row-major grid: a contiguous row range)."""
    plane = 1
    for g in grid[1:]:
        plane *= int(g)
    lo, hi = shard_planes(int(grid[0]), rank, world)
    return x.narrow(1, lo * plane, (hi - lo) * plane).contiguous()


def exchange_halo(x: torch.Tensor, plane: int, r: int, group=None):
    """Halo exchange of the conv front-ends of the structured meshes.

    ``x [B, n_local, dim]`` holds this rank's slab of ``n_local / plane`` grid planes.  Returns ``(x_ext, halo_lo, halo_hi)``:
    the slab preceded by the last ``r`` planes of rank - 1 and followed by the first ``r`` planes of rank + 1 (``r = kernel // 2``;
    nothing at the outer boundaries, where the conv's zero padding is the true one) -- the input of haet_slice_reduce_halo.
    One send and one receive per neighbour (NCCL point-to-point on GPUs, gloo on CPU)."""
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    B, n, dim = x.shape
    if n % plane or n // plane < r:
        raise ValueError(f"a slab must hold whole grid planes and at least kernel // 2 = {r} of them (got {n} rows, {plane} per plane)")
    lo = r if rank > 0 else 0
    hi = r if rank < world - 1 else 0
    ext = x.new_empty(B, n + (lo + hi) * plane, dim)
    ext[:, lo * plane:lo * plane + n] = x
    if world == 1 or r == 0:
        return ext, lo, hi
    peer = (lambda k: dist.get_global_rank(group, k)) if group is not None else (lambda k: k)
    ops, recv_lo, recv_hi, keep = [], None, None, []
    if rank > 0:
        first = x[:, :r * plane].contiguous()
        recv_lo = x.new_empty(B, r * plane, dim)
        keep.append(first)
        ops += [dist.P2POp(dist.isend, first, peer(rank - 1), group), dist.P2POp(dist.irecv, recv_lo, peer(rank - 1), group)]
    if rank < world - 1:
        last = x[:, n - r * plane:].contiguous()
        recv_hi = x.new_empty(B, r * plane, dim)
        keep.append(last)
        ops += [dist.P2POp(dist.isend, last, peer(rank + 1), group), dist.P2POp(dist.irecv, recv_hi, peer(rank + 1), group)]
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    if recv_lo is not None:
        ext[:, :lo * plane] = recv_lo
    if recv_hi is not None:
        ext[:, lo * plane + n:] = recv_hi
    return ext, lo, hi


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
