"""Drop-in for the reference's top-level ``balltree`` module (models/components/balltree/balltree.pyx).

``build_balltree_with_rotations(data, batch_idx, strides, ball_sizes, angle)`` (balltree.pyx:598-662)
and ``build_balltree(data, batch_idx)`` (:525-550) with the same arguments, return values, dtypes
(int64 indices, bool mask) and error behaviour -- but built on the GPU by ``haet_balltree_build``
with no device->host->device round trip.  Restriction: all trees must have the same
number of points (always true inside the HAET layer: B*heads trees of slice_num tokens).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib


def _uniform_pts(batch_idx: Optional[torch.Tensor], T: int) -> int:
    """Points per tree from a sorted batch index (balltree.pyx:444-472).  One host sync; the HAET layer
    passes pts_per_tree directly and never calls this."""
    if batch_idx is None:
        return T
    if batch_idx.dim() != 1 or batch_idx.shape[0] != T:
        raise ValueError("batch_idx must be a 1D tensor with one entry per point")
    if T == 0:
        raise ValueError("empty input")
    b = batch_idx.detach().to("cpu", torch.int64)
    if int(b[0]) != 0:
        raise ValueError("Batch indices must start from 0")
    nb = int(b[-1]) + 1
    counts = torch.bincount(b, minlength=nb)
    if not bool((b[1:] >= b[:-1]).all()) or not bool((counts == counts[0]).all()):
        raise ValueError("haet_b200.balltree needs sorted batch_idx with equally sized trees")
    return int(counts[0])


def balltree_device(pos: torch.Tensor, n_trees: int, pts: int, strides: Sequence[int], ball_sizes: Sequence[int],
                    angle: float):
    """Raw call: pos [n_trees*pts, d] fp32 CUDA -> (tree_idx int64, tree_mask bool, [rot_idx | None])."""
    assert len(strides) == len(ball_sizes) - 1, "Strides must be one less than ball sizes"
    if not pos.is_cuda:
        raise _lib.HaetError("haet_b200.balltree runs on CUDA tensors only (no CPU fallback)")
    l = _lib.lib()
    d = pos.shape[1]
    slots = l.haet_balltree_slots(n_trees, pts)
    if slots < 0:
        raise ValueError("every tree needs at least 2 points")
    dev = pos.device
    nl = len(ball_sizes)
    tree_idx = torch.empty(slots, dtype=torch.int64, device=dev)
    tree_mask = torch.empty(slots, dtype=torch.bool, device=dev)
    rots: List[Optional[torch.Tensor]] = [None] * nl
    if angle > 0:
        cur = slots
        for i in range(nl):
            rots[i] = torch.empty(cur, dtype=torch.int64, device=dev)
            if i < nl - 1:
                if strides[i] < 1 or cur % strides[i]:
                    raise ValueError(f"stride {strides[i]} does not divide {cur} leaves")
                cur //= strides[i]
    rot_ptrs = (C.c_void_p * nl)(*[(r.data_ptr() if r is not None else None) for r in rots])
    st = (C.c_int32 * nl)(*(list(strides) + [1]))
    bs = (C.c_int32 * nl)(*list(ball_sizes))
    stream = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(l.haet_balltree_build(pos.data_ptr(), n_trees, pts, d, st, bs, nl, float(angle), tree_idx.data_ptr(),
                                     tree_mask.data_ptr(), rot_ptrs, stream))
    return tree_idx, tree_mask, rots


def build_balltree_with_rotations(data: torch.Tensor, batch_idx: torch.Tensor, strides: list, ball_sizes: list,
                                  angle: float = 45.0) -> Tuple[torch.Tensor, torch.Tensor, list]:
    if data.dim() != 2:
        raise ValueError("data must be a 2D tensor")
    pos = data.detach().contiguous().float()
    T = pos.shape[0]
    pts = _uniform_pts(batch_idx, T)
    return balltree_device(pos, T // pts, pts, strides, ball_sizes, angle)


def build_balltree(data: torch.Tensor, batch_idx: torch.Tensor):
    ti, tm, _ = build_balltree_with_rotations(data, batch_idx, [], [1], 0.0)
    return ti, tm
