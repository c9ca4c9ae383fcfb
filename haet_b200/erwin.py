"""Host-side mirror of the reference Erwin transformer (models/components/erwinflash/).

The classes below carry the parameters under the reference's names and shapes so that
reference checkpoints load with ``strict=True``:
    ErwinTransformer        erwin_flash.py:15-136
    BasicLayer              components/layers.py:35-70
    ErwinTransformerBlock   components/layers.py:14-28
    BallMSA                 components/attention.py:15-33
    SwiGLU                  components/mlp.py
    BallPooling/Unpooling   components/pooling.py:21-23, 54-56
They hold no torch arithmetic: ``ErwinTransformer.forward`` packs the parameters into one fp32
buffer (order = the walk of csrc/erwin.cu ``Walker``) and calls ``haet_balltree_build`` +
``haet_erwin_forward`` through the C-ABI.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.nn as nn

from . import _lib


class SwiGLU(nn.Module):
    def __init__(self, in_dim: int, dim: int):
        super().__init__()
        self.w1 = nn.Linear(in_dim, dim)
        self.w2 = nn.Linear(in_dim, dim)
        self.w3 = nn.Linear(dim, in_dim)


class BallMSA(nn.Module):
    def __init__(self, dim: int, num_heads: int, ball_size: int, dimensionality: int = 3):
        super().__init__()
        self.num_heads = num_heads
        self.ball_size = ball_size
        self.head_dim = dim // num_heads
        assert self.head_dim * num_heads == dim, "dim must be divisible by num_heads"
        self.qkv = nn.Linear(dim, 3 * dim)
        self.proj = nn.Linear(dim, dim)
        self.pe_proj = nn.Linear(dimensionality, dim)
        # the flash variant never applies the distance bias; ErwinTransformer(variant="erwin") does
        self.sigma_att = nn.Parameter(-1 + 0.01 * torch.randn((1, num_heads, 1, 1)))


class ErwinTransformerBlock(nn.Module):
    def __init__(self, dim: int, num_heads: int, ball_size: int, mlp_ratio: int, dimensionality: int = 3,
                 variant: str = "flash"):
        super().__init__()
        self.variant = variant
        self.ball_size = ball_size
        self.norm1 = nn.RMSNorm(dim)
        self.norm2 = nn.RMSNorm(dim)
        self.BMSA = BallMSA(dim, num_heads, ball_size, dimensionality)
        self.swiglu = SwiGLU(dim, dim * mlp_ratio)

    def tensors(self):
        a, s = self.BMSA, self.swiglu
        qw, qb = a.qkv.weight, a.qkv.bias
        extra = []
        if self.variant == "erwin":
            # the non-flash copy lays the qkv output columns out as (H E K) (benchmarks/02-*/models/components/erwin/
            # components/attention.py:42-48); the kernels read (three, head, head_dim): permute the ROWS of the weight
            H, E = a.num_heads, a.head_dim
            qw = qw.detach().reshape(H, E, 3, -1).permute(2, 0, 1, 3).reshape(3 * H * E, -1)
            qb = qb.detach().reshape(H, E, 3).permute(2, 0, 1).reshape(-1)
            extra = [a.sigma_att.detach().reshape(-1)]   # distance bias sigma_att[h] (attention.py:28-32)
        return [self.norm1.weight, self.norm2.weight, a.pe_proj.weight, a.pe_proj.bias, qw, qb,
                a.proj.weight, a.proj.bias, s.w1.weight, s.w1.bias, s.w2.weight, s.w2.bias, s.w3.weight, s.w3.bias] + extra


class _PoolBase(nn.Module):
    def tensors(self):
        bn = self.norm
        if bn.training or not bn.track_running_stats:
            # train-mode semantics (haet_erwin_cfg.bn_train): the kernels normalise with batch statistics, the slots carry (gamma, beta)
            return [self.proj.weight, self.proj.bias, bn.weight.detach().float(), bn.bias.detach().float()]
        scale = bn.weight.detach().double() / torch.sqrt(bn.running_var.detach().double() + bn.eps)
        shift = bn.bias.detach().double() - bn.running_mean.detach().double() * scale
        return [self.proj.weight, self.proj.bias, scale.float(), shift.float()]


class BallPooling(_PoolBase):
    def __init__(self, in_dim: int, out_dim: int, stride: int, dimensionality: int = 3):
        super().__init__()
        self.stride = stride
        self.proj = nn.Linear(stride * in_dim + stride * dimensionality, out_dim)
        self.norm = nn.BatchNorm1d(out_dim)


class BallUnpooling(_PoolBase):
    def __init__(self, in_dim: int, out_dim: int, stride: int, dimensionality: int = 3):
        super().__init__()
        self.stride = stride
        self.proj = nn.Linear(in_dim + stride * dimensionality, stride * out_dim)
        self.norm = nn.BatchNorm1d(out_dim)


class BasicLayer(nn.Module):
    def __init__(self, direction, depth: int, stride, in_dim: int, out_dim: int, num_heads: int, ball_size: int,
                 mlp_ratio: int, rotate: bool, dimensionality: int = 3, variant: str = "flash"):
        super().__init__()
        hidden_dim = in_dim if direction == "down" else out_dim
        self.blocks = nn.ModuleList([ErwinTransformerBlock(hidden_dim, num_heads, ball_size, mlp_ratio, dimensionality,
                                                           variant) for _ in range(depth)])
        self.rotate = [i % 2 for i in range(depth)] if rotate else [False] * depth
        self.direction = direction
        if direction == "down" and stride is not None:
            self.pool = BallPooling(hidden_dim, out_dim, stride, dimensionality)
        elif direction == "up" and stride is not None:
            self.unpool = BallUnpooling(in_dim, hidden_dim, stride, dimensionality)


class ErwinTransformer(nn.Module):
    """Same constructor as the reference (erwin_flash.py:42-58); ``embed`` must be False and ``mp_steps`` 0,
    the only mode the HAET layer uses (IrregularMesh.py:61-62).

    ``variant`` (extension): "flash" = models/components/erwinflash (no distance bias, qkv columns
    (three, head, head_dim)) -- what the top-level HAET models import; "erwin" = the non-flash copy under
    benchmarks/02-Car-Design-ShapeNetCar/models/components/erwin: ``sigma_att[h] * cdist(pos, pos)`` added to the per-ball
    logits and qkv columns laid out (H E K).  Parameter names are the same in both, checkpoints are not interchangeable."""

    def __init__(self, c_in: int, c_hidden: List, ball_sizes: List, enc_num_heads: List, enc_depths: List,
                 dec_num_heads: List, dec_depths: List, strides: List, rotate: int, decode: bool = True,
                 mlp_ratio: int = 4, dimensionality: int = 3, mp_steps: int = 3, embed: bool = True,
                 variant: str = "flash"):
        super().__init__()
        if variant not in ("flash", "erwin"):
            raise ValueError("variant must be 'flash' (erwinflash) or 'erwin' (distance-bias copy)")
        self.variant = variant
        assert len(enc_num_heads) == len(enc_depths) == len(ball_sizes)
        assert len(dec_num_heads) == len(dec_depths) == len(strides)
        assert len(strides) == len(ball_sizes) - 1
        if embed or mp_steps:
            raise NotImplementedError("haet_b200 implements the HAET hot path only: embed=False, mp_steps=0 "
                                      "(ErwinEmbedding/MPNN are never executed by the HAET layer)")
        if c_in != c_hidden[0]:
            raise ValueError("embed=False needs c_in == c_hidden[0]")
        self.rotate, self.decode = rotate, decode
        self.ball_sizes, self.strides = list(ball_sizes), list(strides)
        self.mp_steps, self.use_embedding = mp_steps, embed
        self.dimensionality, self.mlp_ratio = dimensionality, mlp_ratio
        num_layers = len(enc_depths) - 1
        self.encoder = nn.ModuleList(
            BasicLayer("down", enc_depths[i], strides[i], c_hidden[i], c_hidden[i + 1], enc_num_heads[i], ball_sizes[i],
                       mlp_ratio, rotate > 0, dimensionality, variant) for i in range(num_layers))
        self.bottleneck = BasicLayer(None, enc_depths[-1], None, c_hidden[-1], c_hidden[-1], enc_num_heads[-1],
                                     ball_sizes[-1], mlp_ratio, rotate > 0, dimensionality, variant)
        if decode:
            self.decoder = nn.ModuleList(
                BasicLayer("up", dec_depths[i], strides[i], c_hidden[i + 1], c_hidden[i], dec_num_heads[i], ball_sizes[i],
                           mlp_ratio, rotate > 0, dimensionality, variant) for i in range(num_layers - 1, -1, -1))
        self.in_dim, self.out_dim = c_in, c_hidden[0]
        self.cfg = _lib.make_erwin_cfg(c_hidden, ball_sizes, enc_num_heads, enc_depths, dec_num_heads, dec_depths,
                                       strides, rotate, decode, mlp_ratio, dimensionality, variant == "erwin")
        self._packed: Optional[torch.Tensor] = None
        self._packed_key = None
        self.apply(self._init_weights)

    @staticmethod
    def _init_weights(m):   # erwin_flash.py:126-136
        if isinstance(m, nn.Linear):
            nn.init.trunc_normal_(m.weight, mean=0.0, std=0.02, a=-2.0, b=2.0)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)

    # ---- packing --------------------------------------------------------------------------
    def _tensor_walk(self):
        out = []
        for layer in self.encoder:
            for blk in layer.blocks:
                out += blk.tensors()
            if layer.pool.stride != 1:
                out += layer.pool.tensors()
        for blk in self.bottleneck.blocks:
            out += blk.tensors()
        if self.decode:
            for layer in self.decoder:
                out += layer.unpool.tensors()
                for blk in layer.blocks:
                    out += blk.tensors()
        return out

    def _state_key(self, device):
        return (str(device), self.training,
                tuple((t.data_ptr(), t._version) for t in list(self.parameters()) + list(self.buffers())))

    def packed_weights(self, device) -> torch.Tensor:
        """One flat fp32 device buffer, every tensor padded to a multiple of 4 floats (csrc/erwin.cu Walker).
        Rebuilt only when a parameter/buffer changed (version counters)."""
        key = self._state_key(device)
        if self._packed is not None and self._packed_key == key:
            return self._packed
        chunks = []
        for t in self._tensor_walk():
            f = t.detach().reshape(-1).to(device=device, dtype=torch.float32)
            pad = (-f.numel()) % 4
            chunks.append(f if pad == 0 else torch.cat([f, f.new_zeros(pad)]))
        packed = torch.cat(chunks).contiguous()
        want = _lib.lib().haet_erwin_weight_floats(C.byref(self.cfg))
        if packed.numel() != want:
            raise _lib.HaetError(f"packed Erwin weights have {packed.numel()} floats, library expects {want}")
        self._packed, self._packed_key = packed, key
        return packed

    # ---- train-mode BatchNorm (forward only) ------------------------------------------------
    def batch_norms(self):
        """The BatchNorm1d modules in the order the kernels run them: encoder pools (stride != 1), then decoder unpools."""
        out = [layer.pool.norm for layer in self.encoder if layer.pool.stride != 1]
        if self.decode:
            out += [layer.unpool.norm for layer in self.decoder]
        return out

    @torch.no_grad()
    def update_running_stats(self, norms, stats: torch.Tensor, n_slots: int):
        """nn.BatchNorm1d's train-mode buffer update from the batch (mean, biased var) the kernels report: running_var takes the
        UNBIASED variance, momentum None means a cumulative average (torch/nn/modules/batchnorm.py)."""
        rows, n = [], n_slots
        for i, layer in enumerate(self.encoder):          # rows entering each BatchNorm
            n //= self.strides[i]
            if layer.pool.stride != 1:
                rows.append(n)
        if self.decode:
            for i in range(len(self.strides) - 1, -1, -1):
                rows.append(n * self.strides[i])
                n *= self.strides[i]
        off = 0
        for bn, cnt in zip(norms, rows):
            Cn = bn.num_features
            mean, var = stats[off:off + Cn], stats[off + Cn:off + 2 * Cn]
            off += 2 * Cn
            if not bn.track_running_stats or bn.running_mean is None:
                continue
            bn.num_batches_tracked += 1
            f = 1.0 / float(bn.num_batches_tracked) if bn.momentum is None else float(bn.momentum)
            unbiased = var * (cnt / max(cnt - 1, 1))
            bn.running_mean.mul_(1 - f).add_(mean.to(bn.running_mean.dtype), alpha=f)
            bn.running_var.mul_(1 - f).add_(unbiased.to(bn.running_var.dtype), alpha=f)

    # ---- forward --------------------------------------------------------------------------
    def forward(self, node_features: torch.Tensor, node_positions: torch.Tensor, batch_idx: torch.Tensor = None,
                pts_per_tree: Optional[int] = None, **kwargs):
        """erwin_flash.py:138-195.  Trees must all have the same number of points (the HAET layer builds
        B*heads trees of slice_num tokens each); pass ``pts_per_tree`` to skip deriving it from batch_idx."""
        # train mode is FORWARD ONLY (partial): BatchNorm1d normalises with batch statistics and updates its
        # running buffers like torch; there is no backward pass, the result carries no grad_fn
        from .balltree import _uniform_pts, balltree_device
        x = node_features.contiguous().float()
        pos = node_positions.contiguous().float()
        T = x.shape[0]
        pts = pts_per_tree if pts_per_tree is not None else _uniform_pts(batch_idx, T)
        tree_idx, tree_mask, rots = balltree_device(pos, T // pts, pts, self.strides, self.ball_sizes, float(self.rotate))
        return erwin_device(self, x, pos, tree_idx, tree_mask, rots, T // pts)


def erwin_device(erwin: ErwinTransformer, x, pos, tree_idx, tree_mask, rots, n_trees: int = 1):
    """haet_erwin_forward_trees: with the tree count the whole pass is one cooperative kernel with per-tree barriers."""
    l = _lib.lib()
    dev = x.device
    w = erwin.packed_weights(dev)
    n_slots = tree_idx.numel()
    ws_bytes = l.haet_erwin_workspace_bytes(C.byref(erwin.cfg), n_slots)
    ws = torch.empty(max(ws_bytes, 4) // 4, dtype=torch.float32, device=dev)
    out = torch.empty_like(x)
    cfg, stats = erwin.cfg, None
    norms = erwin.batch_norms()
    batch = [bn.training or not bn.track_running_stats for bn in norms]
    if any(batch) and not all(batch):
        raise NotImplementedError("BatchNorm layers in mixed train / eval state are not supported")
    if not any(batch):
        norms = []
    if norms:
        if len({float(bn.eps) for bn in norms}) != 1:
            raise NotImplementedError("train-mode BatchNorm: all pooling layers must share one eps")
        cfg = _lib.ErwinCfg.from_buffer_copy(erwin.cfg)
        stats = torch.zeros(2 * sum(bn.num_features for bn in norms), dtype=torch.float32, device=dev)
        cfg.bn_train, cfg.bn_eps, cfg.bn_stats_d = 1, float(norms[0].eps), stats.data_ptr()
    nl = erwin.cfg.n_levels
    rot_ptrs = (C.c_void_p * nl)(*[(r.data_ptr() if r is not None else None) for r in rots])
    st = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(l.haet_erwin_forward_trees(C.byref(cfg), x.data_ptr(), pos.data_ptr(), x.shape[0], max(int(n_trees), 1),
                                          tree_idx.data_ptr(), tree_mask.data_ptr(), n_slots, rot_ptrs, w.data_ptr(),
                                          ws.data_ptr(), out.data_ptr(), st))
    if norms:
        erwin.update_running_stats(norms, stats, n_slots)
    return out
