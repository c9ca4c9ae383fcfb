"""Host-side mirrors of the reference model constructors and the fused ``Model.forward``.

    HAETransolver_Irregular_Mesh      models/HAETransolver_Irregular_Mesh.py:136-376
    HAETransolver_Structured_Mesh_2D  models/HAETransolver_Structured_Mesh_2D.py (forward :384-410)
    HAETransolver_Structured_Mesh_3D  models/HAETransolver_Structured_Mesh_3D.py (forward :418-461)

Same constructor arguments, attribute names, parameter names and shapes (``load_state_dict(strict=True)`` both ways) and
``forward(x, fx, T=None)``.  The classes hold parameters; the arithmetic runs in libhaet_b200.so:

    prologue   preprocess MLP + placeholder + time embedding      haet_model_prologue / haet_time_embed
    blocks     fx = Attn(ln_1(fx)) + fx; fx = mlp(ln_2(fx)) + fx   haet_slice_reduce_ln .. haet_deslice_out_res, haet_block_mlp
    epilogue   last block: mlp2(ln_3(fx))                          haet_head_ln_linear

``fused_model_forward`` works on any module with the reference's attribute names, so ``patch_model(reference_model,
fuse_blocks=True)`` gives a reference-constructed model the same path.  Forward / eval only; CUDA tensors only.  Shapes the
fused kernels do not cover (n_hidden != 256, activations other than GELU, ...) run the same modules as torch ops on the GPU.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .physics_attention import (Physics_Attention_Irregular_Mesh, Physics_Attention_Structured_Mesh_2D,
                                Physics_Attention_Structured_Mesh_3D, _PhysicsAttentionBase, block_mlp)

_ACT = {"gelu": nn.GELU, "tanh": nn.Tanh, "sigmoid": nn.Sigmoid, "relu": nn.ReLU, "softplus": nn.Softplus, "ELU": nn.ELU,
        "silu": nn.SiLU}


class MLP(nn.Module):
    """models/components/mlp.py:22-96 (attribute layout ``linear_pre``, ``linears``, ``linear_post``)."""

    def __init__(self, n_input, n_hidden, n_output, n_layers=1, act="gelu", res=True):
        super().__init__()
        if act == "leaky_relu":
            make = lambda: nn.LeakyReLU(0.1)
        elif act in _ACT:
            make = _ACT[act]
        else:
            raise NotImplementedError
        self.n_input, self.n_hidden, self.n_output, self.n_layers, self.res = n_input, n_hidden, n_output, n_layers, res
        self.linear_pre = nn.Sequential(nn.Linear(n_input, n_hidden), make())
        self.linear_post = nn.Linear(n_hidden, n_output)
        self.linears = nn.ModuleList([nn.Sequential(nn.Linear(n_hidden, n_hidden), make()) for _ in range(n_layers)])

    def forward(self, x):
        x = self.linear_pre(x)
        for lin in self.linears:
            x = lin(x) + x if self.res else lin(x)
        return self.linear_post(x)


class TransolverErwinBlock(nn.Module):
    """HAETransolver_Irregular_Mesh.py:16-134 (2D / 3D: the same block around the structured attention)."""

    def __init__(self, attn_cls, num_heads, hidden_dim, dropout, act="gelu", mlp_ratio=4, last_layer=False, out_dim=1,
                 slice_num=32, **attn_kw):
        super().__init__()
        self.last_layer = last_layer
        self.ln_1 = nn.LayerNorm(hidden_dim)
        self.Attn = attn_cls(hidden_dim, heads=num_heads, dim_head=hidden_dim // num_heads, dropout=dropout,
                             slice_num=slice_num, mlp_ratio=mlp_ratio, **attn_kw)
        self.ln_2 = nn.LayerNorm(hidden_dim)
        self.mlp = MLP(hidden_dim, hidden_dim * mlp_ratio, hidden_dim, n_layers=0, res=False, act=act)
        if last_layer:
            self.ln_3 = nn.LayerNorm(hidden_dim)
            self.mlp2 = nn.Linear(hidden_dim, out_dim)

    def forward(self, fx):
        return fused_block_forward(self, fx)


def fused_block_forward(blk, fx, head: bool = True):
    """TransolverErwinBlock.forward (HAETransolver_Irregular_Mesh.py:114-134): the attention half as one fused call, the
    MLP half as the fused GEMM pair, the last block's ``mlp2(ln_3(fx))`` as one pass (``head=False`` leaves it to the caller)."""
    fx = blk.Attn.forward_block(fx, blk.ln_1)
    fx = block_mlp(fx, blk.ln_2, blk.mlp)
    if blk.last_layer and head:
        return head_ln_linear(fx, blk.ln_3, blk.mlp2)
    return fx


@torch.no_grad()
def head_ln_linear(fx: torch.Tensor, ln: nn.LayerNorm, lin: nn.Linear) -> torch.Tensor:
    """``lin(ln(fx))`` with a handful of output channels in one pass over fx (haet_head_ln_linear); other shapes run the modules."""
    if not fx.is_cuda:
        raise _lib.HaetError("haet_b200 runs on CUDA tensors only (no CPU fallback)")
    dim, out_dim = lin.in_features, lin.out_features
    if (dim % 128 or dim > 1024 or out_dim > 8 or ln.weight is None or ln.bias is None or lin.bias is None
            or tuple(ln.normalized_shape) != (dim,)):
        return lin(ln(fx))
    l = _lib.lib()
    x = fx.contiguous().float()
    M = x.numel() // dim
    f = lambda t: t.detach().float().contiguous()
    g, b, w, wb = f(ln.weight), f(ln.bias), f(lin.weight), f(lin.bias)
    y = torch.empty(*x.shape[:-1], out_dim, dtype=torch.float32, device=x.device)
    _lib.check(l.haet_head_ln_linear(x.data_ptr(), M, dim, g.data_ptr(), b.data_ptr(), float(ln.eps), w.data_ptr(),
                                     wb.data_ptr(), out_dim, y.data_ptr(), torch.cuda.current_stream(x.device).cuda_stream))
    return y if fx.dtype == torch.float32 else y.to(fx.dtype)


def timestep_embedding(timesteps, dim, max_period=10000):
    """models/components/embedding.py:72-99 (torch composition, used on the unfused route only)."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32) / half).to(timesteps.device)
    args = timesteps[:, None].float() * freqs[None]
    emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    if dim % 2:
        emb = torch.cat([emb, torch.zeros_like(emb[:, :1])], dim=-1)
    return emb


def _is_plain_gelu_mlp(mlp) -> bool:
    pre, post = getattr(mlp, "linear_pre", None), getattr(mlp, "linear_post", None)
    return (isinstance(pre, nn.Sequential) and len(pre) == 2 and isinstance(pre[0], nn.Linear)
            and isinstance(pre[1], nn.GELU) and getattr(pre[1], "approximate", "none") == "none"
            and isinstance(post, nn.Linear) and len(getattr(mlp, "linears", ())) == 0
            and pre[0].bias is not None and post.bias is not None)


@torch.no_grad()
def model_prologue(model, x: torch.Tensor, fx, T=None) -> torch.Tensor:
    """Everything of ``Model.forward`` in front of the blocks (HAETransolver_Irregular_Mesh.py:351-370; 2D :384-404; 3D
    :436-451): unified-position features, ``preprocess(cat(x, fx))``, placeholder, time embedding."""
    if not x.is_cuda:
        raise _lib.HaetError("haet_b200 runs on CUDA tensors only (no CPU fallback)")
    l = _lib.lib()
    dev = x.device
    B = x.shape[0]
    structured = hasattr(model, "H")
    ref = 0
    if model.unified_pos:
        if structured:   # precomputed grid-to-reference distances, the same for every batch item
            pos = model.pos.to(dev)
            x = pos.reshape(1, -1, pos.shape[-1]).expand(B, -1, -1)
        else:
            ref = int(model.ref)
    N = x.shape[1]
    # the structured models add the placeholder only when there are no function values (2D :399-402)
    use_placeholder = (not structured) or fx is None
    pre = model.preprocess
    n_hidden = pre.linear_post.out_features
    hidden = pre.linear_pre[0].out_features if isinstance(pre.linear_pre, nn.Sequential) else 0
    Ka = x.shape[-1]
    Kf = fx.shape[-1] if fx is not None else 0
    f = lambda t: t.detach().float().contiguous()
    wsz = -1
    if _is_plain_gelu_mlp(pre) and (ref == 0 or Ka == 2):
        wsz = l.haet_model_prologue_workspace_bytes(B * N, Ka, Kf, ref, hidden, n_hidden)
    if wsz < 0:   # unfused: the reference's own composition as torch ops on the GPU
        feat = model.get_grid(x, B) if ref else x
        h = pre(torch.cat((feat, fx), -1) if fx is not None else feat)
        if use_placeholder:
            h = h + model.placeholder[None, None, :]
        if T is not None:
            h = h + model.time_fc(timestep_embedding(T, n_hidden).repeat(1, N, 1))
        return h
    st = torch.cuda.current_stream(dev).cuda_stream
    xa = x.contiguous().float()
    xf = fx.contiguous().float() if fx is not None else None
    w1, b1, w2, b2 = f(pre.linear_pre[0].weight), f(pre.linear_pre[0].bias), f(pre.linear_post.weight), f(pre.linear_post.bias)
    add, per_batch = (f(model.placeholder) if use_placeholder else None), 0
    if T is not None:
        t = T.reshape(-1).float().contiguous()
        if t.numel() != B:
            raise ValueError(f"T must hold one time per batch item, got {tuple(T.shape)} for batch {B}")
        tw1, tb1, tw2, tb2 = f(model.time_fc[0].weight), f(model.time_fc[0].bias), f(model.time_fc[2].weight), f(model.time_fc[2].bias)
        vec = torch.empty(B, n_hidden, dtype=torch.float32, device=dev)
        _lib.check(l.haet_time_embed(t.data_ptr(), B, n_hidden, tw1.data_ptr(), tb1.data_ptr(), tw2.data_ptr(), tb2.data_ptr(),
                                     add.data_ptr() if add is not None else None, vec.data_ptr(), st))
        add, per_batch = vec, 1
    work = torch.empty(max(wsz, 4), dtype=torch.uint8, device=dev)
    out = torch.empty(B, N, n_hidden, dtype=torch.float32, device=dev)
    _lib.check(l.haet_model_prologue(xa.data_ptr(), Ka, xf.data_ptr() if xf is not None else None, Kf, ref, B, N, hidden,
                                     n_hidden, w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(),
                                     add.data_ptr() if add is not None else None, per_batch, out.data_ptr(),
                                     work.data_ptr(), st))
    return out


@torch.no_grad()
def fused_model_forward(model, x, fx, T=None):
    """``Model.forward(x, fx, T)`` of the three HAETransolver models on the fused kernels (see the module docstring)."""
    if model.training:
        raise NotImplementedError("haet_b200 is forward-only / eval mode; call .eval()")
    h = model_prologue(model, x, fx, T)
    for blk in model.blocks:
        if isinstance(getattr(blk, "Attn", None), _PhysicsAttentionBase):
            h = fused_block_forward(blk, h)
        else:
            h = blk(h)
    return h


def _grid_distances(sizes, ref):
    """get_grid of the structured models (HAETransolver_Structured_Mesh_2D.py:330-382, _3D.py:340-416): distances of
    every point of the linspace(0, 1) mesh to every point of the ref^nd reference grid -> [*sizes, ref^nd] (fp32)."""
    nd = len(sizes)
    axes = [torch.tensor(np.linspace(0, 1, s), dtype=torch.float) for s in sizes]
    raxes = [torch.tensor(np.linspace(0, 1, ref), dtype=torch.float) for _ in sizes]
    grid = torch.stack(torch.meshgrid(*axes, indexing="ij"), dim=-1)        # [*sizes, nd]
    gref = torch.stack(torch.meshgrid(*raxes, indexing="ij"), dim=-1)       # [ref]*nd + [nd]
    diff = grid.reshape(*sizes, *([1] * nd), nd) - gref.reshape(*([1] * nd), *([ref] * nd), nd)
    return torch.sqrt(torch.sum(diff ** 2, dim=-1)).reshape(*sizes, ref ** nd).contiguous()


class _ModelBase(nn.Module):
    def _build(self, attn_cls, pos_features, space_dim, n_layers, n_hidden, dropout, n_head, Time_Input, act, mlp_ratio,
               fun_dim, out_dim, slice_num, ref, unified_pos, attn_kw):
        self.ref, self.unified_pos, self.Time_Input = ref, unified_pos, Time_Input
        self.n_hidden, self.space_dim = n_hidden, space_dim
        n_in = fun_dim + (pos_features if unified_pos else space_dim)
        self.preprocess = MLP(n_in, n_hidden * 2, n_hidden, n_layers=0, res=False, act=act)
        if Time_Input:
            self.time_fc = nn.Sequential(nn.Linear(n_hidden, n_hidden), nn.SiLU(), nn.Linear(n_hidden, n_hidden))
        self.blocks = nn.ModuleList([
            TransolverErwinBlock(attn_cls, num_heads=n_head, hidden_dim=n_hidden, dropout=dropout, act=act,
                                 mlp_ratio=mlp_ratio, out_dim=out_dim, slice_num=slice_num,
                                 last_layer=(i == n_layers - 1), **attn_kw) for i in range(n_layers)])
        self.apply(self._init_weights)
        self.placeholder = nn.Parameter((1 / n_hidden) * torch.rand(n_hidden, dtype=torch.float))

    @staticmethod
    def _init_weights(m):   # HAETransolver_Irregular_Mesh.py:275-295
        if isinstance(m, nn.Linear):
            nn.init.trunc_normal_(m.weight, std=0.02)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, (nn.LayerNorm, nn.BatchNorm1d)):
            nn.init.constant_(m.bias, 0)
            nn.init.constant_(m.weight, 1.0)

    def forward(self, x, fx, T=None):
        return fused_model_forward(self, x, fx, T)


def _erwin_kw(c_hidden, ball_sizes, enc_num_heads, enc_depths, dec_num_heads, dec_depths, strides, rotate, decode, mp_steps,
              embed):
    return dict(c_hidden=c_hidden, ball_sizes=ball_sizes, enc_num_heads=enc_num_heads, enc_depths=enc_depths,
                dec_num_heads=dec_num_heads, dec_depths=dec_depths, strides=strides, rotate=rotate, decode=decode,
                mp_steps=mp_steps, embed=embed)


class HAETransolver_Irregular_Mesh(_ModelBase):
    """``Model`` of models/HAETransolver_Irregular_Mesh.py (constructor :157-184)."""

    def __init__(self, space_dim=1, n_layers=5, n_hidden=256, dropout=0.0, n_head=8, Time_Input=False, act="gelu",
                 mlp_ratio=1, fun_dim=1, out_dim=1, slice_num=32, ref=8, unified_pos=False, c_hidden=None, ball_sizes=None,
                 enc_num_heads=None, enc_depths=None, dec_num_heads=None, dec_depths=None, strides=None, rotate=45,
                 decode=True, mp_steps=0, embed=False):
        super().__init__()
        self.__name__ = "Transolver_1D"
        self._build(Physics_Attention_Irregular_Mesh, ref * ref, space_dim, n_layers, n_hidden, dropout, n_head, Time_Input,
                    act, mlp_ratio, fun_dim, out_dim, slice_num, ref, unified_pos,
                    _erwin_kw(c_hidden, ball_sizes, enc_num_heads, enc_depths, dec_num_heads, dec_depths, strides, rotate,
                              decode, mp_steps, embed))

    def get_grid(self, x, batchsize=1):
        """:297-336 on x's own device (the reference hard-codes .cuda())."""
        g = torch.tensor(np.linspace(0, 1, self.ref), dtype=torch.float, device=x.device)
        gref = torch.stack(torch.meshgrid(g, g, indexing="ij"), dim=-1).reshape(1, self.ref * self.ref, 2)
        return torch.sqrt(torch.sum((x[:, :, None, :] - gref[:, None, :, :]) ** 2, dim=-1)).contiguous()


class HAETransolver_Structured_Mesh_2D(_ModelBase):
    """``Model`` of models/HAETransolver_Structured_Mesh_2D.py."""

    def __init__(self, space_dim=1, n_layers=5, n_hidden=256, dropout=0.0, n_head=8, Time_Input=False, act="gelu",
                 mlp_ratio=1, fun_dim=1, out_dim=1, slice_num=32, ref=8, unified_pos=False, H=85, W=85, c_hidden=None,
                 ball_sizes=None, enc_num_heads=None, enc_depths=None, dec_num_heads=None, dec_depths=None, strides=None,
                 rotate=45, decode=True, mp_steps=0, embed=False):
        super().__init__()
        self.__name__ = "Transolver_2D"
        self.H, self.W = H, W
        if unified_pos:
            self.pos = _grid_distances((H, W), ref)[None]
        self._build(Physics_Attention_Structured_Mesh_2D, ref * ref, space_dim, n_layers, n_hidden, dropout, n_head,
                    Time_Input, act, mlp_ratio, fun_dim, out_dim, slice_num, ref, unified_pos,
                    dict(H=H, W=W, **_erwin_kw(c_hidden, ball_sizes, enc_num_heads, enc_depths, dec_num_heads, dec_depths,
                                               strides, rotate, decode, mp_steps, embed)))


class HAETransolver_Structured_Mesh_3D(_ModelBase):
    """``Model`` of models/HAETransolver_Structured_Mesh_3D.py."""

    def __init__(self, space_dim=1, n_layers=5, n_hidden=256, dropout=0.0, n_head=8, Time_Input=False, act="gelu",
                 mlp_ratio=1, fun_dim=1, out_dim=1, slice_num=32, ref=8, unified_pos=False, H=32, W=32, D=32, c_hidden=None,
                 ball_sizes=None, enc_num_heads=None, enc_depths=None, dec_num_heads=None, dec_depths=None, strides=None,
                 rotate=45, decode=True, mp_steps=0, embed=False):
        super().__init__()
        self.__name__ = "Transolver_3D"
        self.use_checkpoint = False
        self.H, self.W, self.D = H, W, D
        if unified_pos:
            self.pos = _grid_distances((H, W, D), ref)[None]
        self._build(Physics_Attention_Structured_Mesh_3D, ref ** 3, space_dim, n_layers, n_hidden, dropout, n_head,
                    Time_Input, act, mlp_ratio, fun_dim, out_dim, slice_num, ref, unified_pos,
                    dict(H=H, W=W, D=D, **_erwin_kw(c_hidden, ball_sizes, enc_num_heads, enc_depths, dec_num_heads,
                                                    dec_depths, strides, rotate, decode, mp_steps, embed)))
