"""Drop-in PhysicsAttention modules (reference: models/PhysicsAttention/{IrregularMesh,StructuredMesh2D,
StructuredMesh3D}.py).  Same constructor arguments, attribute and parameter names, ``forward(x)``.

forward (IrregularMesh.py:141-235) is five C-ABI calls into libhaet_b200.so, all enqueued on the current
CUDA stream with no host synchronisation:
    haet_slice_reduce  -> [all_gather of the packed slice state when point-sharded]
    haet_slice_merge   -> haet_balltree_build -> haet_erwin_forward -> haet_deslice_out
Forward / eval only (no autograd graph is recorded); there is no CPU or PyTorch fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
import torch.nn as nn

from . import _lib
from .balltree import balltree_device
from .erwin import ErwinTransformer, erwin_device


class _PhysicsAttentionBase(nn.Module):
    conv_nd = 0

    def _setup(self, dim, heads, dim_head, dropout, slice_num, base_temp, epsilon, c_hidden, ball_sizes,
               enc_num_heads, enc_depths, dec_num_heads, dec_depths, strides, rotate, decode, mlp_ratio, mp_steps,
               embed, erwin_variant="flash"):
        inner_dim = dim_head * heads
        self.dim_head, self.heads = dim_head, heads
        self.epsilon, self.base_temp = epsilon, base_temp
        self.dim, self.slice_num = dim, slice_num
        self.in_project_slice = nn.Linear(dim_head, slice_num)
        self.ada_temp_linear = nn.Linear(dim_head, 1)
        torch.nn.init.orthogonal_(self.in_project_slice.weight)
        # defaults of IrregularMesh.py:105-118
        if c_hidden is None:
            c_hidden = [dim_head, dim_head * 2]
        if ball_sizes is None:
            ball_sizes = [min(32, slice_num), min(16, slice_num // 2)]
        if enc_num_heads is None:
            enc_num_heads = [heads // 2, heads]
        if enc_depths is None:
            enc_depths = [2, 2]
        if dec_num_heads is None:
            dec_num_heads = [heads // 2]
        if dec_depths is None:
            dec_depths = [2]
        if strides is None:
            strides = [2]
        self.erwin = ErwinTransformer(c_in=dim_head, c_hidden=c_hidden, ball_sizes=ball_sizes,
                                      enc_num_heads=enc_num_heads, enc_depths=enc_depths, dec_num_heads=dec_num_heads,
                                      dec_depths=dec_depths, strides=strides, rotate=rotate, decode=decode,
                                      mlp_ratio=mlp_ratio, dimensionality=self.dimensionality, mp_steps=mp_steps,
                                      embed=embed, variant=erwin_variant)
        self.to_out = nn.Sequential(nn.Linear(inner_dim, dim), nn.Dropout(dropout))
        # point-sharded multi-GPU: set to a torch.distributed group whose ranks hold disjoint point ranges
        self.point_group = None
        self._px_key, self._px = None, None   # peer-memory exchange of the slice state (distributed.PeerExchange)
        self._wx_cache = None
        self.last_taps = None   # filled when keep_taps is set (tests)
        self.keep_taps = False

    # weights in the layout the kernels read -------------------------------------------------------------
    def _wx(self):
        w = self.in_project_x.weight
        if w.dim() == 2:
            return w.detach().contiguous()
        key = (w.data_ptr(), w._version, str(w.device))
        if self._wx_cache is None or self._wx_cache[0] != key:
            # torch [inner, dim, k, k(, k)] -> tap-major [inner, taps, dim]
            inner, dim = w.shape[:2]
            self._wx_cache = (key, w.detach().reshape(inner, dim, -1).permute(0, 2, 1).contiguous())
        return self._wx_cache[1]

    def _shape(self, B, N, planes=None):
        grid = (1, 1, 1)
        if self.conv_nd == 2:
            grid = (self.H, self.W, 1)
        elif self.conv_nd == 3:
            grid = (self.H, self.W, self.D)
        if planes is not None:   # slab sharding: this rank's planes of the first grid axis
            grid = (planes,) + grid[1:]
        k = self.in_project_x.weight.shape[-1] if self.conv_nd else 1
        return _lib.make_shape(B, N, self.dim, self.heads, self.dim_head, self.slice_num, self.conv_nd, grid, k,
                               self.dimensionality, self.base_temp, self.epsilon)

    @torch.no_grad()
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self._forward(x, None, False)

    @torch.no_grad()
    def forward_block(self, fx: torch.Tensor, ln_1: nn.LayerNorm) -> torch.Tensor:
        """``self(ln_1(fx)) + fx`` -- the attention half of TransolverErwinBlock.forward
        (models/HAETransolver_Irregular_Mesh.py:125) with the LayerNorm fused into the in-projection kernel and the
        residual add into the to_out epilogue (haet_slice_reduce_ln / haet_deslice_out_res)."""
        if tuple(ln_1.normalized_shape) != (self.dim,) or ln_1.weight is None or ln_1.bias is None:
            raise ValueError("forward_block expects an affine nn.LayerNorm over the hidden dimension")
        return self._forward(fx, ln_1, True)

    def _forward(self, x: torch.Tensor, ln_1: Optional[nn.LayerNorm], residual: bool) -> torch.Tensor:
        if not x.is_cuda:
            raise _lib.HaetError("haet_b200 PhysicsAttention runs on CUDA tensors only (no CPU fallback)")
        if x.dim() != 3 or x.shape[2] != self.dim:
            raise ValueError(f"expected x of shape [B, N, {self.dim}], got {tuple(x.shape)}")
        l = _lib.lib()
        in_dtype = x.dtype
        x = x.contiguous().float()
        B, N, _ = x.shape
        dev = x.device
        if B == 0:   # nothing to do (the C-ABI returns HAET_OK for an empty batch too)
            return x.new_empty((0, N, self.dim), dtype=in_dtype)
        f32 = dict(dtype=torch.float32, device=dev)
        slab = self.conv_nd != 0 and self.point_group is not None   # structured mesh sharded by slabs of planes + halo exchange
        if slab:
            plane = self.W * (self.D if self.conv_nd == 3 else 1)
            if N % plane:
                raise ValueError(f"slab sharding: a rank's {N} points are not whole planes of {plane} points")
            sh = self._shape(B, N, N // plane)
        else:
            sh = self._shape(B, N)
        psh = C.byref(sh)
        st = torch.cuda.current_stream(dev).cuda_stream
        wx = self._wx().float()
        bx = self.in_project_x.bias.detach().float().contiguous()
        ws = self.in_project_slice.weight.detach().float().contiguous()
        bs = self.in_project_slice.bias.detach().float().contiguous()
        wt = self.ada_temp_linear.weight.detach().float().contiguous().view(-1)
        bt = self._bt()
        wo = self.to_out[0].weight.detach().float().contiguous()
        bo = self.to_out[0].bias.detach().float().contiguous()

        H, G, Cc = self.heads, self.slice_num, self.dim_head
        T = B * H * G
        spill = torch.empty(max(l.haet_spill_bytes(psh), 4), dtype=torch.uint8, device=dev)
        if slab:
            from .distributed import exchange_halo
            x_ext, h_lo, h_hi = exchange_halo(x, plane, int(self.in_project_x.weight.shape[-1]) // 2, self.point_group)
            slice_ws = l.haet_slice_halo_workspace_bytes(psh, h_lo, h_hi, int(ln_1 is not None))
            if slice_ws < 0:
                _lib.check(int(slice_ws))
        else:
            slice_ws = l.haet_slice_ln_workspace_bytes(psh) if ln_1 is not None else l.haet_slice_workspace_bytes(psh)
        wsz = max(slice_ws, l.haet_deslice_workspace_bytes(psh), 4)
        work = torch.empty(wsz, dtype=torch.uint8, device=dev)
        partial = torch.empty(l.haet_partial_floats(psh), **f32)
        if slab:
            lg = ln_1.weight.detach().float().contiguous() if ln_1 is not None else None
            lb = ln_1.bias.detach().float().contiguous() if ln_1 is not None else None
            _lib.check(l.haet_slice_reduce_halo(psh, h_lo, h_hi, x_ext.data_ptr(), lg.data_ptr() if lg is not None else None,
                                                lb.data_ptr() if lb is not None else None, float(ln_1.eps) if ln_1 is not None else 0.0,
                                                wx.data_ptr(), bx.data_ptr(), ws.data_ptr(), bs.data_ptr(), wt.data_ptr(), bt,
                                                spill.data_ptr(), partial.data_ptr(), work.data_ptr(), st))
        elif ln_1 is not None:
            lg = ln_1.weight.detach().float().contiguous()
            lb = ln_1.bias.detach().float().contiguous()
            _lib.check(l.haet_slice_reduce_ln(psh, x.data_ptr(), lg.data_ptr(), lb.data_ptr(), float(ln_1.eps),
                                              wx.data_ptr(), bx.data_ptr(), ws.data_ptr(), bs.data_ptr(), wt.data_ptr(),
                                              bt, spill.data_ptr(), partial.data_ptr(), work.data_ptr(), st))
        else:
            _lib.check(l.haet_slice_reduce(psh, x.data_ptr(), wx.data_ptr(), bx.data_ptr(), ws.data_ptr(),
                                           bs.data_ptr(), wt.data_ptr(), bt, spill.data_ptr(), partial.data_ptr(),
                                           work.data_ptr(), st))
        parts, n_ranks = partial, 1
        if self.point_group is not None:
            parts, n_ranks = self._exchange(sh, partial)
        eidetic = torch.empty(T, Cc, **f32)
        lse2 = torch.empty(T, **f32)
        pos = torch.empty(T, self.dimensionality, **f32)
        _lib.check(l.haet_slice_merge(psh, parts.data_ptr(), n_ranks, eidetic.data_ptr(), lse2.data_ptr(),
                                      pos.data_ptr(), st))
        # self.training: forward only (partial) -- Erwin's BatchNorm1d layers use batch statistics and update their
        # running buffers (erwin.py); the result has no grad_fn, there is no backward pass
        tree_idx, tree_mask, rots = balltree_device(pos, B * H, G, self.erwin.strides, self.erwin.ball_sizes,
                                                    float(self.erwin.rotate))
        states = erwin_device(self.erwin, eidetic, pos, tree_idx, tree_mask, rots, B * H)
        out = torch.empty(B, N, self.dim, **f32)
        if residual and self.to_out[1].p > 0 and self.to_out[1].training:
            raise NotImplementedError("forward_block is eval-mode only (dropout sits between to_out and the residual)")
        _lib.check(l.haet_deslice_out_res(psh, spill.data_ptr(), lse2.data_ptr(), states.data_ptr(), ws.data_ptr(),
                                          bs.data_ptr(), wt.data_ptr(), bt, wo.data_ptr(), bo.data_ptr(),
                                          x.data_ptr() if residual else None, out.data_ptr(), work.data_ptr(), st))
        if self.keep_taps:
            self.last_taps = dict(S=eidetic.view(B, H, G, Cc), lse2=lse2.view(B, H, G), pos=pos, tree_idx=tree_idx,
                                  tree_mask=tree_mask, rot_idx=rots, S_out=states)
        out = self.to_out[1](out)   # nn.Dropout: identity in eval / p == 0 (IrregularMesh.py:139,235)
        return out if in_dtype == torch.float32 else out.to(in_dtype)

    def _exchange(self, sh, partial):
        """Every rank's packed slice state, rank-major: one kernel over NVLink peer memory when it can be set up (and
        HAET_NCCL_EXCHANGE is not set), else an NCCL (or gloo) all-gather."""
        import os
        from .distributed import PeerExchange, gather_partials
        key = (id(self.point_group), sh.B, sh.slice_num, sh.heads, sh.dim_head, str(partial.device))
        if self._px_key != key:
            self._px_key, self._px = key, None
            if partial.is_cuda and not os.environ.get("HAET_NCCL_EXCHANGE"):
                self._px = PeerExchange.create(sh, self.point_group, partial.device)
        if self._px is not None:
            return self._px(sh, partial)
        return gather_partials(partial, self.point_group)

    def _bt(self) -> float:
        """ada_temp_linear.bias as a host scalar, cached per parameter version (one sync when it changes)."""
        b = self.ada_temp_linear.bias
        key = (b.data_ptr(), b._version)
        c = getattr(self, "_bt_cache", None)
        if c is None or c[0] != key:
            c = (key, float(b.detach().float().cpu()))
            self._bt_cache = c
        return c[1]


class Physics_Attention_Irregular_Mesh(_PhysicsAttentionBase):
    """IrregularMesh.py:17-139."""

    def __init__(self, dim, heads=8, dim_head=64, dropout=0.0, slice_num=64, base_temp=0.5, epsilon=1e-6,
                 c_hidden=None, ball_sizes=None, enc_num_heads=None, enc_depths=None, dec_num_heads=None,
                 dec_depths=None, strides=None, rotate=1, decode=True, mlp_ratio=4, mp_steps=0, embed=False, *,
                 erwin_variant="flash"):
        super().__init__()
        self.dimensionality = 3
        self.in_project_x = nn.Linear(dim, dim_head * heads)
        self._setup(dim, heads, dim_head, dropout, slice_num, base_temp, epsilon, c_hidden, ball_sizes, enc_num_heads,
                    enc_depths, dec_num_heads, dec_depths, strides, rotate, decode, mlp_ratio, mp_steps, embed,
                    erwin_variant)


class Physics_Attention_Structured_Mesh_2D(_PhysicsAttentionBase):
    """StructuredMesh2D.py:18-156: Conv2d(dim, inner, kernel, 1, kernel // 2) front-end on the H x W grid."""
    conv_nd = 2

    def __init__(self, dim, heads=8, dim_head=64, dropout=0.0, slice_num=64, H=101, W=31, kernel=3, base_temp=0.5,
                 epsilon=1e-6, c_hidden=None, ball_sizes=None, enc_num_heads=None, enc_depths=None,
                 dec_num_heads=None, dec_depths=None, strides=None, rotate=1, decode=True, mlp_ratio=4, mp_steps=0,
                 embed=False, *, erwin_variant="flash"):
        super().__init__()
        self.H, self.W = H, W
        self.dimensionality = 2
        self.in_project_x = nn.Conv2d(dim, dim_head * heads, kernel, 1, kernel // 2)
        self._setup(dim, heads, dim_head, dropout, slice_num, base_temp, epsilon, c_hidden, ball_sizes, enc_num_heads,
                    enc_depths, dec_num_heads, dec_depths, strides, rotate, decode, mlp_ratio, mp_steps, embed,
                    erwin_variant)


class Physics_Attention_Structured_Mesh_3D(_PhysicsAttentionBase):
    """StructuredMesh3D.py:19-160: Conv3d front-end on the H x W x D grid."""
    conv_nd = 3

    def __init__(self, dim, heads=8, dim_head=64, dropout=0.0, slice_num=32, H=32, W=32, D=32, kernel=3,
                 base_temp=0.5, epsilon=1e-6, c_hidden=None, ball_sizes=None, enc_num_heads=None, enc_depths=None,
                 dec_num_heads=None, dec_depths=None, strides=None, rotate=1, decode=True, mlp_ratio=4, mp_steps=0,
                 embed=False, *, erwin_variant="flash"):
        super().__init__()
        self.H, self.W, self.D = H, W, D
        self.dimensionality = 3
        self.in_project_x = nn.Conv3d(dim, dim_head * heads, kernel, 1, kernel // 2)
        self._setup(dim, heads, dim_head, dropout, slice_num, base_temp, epsilon, c_hidden, ball_sizes, enc_num_heads,
                    enc_depths, dec_num_heads, dec_depths, strides, rotate, decode, mlp_ratio, mp_steps, embed,
                    erwin_variant)


@torch.no_grad()
def block_mlp(fx: torch.Tensor, ln_2: nn.LayerNorm, mlp: nn.Module) -> torch.Tensor:
    """``mlp(ln_2(fx)) + fx`` (models/HAETransolver_Irregular_Mesh.py:128) through haet_block_mlp when ``mlp`` is the
    reference's ``MLP(n_layers=0, act='gelu')`` (models/components/mlp.py:67-96: ``linear_pre = Sequential(Linear, GELU)``,
    ``linear_post = Linear``) of a supported shape (dim 256, hidden a multiple of 256); any other MLP runs as the
    block's own torch modules on the GPU."""
    if not fx.is_cuda:
        raise _lib.HaetError("haet_b200.block_mlp runs on CUDA tensors only (no CPU fallback)")
    pre, post = getattr(mlp, "linear_pre", None), getattr(mlp, "linear_post", None)
    ok = (isinstance(pre, nn.Sequential) and len(pre) == 2 and isinstance(pre[0], nn.Linear)
          and isinstance(pre[1], nn.GELU) and getattr(pre[1], "approximate", "none") == "none"
          and isinstance(post, nn.Linear) and len(getattr(mlp, "linears", ())) == 0
          and ln_2.weight is not None and ln_2.bias is not None and pre[0].bias is not None and post.bias is not None)
    if ok:
        l = _lib.lib()
        dim, hidden = pre[0].in_features, pre[0].out_features
        M = fx.numel() // fx.shape[-1]
        wsz = l.haet_block_mlp_workspace_bytes(M, dim, hidden) if fx.shape[-1] == dim == post.out_features else -1
        if wsz >= 0:
            in_dtype = fx.dtype
            x = fx.contiguous().float()
            f = lambda t: t.detach().float().contiguous()
            w1, b1, w2, b2, g, b = f(pre[0].weight), f(pre[0].bias), f(post.weight), f(post.bias), f(ln_2.weight), f(ln_2.bias)
            work = torch.empty(max(wsz, 4), dtype=torch.uint8, device=x.device)
            out = torch.empty_like(x)
            st = torch.cuda.current_stream(x.device).cuda_stream
            _lib.check(l.haet_block_mlp(x.data_ptr(), M, dim, hidden, g.data_ptr(), b.data_ptr(), float(ln_2.eps),
                                        w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), out.data_ptr(),
                                        work.data_ptr(), st))
            return out if in_dtype == torch.float32 else out.to(in_dtype)
    return mlp(ln_2(fx)) + fx


def patch_model(model: nn.Module, fuse_blocks: bool = False, fuse_io: bool = None) -> nn.Module:
    """Swap every reference ``Physics_Attention_*`` inside ``model`` (a reference-constructed HAETransolver_*)
    for the haet_b200 module of the same name, carrying the state_dict over (strict).  With ``fuse_blocks`` every
    block that holds one (attributes ``ln_1, Attn, ln_2, mlp, last_layer`` as in TransolverErwinBlock) also gets the
    fused block forward (models.fused_block_forward; eval / inference only), and with ``fuse_io`` (default: the same as
    ``fuse_blocks``) a model with the reference's ``preprocess / placeholder / blocks`` attributes gets the fused
    ``forward(x, fx, T)`` (models.fused_model_forward: prologue and last-layer head kernels)."""
    import types
    from . import models as _models
    mapping = {c.__name__: c for c in (Physics_Attention_Irregular_Mesh, Physics_Attention_Structured_Mesh_2D,
                                       Physics_Attention_Structured_Mesh_3D)}
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            cls = mapping.get(type(child).__name__)
            if cls is None or isinstance(child, _PhysicsAttentionBase):
                continue
            new = from_reference(child, cls)
            setattr(parent, name, new)
    if fuse_blocks:
        for blk in model.modules():
            if isinstance(getattr(blk, "Attn", None), _PhysicsAttentionBase) and \
                    all(hasattr(blk, a) for a in ("ln_1", "ln_2", "mlp", "last_layer")):
                blk.forward = types.MethodType(_models.fused_block_forward, blk)
    if (fuse_blocks if fuse_io is None else fuse_io) and all(hasattr(model, a) for a in ("preprocess", "placeholder", "blocks",
                                                                                         "unified_pos")):
        model.forward = types.MethodType(_models.fused_model_forward, model)
    return model


def from_reference(ref: nn.Module, cls=None) -> _PhysicsAttentionBase:
    """Build the haet_b200 twin of a reference PhysicsAttention module instance."""
    cls = cls or {"Physics_Attention_Irregular_Mesh": Physics_Attention_Irregular_Mesh,
                  "Physics_Attention_Structured_Mesh_2D": Physics_Attention_Structured_Mesh_2D,
                  "Physics_Attention_Structured_Mesh_3D": Physics_Attention_Structured_Mesh_3D}[type(ref).__name__]
    e = ref.erwin
    n_enc = len(e.encoder)
    enc_layers = list(e.encoder) + [e.bottleneck]
    c_hidden = [l.blocks[0].norm1.weight.shape[0] if len(l.blocks) else None for l in enc_layers]
    kw = dict(dim=ref.to_out[0].weight.shape[0], heads=ref.heads, dim_head=ref.dim_head, dropout=ref.to_out[1].p,
              slice_num=ref.in_project_slice.weight.shape[0], base_temp=ref.base_temp, epsilon=ref.epsilon,
              c_hidden=c_hidden, ball_sizes=list(e.ball_sizes),
              enc_num_heads=[l.blocks[0].BMSA.num_heads for l in enc_layers],
              enc_depths=[len(l.blocks) for l in enc_layers],
              dec_num_heads=[l.blocks[0].BMSA.num_heads for l in reversed(list(e.decoder))] if e.decode else [1] * n_enc,
              dec_depths=[len(l.blocks) for l in reversed(list(e.decoder))] if e.decode else [0] * n_enc,
              strides=list(e.strides), rotate=e.rotate, decode=e.decode,
              mlp_ratio=enc_layers[0].blocks[0].swiglu.w1.weight.shape[0] // c_hidden[0])
    if cls.conv_nd == 2:
        kw.update(H=ref.H, W=ref.W, kernel=ref.in_project_x.weight.shape[-1])
    elif cls.conv_nd == 3:
        kw.update(H=ref.H, W=ref.W, D=ref.D, kernel=ref.in_project_x.weight.shape[-1])
    new = cls(**kw)
    new.load_state_dict(ref.state_dict(), strict=True)
    new.to(next(ref.parameters()).device)
    new.train(ref.training)
    return new
