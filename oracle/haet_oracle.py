"""CPU restatement of HAET's per-layer hot path.  TEST INFRASTRUCTURE ONLY.

This file is the checker: only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import it.  The product (``haet_b200``)
never does, and fails loudly when its CUDA library is missing.

It restates, in plain torch-on-CPU arithmetic (fp32 by default, fp64 on request) driven by a
``state_dict`` with the reference's parameter names:

  * Physics_Attention_{Irregular_Mesh,Structured_Mesh_2D,Structured_Mesh_3D}.forward
      /root/reference/models/PhysicsAttention/IrregularMesh.py:141-235
      /root/reference/models/PhysicsAttention/StructuredMesh2D.py:173-189 (conv front-end)
      /root/reference/models/PhysicsAttention/StructuredMesh3D.py:177-193 (conv front-end)
  * ErwinTransformer.forward      models/components/erwinflash/erwin_flash.py:138-195
  * BasicLayer / ErwinTransformerBlock  .../components/layers.py:30-32,72-92
  * BallMSA (+compute_rel_pos)     .../components/attention.py:35-91  (flash-attn, a
      third-party fp16 CUDA library with no pinned version, replaced by its published
      arithmetic softmax(QK^T/sqrt(d))V in the working dtype -- SURVEY F9/F10)
  * SwiGLU                         .../components/mlp.py:17-18
  * BallPooling / BallUnpooling    .../components/pooling.py:25-40,58-72
  * build_balltree_with_rotations  models/components/balltree/balltree.pyx:598-662
      through ``oracle/balltree_oracle.c`` (ctypes)

Parity status: PINNED.  ``tests/test_oracle_vs_reference.py`` runs the unmodified reference
(imported by ``oracle/ref_loader.py``) beside this file where /root/reference exists, and
``tests/golden/*.npz`` (made by ``tests/golden/make_golden.py`` from the reference itself)
pin it everywhere else.
"""
from __future__ import annotations

import ctypes
import math
import os
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(HERE, "libballtree_oracle.so")
        if not os.path.exists(path):
            from . import build_ref
            build_ref.build_c_oracle()
        lib = ctypes.CDLL(path)
        lib.haet_oracle_tree_slots.restype = ctypes.c_int64
        _LIB = lib
    return _LIB


def _offsets_from_batch_idx(batch_idx: np.ndarray) -> np.ndarray:
    """balltree.pyx:444-472 -- offsets from the change points of a sorted batch index."""
    if batch_idx.ndim != 1 or len(batch_idx) == 0:
        raise ValueError("batch_idx must be a non-empty 1D array")
    if batch_idx[0] != 0:
        raise ValueError("Batch indices must start from 0")
    n = len(batch_idx)
    change = np.where(batch_idx[1:] != batch_idx[:-1])[0] + 1
    nb = int(batch_idx[-1]) + 1
    off = np.zeros(nb + 1, dtype=np.int64)
    off[1:nb] = change
    off[nb] = n
    return off


def _p(a: np.ndarray, t):
    return a.ctypes.data_as(ctypes.POINTER(t))


def balltree_build_with_rotations(pos: torch.Tensor, batch_idx: torch.Tensor, strides: Sequence[int],
                                  ball_sizes: Sequence[int], angle: float = 45.0):
    """balltree.pyx:598-662.  Returns (tree_idx int64, tree_mask bool, [rot_idx per level | None])."""
    assert len(strides) == len(ball_sizes) - 1, "Strides must be one less than ball sizes"
    if pos.dim() != 2 or batch_idx.dim() != 1 or pos.shape[0] != batch_idx.shape[0]:
        raise ValueError("bad shapes")
    p = np.ascontiguousarray(pos.detach().cpu().to(torch.float32).numpy())
    off = _offsets_from_batch_idx(batch_idx.detach().cpu().numpy().astype(np.int64))
    n, d = p.shape
    nt = len(off) - 1
    lib = _lib()
    total = lib.haet_oracle_tree_slots(_p(off, ctypes.c_int64), ctypes.c_int64(nt))
    if total < 0:
        raise ValueError("every tree needs at least 2 points")
    nl = len(ball_sizes)
    tree_idx = np.empty(total, dtype=np.int64)
    tree_mask = np.empty(total, dtype=np.uint8)
    rot = np.empty(total * nl, dtype=np.int64)
    rot_off = np.zeros(nl + 1, dtype=np.int64)
    st = np.asarray(list(strides) + [1], dtype=np.int32)
    bs = np.asarray(list(ball_sizes), dtype=np.int32)
    rc = lib.haet_oracle_build_with_rotations(
        _p(p, ctypes.c_float), ctypes.c_int64(n), ctypes.c_int(d), _p(off, ctypes.c_int64),
        ctypes.c_int64(nt), _p(st, ctypes.c_int32), _p(bs, ctypes.c_int32), ctypes.c_int(nl),
        ctypes.c_double(float(angle)), _p(tree_idx, ctypes.c_int64), _p(tree_mask, ctypes.c_uint8),
        _p(rot, ctypes.c_int64), _p(rot_off, ctypes.c_int64))
    if rc != 0:
        raise RuntimeError(f"balltree oracle failed with code {rc}")
    if angle > 0:
        rots = [torch.from_numpy(rot[rot_off[i]:rot_off[i + 1]].copy()) for i in range(nl)]
    else:
        rots = [None] * nl
    return torch.from_numpy(tree_idx), torch.from_numpy(tree_mask.astype(np.bool_)), rots


# --------------------------------------------------------------------------------------
# Erwin over the slice tokens
# --------------------------------------------------------------------------------------
def default_erwin_cfg(heads: int, dim_head: int, slice_num: int, dimensionality: int, rotate=45,
                      mlp_ratio: int = 4, **over) -> dict:
    """Defaults of Physics_Attention_*.__init__ (IrregularMesh.py:105-137)."""
    cfg = dict(c_hidden=[dim_head, dim_head * 2],
               ball_sizes=[min(32, slice_num), min(16, slice_num // 2)],
               enc_num_heads=[heads // 2, heads], enc_depths=[2, 2],
               dec_num_heads=[heads // 2], dec_depths=[2], strides=[2],
               rotate=rotate, decode=True, mlp_ratio=mlp_ratio, dimensionality=dimensionality)
    for k, v in over.items():
        if v is not None:
            cfg[k] = v
    return cfg


def _rmsnorm(x, w):
    # nn.RMSNorm(dim) with eps=None -> torch.finfo(dtype).eps (layers.py:24-25)
    eps = torch.finfo(x.dtype).eps
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * w


def _bn(x, sd, pre, training):
    # nn.BatchNorm1d (pooling.py:23,56): eval -> running stats; train -> biased batch stats
    w, b = sd[pre + ".weight"], sd[pre + ".bias"]
    if training:
        mean, var = x.mean(0), x.var(0, unbiased=False)
    else:
        mean, var = sd[pre + ".running_mean"], sd[pre + ".running_var"]
    return (x - mean) / torch.sqrt(var + 1e-5) * w + b


def _ball_msa(x, pos, sd, pre, num_heads, ball_size, variant="flash"):
    """attention.py:35-91 (flash variant: no distance bias, column layout (three, head, hd)).

    variant="erwin": the non-flash copy, benchmarks/02-Car-Design-ShapeNetCar/models/components/erwin/components/
    attention.py:40-53 -- qkv columns laid out (H E K) (:42-48) and the distance bias
    sigma_att * cdist(pos, pos) (:28-32) passed to scaled_dot_product_attention as attn_mask (:49-51), i.e. added
    to q k^T / sqrt(E) before the softmax."""
    T, dim = x.shape
    hd = dim // num_heads
    nb = T // ball_size
    p = pos.view(nb, ball_size, -1)
    rel = (p - p.mean(dim=1, keepdim=True)).view(T, -1)                      # :35-40
    x = x + F.linear(rel, sd[pre + ".pe_proj.weight"], sd[pre + ".pe_proj.bias"])  # :59-62
    qkv = F.linear(x, sd[pre + ".qkv.weight"], sd[pre + ".qkv.bias"])          # :65
    if variant == "erwin":
        qkv = qkv.view(nb, ball_size, num_heads, hd, 3)                        # "(n m) (H E K) -> K n H m E"
        q, k, v = (qkv[..., i].permute(0, 2, 1, 3) for i in range(3))         # [nb, h, n, hd]
        bias = sd[pre + ".sigma_att"].to(x.dtype) * torch.cdist(p, p, p=2).unsqueeze(1)   # [1,H,1,1] * [nb,1,m,m]
        att = torch.softmax(q @ k.transpose(-1, -2) * (1.0 / math.sqrt(hd)) + bias, dim=-1)
        o = (att @ v).permute(0, 2, 1, 3).reshape(T, dim)                      # "n H m E -> (n m) (H E)"
        return F.linear(o, sd[pre + ".proj.weight"], sd[pre + ".proj.bias"])
    qkv = qkv.view(nb, ball_size, 3, num_heads, hd)                            # :68-76
    q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))            # [nb, h, n, hd]
    att = torch.softmax(q @ k.transpose(-1, -2) * (1.0 / math.sqrt(hd)), dim=-1)  # :79-84
    o = (att @ v).permute(0, 2, 1, 3).reshape(T, dim)                          # :87
    return F.linear(o, sd[pre + ".proj.weight"], sd[pre + ".proj.bias"])       # :88


def _swiglu(x, sd, pre):
    """mlp.py:17-18"""
    return F.linear(F.linear(x, sd[pre + ".w2.weight"], sd[pre + ".w2.bias"])
                    * F.silu(F.linear(x, sd[pre + ".w1.weight"], sd[pre + ".w1.bias"])),
                    sd[pre + ".w3.weight"], sd[pre + ".w3.bias"])


def _block(x, pos, sd, pre, num_heads, ball_size, variant="flash"):
    """layers.py:30-32"""
    x = x + _ball_msa(_rmsnorm(x, sd[pre + ".norm1.weight"]), pos, sd, pre + ".BMSA", num_heads, ball_size, variant)
    return x + _swiglu(_rmsnorm(x, sd[pre + ".norm2.weight"]), sd, pre + ".swiglu")


def _blocks(x, pos, rot, sd, pre, depth, num_heads, ball_size, rotate, variant="flash"):
    """layers.py:74-91: odd blocks run on the rotated permutation and are permuted back."""
    inv = torch.argsort(rot) if (rotate and depth > 1) else None
    for i in range(depth):
        if rotate and i % 2 == 1:
            x = _block(x[rot], pos[rot], sd, f"{pre}.blocks.{i}", num_heads, ball_size, variant)[inv]
        else:
            x = _block(x, pos, sd, f"{pre}.blocks.{i}", num_heads, ball_size, variant)
    return x


def erwin_forward(sd: Dict[str, torch.Tensor], cfg: dict, x: torch.Tensor, pos: torch.Tensor,
                  batch_idx: torch.Tensor, tree=None, bn_training: bool = False, taps: Optional[dict] = None):
    """erwin_flash.py:138-195 with embed=False, mp_steps=0 (the only mode HAET uses).

    ``sd`` holds the ``erwin.*`` entries with the prefix stripped.  ``tree`` may carry a
    pre-built (tree_idx, tree_mask, rot_list) triple."""
    strides, balls = cfg["strides"], cfg["ball_sizes"]
    rotate = cfg["rotate"] > 0
    var = cfg.get("variant", "flash")
    if tree is None:
        tree = balltree_build_with_rotations(pos, batch_idx, strides, balls, cfg["rotate"])
    tree_idx, tree_mask, rots = tree
    rots = list(rots)
    if taps is not None:
        taps.update(tree_idx=tree_idx, tree_mask=tree_mask, rot_idx=list(rots))
    nlev = len(cfg["enc_depths"]) - 1
    xs, ps, rs = [], [], []                      # encoder skip state (node.children chain)
    x, p = x[tree_idx], pos[tree_idx]            # :176-181
    for i in range(nlev):                        # encoder :183-185
        rot = rots.pop(0)
        x = _blocks(x, p, rot, sd, f"encoder.{i}", cfg["enc_depths"][i], cfg["enc_num_heads"][i], balls[i], rotate, var)
        xs.append(x); ps.append(p); rs.append(rot)
        s = strides[i]                           # BallPooling pooling.py:25-40
        if s != 1:
            n = x.shape[0] // s
            centers = p.view(n, s, -1).mean(1)
            rel = (p.view(n, s, -1) - centers[:, None]).reshape(n, -1)
            cat = torch.cat([x.view(n, -1), rel], dim=1)
            x = _bn(F.linear(cat, sd[f"encoder.{i}.pool.proj.weight"], sd[f"encoder.{i}.pool.proj.bias"]),
                    sd, f"encoder.{i}.pool.norm", bn_training)
            p = centers
    rot = rots.pop(0)                            # bottleneck :187-188
    x = _blocks(x, p, rot, sd, "bottleneck", cfg["enc_depths"][-1], cfg["enc_num_heads"][-1], balls[-1], rotate, var)
    if not cfg.get("decode", True):
        return x
    for j, i in enumerate(range(nlev - 1, -1, -1)):   # decoder :190-192
        s = strides[i]                           # BallUnpooling pooling.py:58-72
        cx, cp, crot = xs[i], ps[i], rs[i]
        n = x.shape[0]
        rel = (cp.view(n, s, -1) - p[:, None]).reshape(n, -1)
        up = F.linear(torch.cat([x, rel], dim=-1), sd[f"decoder.{j}.unpool.proj.weight"],
                      sd[f"decoder.{j}.unpool.proj.bias"]).view(n * s, -1)
        x = _bn(cx + up, sd, f"decoder.{j}.unpool.norm", bn_training)
        p = cp
        x = _blocks(x, p, crot, sd, f"decoder.{j}", cfg["dec_depths"][i], cfg["dec_num_heads"][i], balls[i], rotate, var)
    return x[tree_mask][torch.argsort(tree_idx[tree_mask])]   # :193


# --------------------------------------------------------------------------------------
# PhysicsAttention layer (slicing -> Erwin -> deslicing)
# --------------------------------------------------------------------------------------
def split_state_dict(sd: Dict[str, torch.Tensor], dtype=torch.float32):
    """Split a Physics_Attention_* state_dict into (own params, erwin.* params w/o prefix)."""
    own, erw = {}, {}
    for k, v in sd.items():
        v = v.detach().cpu()
        if v.is_floating_point():
            v = v.to(dtype)
        if k.startswith("erwin."):
            erw[k[len("erwin."):]] = v
        else:
            own[k] = v
    return own, erw


def in_project(own, x, heads, dim_head, grid=None):
    """IrregularMesh.py:157-164 (Linear) / StructuredMesh2D.py:173-189 / StructuredMesh3D.py:177-193
    (Conv2d/Conv3d k=3 pad=1 on the reshaped grid).  Returns x_proj [B, H, N, C]."""
    B, N, D = x.shape
    w, b = own["in_project_x.weight"], own["in_project_x.bias"]
    if w.dim() == 2:
        xp = F.linear(x, w, b)
    elif w.dim() == 4:
        H, W = grid
        xg = x.reshape(B, H, W, D).permute(0, 3, 1, 2)
        xp = F.conv2d(xg, w, b, stride=1, padding=w.shape[-1] // 2).permute(0, 2, 3, 1).reshape(B, N, -1)
    else:
        H, W, Dz = grid
        xg = x.reshape(B, H, W, Dz, D).permute(0, 4, 1, 2, 3)
        xp = F.conv3d(xg, w, b, stride=1, padding=w.shape[-1] // 2).permute(0, 2, 3, 4, 1).reshape(B, N, -1)
    return xp.reshape(B, N, heads, dim_head).permute(0, 2, 1, 3).contiguous()


def slice_weights_and_states(own, x_proj, base_temp=0.5, epsilon=1e-6):
    """IrregularMesh.py:168-188.  Returns (W [B,H,N,G], S [B,H,G,C], tau [B,H,N,1])."""
    tau = base_temp + F.linear(x_proj, own["ada_temp_linear.weight"], own["ada_temp_linear.bias"]).clamp(-0.4, 0.4)
    c = torch.log(-torch.log(torch.tensor(epsilon, dtype=torch.float32))).to(x_proj.dtype)   # :172 (fp32 scalar)
    logits = F.linear(x_proj, own["in_project_slice.weight"], own["in_project_slice.bias"]) - c
    W = torch.softmax(logits / tau, dim=2)                       # over the N points
    mass = W.sum(2)                                              # :177
    S = torch.einsum("bhnc,bhng->bhgc", x_proj, W)               # :181-183
    S = S / (mass + 1e-5)[:, :, :, None]                         # :186-188
    return W, S, tau


def slice_positions(S_flat, dimensionality):
    """IrregularMesh.py:204-209 -- min/max over ALL rows jointly, first d channels."""
    fmin = S_flat.min(dim=0, keepdim=True)[0]
    fmax = S_flat.max(dim=0, keepdim=True)[0]
    rng = fmax - fmin + 1e-8
    d = dimensionality
    return (S_flat[:, :d] - fmin[:, :d]) / rng[:, :d]


def physics_attention_forward(sd: Dict[str, torch.Tensor], x: torch.Tensor, *, heads: int, dim_head: int,
                              slice_num: int, dimensionality: int = 3, grid=None, base_temp: float = 0.5,
                              epsilon: float = 1e-6, erwin_cfg: Optional[dict] = None, mlp_ratio: int = 4,
                              rotate=45, dtype=torch.float32, bn_training: bool = False,
                              taps: Optional[dict] = None, tree=None, pos_override=None):
    """Whole layer: Physics_Attention_*.forward (IrregularMesh.py:141-235).

    ``taps`` (optional dict) receives the intermediates x_proj, tau, W (only when N*G is small),
    S, pos, tree_idx, tree_mask, rot_idx, S_out.  ``tree`` / ``pos_override`` let a test feed the
    exact positions / tree the device path used (SURVEY 7.3 hard part 2)."""
    own, erw = split_state_dict(sd, dtype)
    x = x.detach().cpu().to(dtype)
    B, N, _ = x.shape
    cfg = erwin_cfg or default_erwin_cfg(heads, dim_head, slice_num, dimensionality, rotate=rotate,
                                         mlp_ratio=mlp_ratio)
    xp = in_project(own, x, heads, dim_head, grid)
    W, S, tau = slice_weights_and_states(own, xp, base_temp, epsilon)
    G = S.shape[2]
    S_flat = S.reshape(B * heads * G, dim_head)
    pos = slice_positions(S_flat, dimensionality) if pos_override is None else pos_override.to(dtype)
    batch_idx = torch.arange(B * heads).repeat_interleave(G)     # :213
    et = {} if taps is not None else None
    # the tree is always built from the fp32 value of pos (balltree.pyx:542 promotes fp32 -> fp64)
    S_out = erwin_forward(erw, cfg, S_flat, pos, batch_idx, tree=tree, bn_training=bn_training, taps=et)
    out_x = torch.einsum("bhgc,bhng->bhnc", S_out.reshape(B, heads, G, dim_head), W)   # :228
    out_x = out_x.permute(0, 2, 1, 3).reshape(B, N, heads * dim_head)                  # :232
    out = F.linear(out_x, own["to_out.0.weight"], own["to_out.0.bias"])                # :235
    if taps is not None:
        taps.update(x_proj=xp, tau=tau, S=S, pos=pos, S_out=S_out, **et)
        if N * G * B * heads <= (1 << 26):
            taps["W"] = W
    return out


# ---- the rest of TransolverErwinBlock around the layer ------------------------------------------------
def block_attention_half(sd, fx, ln_w, ln_b, ln_eps=1e-5, **layer_kw):
    """models/HAETransolver_Irregular_Mesh.py:125  fx = self.Attn(self.ln_1(fx)) + fx"""
    xn = F.layer_norm(fx.float(), (fx.shape[-1],), ln_w.float(), ln_b.float(), ln_eps)
    return physics_attention_forward(sd, xn, **layer_kw) + fx.float()


def block_mlp_half(fx, ln_w, ln_b, w1, b1, w2, b2, ln_eps=1e-5, dtype=torch.float64):
    """models/HAETransolver_Irregular_Mesh.py:128  fx = self.mlp(self.ln_2(fx)) + fx  with
    models/components/mlp.py:67-96 (n_layers = 0, res = False): linear_pre = Linear + nn.GELU() (exact erf), linear_post."""
    x = fx.to(dtype)
    xn = F.layer_norm(x, (x.shape[-1],), ln_w.to(dtype), ln_b.to(dtype), ln_eps)
    h = F.gelu(F.linear(xn, w1.to(dtype), b1.to(dtype)))
    return (F.linear(h, w2.to(dtype), b2.to(dtype)) + x).float()


# ---- streamed restatement for the BASELINE-size meshes (W is never materialised) -------------------------------------
def in_project_taps(own, x, heads, dim_head, grid, dtype=torch.float64):
    """Same map as ``in_project`` for the conv front-ends (StructuredMesh2D.py:173-189 / StructuredMesh3D.py:177-193),
    evaluated tap by tap (sum over the k^nd taps of a shifted [N, D] x [D, inner] product) so that a 64^3 x 256 grid needs
    no im2col buffer (F.conv3d on the CPU would allocate N x 27 D doubles).  Returns x_proj [B, N, inner]."""
    w, b = own["in_project_x.weight"].to(dtype), own["in_project_x.bias"].to(dtype)
    B, N, D = x.shape
    if w.dim() == 2:
        return F.linear(x.to(dtype), w, b)
    k = w.shape[-1]
    p = k // 2
    g = tuple(grid) + (1,) * (3 - len(grid))
    nd = len(grid)
    xg = x.to(dtype).reshape(B, g[0], g[1], g[2], D)
    xpad = F.pad(xg, (0, 0, p if nd == 3 else 0, p if nd == 3 else 0, p, p, p, p))   # zero padding = conv padding
    out = b.expand(B, g[0], g[1], g[2], -1).clone()
    for i in range(k):
        for j in range(k):
            for l in range(k if nd == 3 else 1):
                wt = w[:, :, i, j, l] if nd == 3 else w[:, :, i, j]
                out += xpad[:, i:i + g[0], j:j + g[1], l:l + g[2], :] @ wt.t()
    return out.reshape(B, N, -1)


def streamed_slice_state(sd, x, *, heads, dim_head, grid=None, base_temp=0.5, epsilon=1e-6, chunk=16384,
                         dtype=torch.float64):
    """IrregularMesh.py:157-188 on meshes of any size: the points stream through in chunks carrying the running column
    state (m, l, sum e^{z-m} x_proj) of the softmax over the N points, exactly what a two-pass evaluation of
    torch.softmax(dim=2) followed by the einsum computes.  Returns
        x_proj [B, N, H*C], tau [B, N, H], S [B, H, G, C] (= W^T x / (mass + 1e-5), mass = 1 in exact arithmetic, F2),
        lse [B, H, G] = log sum_n exp(z)  (natural log)."""
    own, _ = split_state_dict(sd, dtype)
    x = x.detach().cpu()
    B, N, _ = x.shape
    xp = in_project_taps(own, x, heads, dim_head, grid, dtype) if own["in_project_x.weight"].dim() > 2 else \
        torch.cat([F.linear(x[:, i:i + chunk].to(dtype), own["in_project_x.weight"], own["in_project_x.bias"])
                   for i in range(0, N, chunk)], dim=1)
    ws, bs = own["in_project_slice.weight"], own["in_project_slice.bias"]
    wt, bt = own["ada_temp_linear.weight"], own["ada_temp_linear.bias"]
    c = torch.log(-torch.log(torch.tensor(epsilon, dtype=torch.float32))).to(dtype)          # :172 (fp32 scalar)
    G, C = ws.shape
    m = torch.full((B, heads, G), -float("inf"), dtype=dtype)
    l = torch.zeros(B, heads, G, dtype=dtype)
    S = torch.zeros(B, heads, G, C, dtype=dtype)
    tau_all = torch.empty(B, N, heads, dtype=dtype)
    for i in range(0, N, chunk):
        xc = xp[:, i:i + chunk].reshape(B, -1, heads, C)                                        # [B, n, H, C]
        tau = base_temp + F.linear(xc, wt, bt).clamp(-0.4, 0.4)                                 # :168  [B, n, H, 1]
        tau_all[:, i:i + chunk] = tau[..., 0]
        z = (F.linear(xc, ws, bs) - c) / tau                                                    # :173-174  [B, n, H, G]
        mc = z.amax(dim=1)                                                                      # [B, H, G]
        mn = torch.maximum(m, mc)
        sc = torch.exp(m - mn)
        sc[m == -float("inf")] = 0
        e = torch.exp(z - mn[:, None])
        l = l * sc + e.sum(1)
        S = S * sc[..., None] + torch.einsum("bnhg,bnhc->bhgc", e, xc)
        m = mn
    return dict(x_proj=xp, tau=tau_all, S=S / l[..., None] / (1 + 1e-5), lse=m + torch.log(l))


def streamed_output_rows(sd, st, S_out, rows, *, heads, dim_head, epsilon=1e-6, dtype=torch.float64):
    """IrregularMesh.py:221-235 for the sampled point rows ``rows`` (indices into N) of every batch item:
    W[n, g] = exp(z[n, g] - lse[g]) with the global normaliser from ``streamed_slice_state``, out = to_out(W . S')."""
    own, _ = split_state_dict(sd, dtype)
    xp, tau, lse = st["x_proj"], st["tau"], st["lse"]
    B = xp.shape[0]
    ws, bs = own["in_project_slice.weight"], own["in_project_slice.bias"]
    G, C = ws.shape
    c = torch.log(-torch.log(torch.tensor(epsilon, dtype=torch.float32))).to(dtype)
    xc = xp[:, rows].reshape(B, -1, heads, C)
    z = (F.linear(xc, ws, bs) - c) / tau[:, rows][..., None]
    W = torch.exp(z - lse[:, None])                                                            # [B, n, H, G]
    o = torch.einsum("bnhg,bhgc->bnhc", W, S_out.to(dtype).reshape(B, heads, G, C)).reshape(B, len(rows), heads * C)
    return F.linear(o, own["to_out.0.weight"], own["to_out.0.bias"])


# ---- Model.forward around the blocks --------------------------------------------------------------------
def timestep_embedding(timesteps, dim, max_period=10000):
    """models/components/embedding.py:72-99"""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(start=0, end=half, dtype=torch.float32) / half)
    args = timesteps[:, None].float() * freqs[None]
    emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    if dim % 2:
        emb = torch.cat([emb, torch.zeros_like(emb[:, :1])], dim=-1)
    return emb


def grid_distances(x, ref):
    """get_grid of the irregular model (HAETransolver_Irregular_Mesh.py:297-336): distances of the 2-D points x [B, N, 2]
    to the ref x ref reference grid on [0, 1]^2, reference point k = (k // ref, k % ref)."""
    g = torch.tensor(np.linspace(0, 1, ref), dtype=torch.float)
    gx = g.reshape(1, ref, 1, 1).repeat([1, 1, ref, 1])
    gy = g.reshape(1, 1, ref, 1).repeat([1, ref, 1, 1])
    gref = torch.cat((gx, gy), dim=-1).reshape(1, ref * ref, 2)
    return torch.sqrt(torch.sum((x[:, :, None, :] - gref[:, None, :, :]) ** 2, dim=-1))


def structured_grid_distances(sizes, ref):
    """get_grid of the structured models (HAETransolver_Structured_Mesh_2D.py:330-382; _3D.py:340-416) -> [1, N, ref^nd]."""
    nd = len(sizes)
    axes = [torch.tensor(np.linspace(0, 1, s), dtype=torch.float) for s in sizes]
    raxes = [torch.tensor(np.linspace(0, 1, ref), dtype=torch.float) for _ in sizes]
    grid = torch.stack(torch.meshgrid(*axes, indexing="ij"), dim=-1).reshape(-1, 1, nd)
    gref = torch.stack(torch.meshgrid(*raxes, indexing="ij"), dim=-1).reshape(1, -1, nd)
    return torch.sqrt(torch.sum((grid - gref) ** 2, dim=-1))[None]


def _mlp0(sd, pre, x):
    """components/mlp.py:75-96 with n_layers = 0, act = gelu: linear_post(gelu(linear_pre(x)))"""
    return F.linear(F.gelu(F.linear(x, sd[pre + ".linear_pre.0.weight"], sd[pre + ".linear_pre.0.bias"])),
                    sd[pre + ".linear_post.weight"], sd[pre + ".linear_post.bias"])


def model_forward(sd, x, fx, T=None, *, variant="irregular", n_layers, n_head, slice_num, mlp_ratio=1, ref=8,
                  unified_pos=False, grid=None, rotate=45, layer_hook=None):
    """Model.forward(x, fx, T) of HAETransolver_Irregular_Mesh (:338-376) / _Structured_Mesh_2D (:384-410) / _3D (:418-461)
    driven by the model's state_dict (fp32, CPU).  ``layer_hook(i)`` may return dict(tree=..., pos_override=...) for block i
    (the device's positions / tree, SURVEY 7.3 hard part 2)."""
    sd = {k: v.detach().cpu().float() if v.is_floating_point() else v.detach().cpu() for k, v in sd.items()}
    x = x.detach().cpu().float()
    B = x.shape[0]
    structured = variant != "irregular"
    if unified_pos:
        x = structured_grid_distances(grid, ref).expand(B, -1, -1) if structured else grid_distances(x, ref)
    N = x.shape[1]
    if fx is not None:
        h = _mlp0(sd, "preprocess", torch.cat((x, fx.detach().cpu().float()), -1))
        if not structured:
            h = h + sd["placeholder"][None, None, :]
    else:
        h = _mlp0(sd, "preprocess", x) + sd["placeholder"][None, None, :]
    n_hidden = h.shape[-1]
    if T is not None:
        emb = timestep_embedding(T.detach().cpu(), n_hidden).repeat(1, N, 1)
        h = h + F.linear(F.silu(F.linear(emb, sd["time_fc.0.weight"], sd["time_fc.0.bias"])), sd["time_fc.2.weight"],
                         sd["time_fc.2.bias"])
    dim_head = n_hidden // n_head
    dimensionality = 2 if variant == "2d" else 3
    for i in range(n_layers):
        p = f"blocks.{i}."
        attn_sd = {k[len(p) + 5:]: v for k, v in sd.items() if k.startswith(p + "Attn.")}
        kw = dict(heads=n_head, dim_head=dim_head, slice_num=slice_num, dimensionality=dimensionality, grid=grid,
                  rotate=rotate, mlp_ratio=mlp_ratio)
        if layer_hook is not None:
            kw.update(layer_hook(i) or {})
        h = block_attention_half(attn_sd, h, sd[p + "ln_1.weight"], sd[p + "ln_1.bias"], **kw)           # :125
        h = block_mlp_half(h, sd[p + "ln_2.weight"], sd[p + "ln_2.bias"], sd[p + "mlp.linear_pre.0.weight"],
                           sd[p + "mlp.linear_pre.0.bias"], sd[p + "mlp.linear_post.weight"],
                           sd[p + "mlp.linear_post.bias"], dtype=torch.float32)                             # :128
        if i == n_layers - 1:                                                                                # :131-132
            h = F.linear(F.layer_norm(h, (n_hidden,), sd[p + "ln_3.weight"], sd[p + "ln_3.bias"], 1e-5),
                         sd[p + "mlp2.weight"], sd[p + "mlp2.bias"])
    return h
