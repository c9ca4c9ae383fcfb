"""Generate the committed golden vectors from the UNMODIFIED reference.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
It imports the reference through oracle/ref_loader.py (shims documented there), runs it on
seeded inputs on the CPU in fp32 and stores inputs, weights and outputs as small .npz files.
The GPU box never runs this script; tests only read the .npz files.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def trained_like_(module, seed):
    """Make the softmax over points peaky and drive tau into both clamps:
    scale in_project_slice / ada_temp_linear, randomise biases and BatchNorm running stats."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for n, p in module.named_parameters():
            if n.startswith("in_project_slice.weight"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.6)
            elif n.startswith("ada_temp_linear.weight"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.5)
            elif n.endswith(".bias") and "norm" not in n:
                p.copy_(torch.randn(p.shape, generator=g) * 0.1)
            elif n.startswith("erwin.") and n.endswith(".weight") and p.dim() == 2:
                p.copy_(torch.randn(p.shape, generator=g) * (1.0 / p.shape[1] ** 0.5))
            elif "norm" in n and n.endswith(".weight"):
                p.copy_(1 + 0.2 * torch.randn(p.shape, generator=g))
            elif "norm" in n and n.endswith(".bias"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
        for n, b in module.named_buffers():
            if n.endswith("running_mean"):
                b.copy_(0.2 * torch.randn(b.shape, generator=g))
            elif n.endswith("running_var"):
                b.copy_(0.5 + torch.rand(b.shape, generator=g))


def save_layer(name, cls_name, kwargs, B, N, seed, trained=True):
    ref = ref_loader.load_reference()
    torch.manual_seed(seed)
    m = getattr(ref.PhysicsAttention, cls_name)(**kwargs).eval()
    if trained:
        trained_like_(m, seed + 1)
    x = torch.randn(B, N, kwargs["dim"])
    x = torch.nn.functional.layer_norm(x, (kwargs["dim"],))
    with torch.no_grad():
        out = m(x)
    blob = {"x": x.numpy(), "out": out.numpy(),
            "cfg": np.frombuffer(json.dumps({"cls": cls_name, "kwargs": kwargs}).encode(), dtype=np.uint8)}
    for k, v in m.state_dict().items():
        blob["sd/" + k] = v.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **blob)
    print(name, "out absmax", float(out.abs().max()), "bytes", os.path.getsize(os.path.join(OUT, name + ".npz")))


def save_layer_train(name, cls_name, kwargs, B, N, seed):
    """The reference layer in TRAIN mode (BatchNorm1d batch statistics in BallPooling / BallUnpooling,
    pooling.py:23,56; dropout 0): weights BEFORE the call, the output, and the BatchNorm buffers AFTER it."""
    ref = ref_loader.load_reference()
    torch.manual_seed(seed)
    m = getattr(ref.PhysicsAttention, cls_name)(**kwargs)
    trained_like_(m, seed + 1)
    x = torch.nn.functional.layer_norm(torch.randn(B, N, kwargs["dim"]), (kwargs["dim"],))
    blob = {"x": x.numpy(), "cfg": np.frombuffer(json.dumps({"cls": cls_name, "kwargs": kwargs}).encode(), dtype=np.uint8)}
    for k, v in m.state_dict().items():
        blob["sd/" + k] = v.clone().numpy()
    m.train()
    with torch.no_grad():
        out = m(x)
    blob["out"] = out.numpy()
    for k, v in m.state_dict().items():
        if "running_" in k or "num_batches_tracked" in k:
            blob["after/" + k] = v.clone().numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **blob)
    print(name, "out absmax", float(out.abs().max()), "bytes", os.path.getsize(os.path.join(OUT, name + ".npz")))


def save_balltree(name, d, n_trees, G, angle, mode, seed):
    bt = ref_loader.load_reference_balltree()
    g = torch.Generator().manual_seed(seed)
    T = n_trees * G
    if mode == "rand":
        pos = torch.rand(T, d, generator=g)
    elif mode == "ties":
        pos = torch.randint(0, 4, (T, d), generator=g).float() / 4
    else:  # sorted / organ-pipe: stresses the introselect pivoting
        base = torch.arange(G).float() / G
        col = torch.cat([base[::2], base[1::2].flip(0)]) if mode == "pipe" else base
        pos = col.repeat(n_trees)[:, None].repeat(1, d).contiguous()
    bidx = torch.arange(n_trees).repeat_interleave(G)
    balls = [min(32, G), min(16, max(G // 2, 1))]
    tree_idx, tree_mask, rots = bt.build_balltree_with_rotations(pos, bidx, [2], balls, angle)
    blob = {"pos": pos.numpy(), "n_trees": np.int64(n_trees), "G": np.int64(G), "angle": np.float64(angle),
            "ball_sizes": np.asarray(balls), "strides": np.asarray([2]),
            "tree_idx": tree_idx.numpy(), "tree_mask": tree_mask.numpy()}
    for i, r in enumerate(rots):
        if r is not None:
            blob[f"rot{i}"] = r.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **blob)


def save_erwin_bias(name, seed, T_trees=4, G=64, d=3):
    """The NON-flash Erwin copy (distance bias, (H E K) qkv columns) on slice-token-shaped input:
    the features fed to the encoder (after the copy's always-on input Linear, which HAET's embed=False has no
    counterpart for), positions, the state_dict without that Linear, and the reference output."""
    mod = ref_loader.load_reference_erwin_nonflash()
    cfg = dict(c_in=16, c_hidden=[16, 32], ball_sizes=[32, 16], enc_num_heads=[2, 4], enc_depths=[2, 2],
               dec_num_heads=[2], dec_depths=[2], strides=[2], rotate=45, decode=True, mlp_ratio=2, dimensionality=d,
               mp_steps=0)
    torch.manual_seed(seed)
    m = mod.ErwinTransformer(**cfg).eval()
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for n, p in m.named_parameters():
            if n.endswith(".weight") and p.dim() == 2:
                p.copy_(torch.randn(p.shape, generator=g) * (1.0 / p.shape[1] ** 0.5))
            elif n.endswith(".bias") and "norm" not in n:
                p.copy_(torch.randn(p.shape, generator=g) * 0.1)
            elif n.endswith("sigma_att"):
                p.copy_(-1 + 0.5 * torch.randn(p.shape, generator=g))     # a bias that matters: positions span [0, 1]
            elif "norm" in n and n.endswith(".weight"):
                p.copy_(1 + 0.2 * torch.randn(p.shape, generator=g))
        for n, b in m.named_buffers():
            if n.endswith("running_mean"):
                b.copy_(0.2 * torch.randn(b.shape, generator=g))
            elif n.endswith("running_var"):
                b.copy_(0.5 + torch.rand(b.shape, generator=g))
    T = T_trees * G
    x_in = torch.randn(T, cfg["c_in"], generator=g)
    pos = torch.rand(T, d, generator=g)
    bidx = torch.arange(T_trees).repeat_interleave(G)
    with torch.no_grad():
        x_emb = m.embed(x_in, pos, None)
        out = m(x_in, pos, bidx)
    cfg_o = {k: v for k, v in cfg.items() if k not in ("mp_steps", "c_in")}
    blob = {"x": x_emb.numpy(), "pos": pos.numpy(), "out": out.numpy(), "pts": np.int64(G),
            "cfg": np.frombuffer(json.dumps(cfg_o).encode(), dtype=np.uint8)}
    for k, v in m.state_dict().items():
        if not k.startswith("embed."):
            blob["sd/" + k] = v.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **blob)
    print(name, "out absmax", float(out.abs().max()))


class _cpu_cuda:
    """get_grid hard-codes ``.cuda()`` (HAETransolver_Irregular_Mesh.py:323).  While a golden is generated on
    this GPU-less machine ``Tensor.cuda`` is the identity: a device move, not arithmetic."""

    def __enter__(self):
        self._orig = torch.Tensor.cuda
        torch.Tensor.cuda = lambda t, *a, **k: t

    def __exit__(self, *a):
        torch.Tensor.cuda = self._orig


def save_model(name, variant, kwargs, B, N, seed, with_fx=True, with_T=False):
    """A whole HAETransolver_* forward(x, fx, T) of the unmodified reference: inputs, state_dict, output."""
    ref = ref_loader.load_reference()
    mod = {"irregular": ref.HAETransolver_Irregular_Mesh, "2d": ref.HAETransolver_Structured_Mesh_2D,
           "3d": ref.HAETransolver_Structured_Mesh_3D}[variant]
    torch.manual_seed(seed)
    with _cpu_cuda():
        m = mod(**kwargs).eval()
        g = torch.Generator().manual_seed(seed + 1)
        for blk in m.blocks:
            trained_like_(blk.Attn, int(torch.randint(0, 1 << 30, (1,), generator=g)))
        with torch.no_grad():
            for n, p in m.named_parameters():
                if ".Attn." in n:
                    continue
                if n.endswith(".weight") and p.dim() == 2:
                    p.copy_(torch.randn(p.shape, generator=g) * (1.0 / p.shape[1] ** 0.5))
                elif n.endswith(".bias"):
                    p.copy_(torch.randn(p.shape, generator=g) * 0.1)
                elif n.endswith(".weight") and p.dim() == 1:
                    p.copy_(1 + 0.2 * torch.randn(p.shape, generator=g))
        x = torch.rand(B, N, kwargs["space_dim"], generator=g)
        fx = torch.randn(B, N, kwargs["fun_dim"], generator=g) if with_fx else None
        T = torch.rand(B, 1, generator=g) * 10 if with_T else None
        with torch.no_grad():
            out = m(x, fx, T)
    blob = {"x": x.numpy(), "out": out.numpy(),
            "cfg": np.frombuffer(json.dumps({"variant": variant, "kwargs": kwargs}).encode(), dtype=np.uint8)}
    if fx is not None:
        blob["fx"] = fx.numpy()
    if T is not None:
        blob["T"] = T.numpy()
    for k, v in m.state_dict().items():
        blob["sd/" + k] = v.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **blob)
    print(name, "out absmax", float(out.abs().max()), "bytes", os.path.getsize(os.path.join(OUT, name + ".npz")))


def save_models():
    save_model("model_irregular_small", "irregular",
               dict(space_dim=3, n_layers=2, n_hidden=64, n_head=4, Time_Input=True, mlp_ratio=2, fun_dim=1, out_dim=4,
                    slice_num=32, rotate=45), B=2, N=700, seed=31, with_T=True)
    save_model("model_irregular_256", "irregular",
               dict(space_dim=3, n_layers=2, n_hidden=256, n_head=8, Time_Input=True, mlp_ratio=1, fun_dim=1, out_dim=4,
                    slice_num=32, rotate=45), B=2, N=900, seed=32, with_T=True)
    save_model("model_irregular_256_upos", "irregular",
               dict(space_dim=2, n_layers=1, n_hidden=256, n_head=8, mlp_ratio=1, fun_dim=0, out_dim=3,
                    slice_num=64, ref=8, unified_pos=True, rotate=45), B=1, N=1100, seed=33, with_fx=False)
    save_model("model_2d_small", "2d",
               dict(space_dim=2, n_layers=2, n_hidden=32, n_head=4, mlp_ratio=2, fun_dim=1, out_dim=1, slice_num=32,
                    ref=4, unified_pos=True, H=17, W=13, rotate=45), B=2, N=17 * 13, seed=34)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "train":   # only the train-mode layer golden (added later than the others)
        save_layer_train("layer_irregular_train", "Physics_Attention_Irregular_Mesh",
                         dict(dim=64, heads=4, dim_head=16, slice_num=100, rotate=45, mlp_ratio=2), B=2, N=900, seed=41)
        sys.exit(0)
    save_erwin_bias("erwin_bias_small", seed=21)
    save_models()
    save_layer("layer_irregular_small", "Physics_Attention_Irregular_Mesh",
               dict(dim=64, heads=4, dim_head=16, slice_num=32, rotate=45, mlp_ratio=2), B=2, N=1031, seed=10)
    save_layer("layer_irregular_g64_fresh", "Physics_Attention_Irregular_Mesh",
               dict(dim=64, heads=8, dim_head=8, slice_num=64, rotate=45, mlp_ratio=1), B=1, N=2050, seed=11,
               trained=False)
    save_layer("layer_2d_small", "Physics_Attention_Structured_Mesh_2D",
               dict(dim=32, heads=4, dim_head=8, slice_num=32, H=17, W=13, rotate=45, mlp_ratio=2), B=2, N=17 * 13,
               seed=12)
    save_layer("layer_3d_small", "Physics_Attention_Structured_Mesh_3D",
               dict(dim=32, heads=4, dim_head=8, slice_num=32, H=6, W=5, D=7, rotate=45, mlp_ratio=2), B=1,
               N=6 * 5 * 7, seed=13)
    k = 0
    for d in (2, 3):
        for (nt, G) in ((8, 32), (3, 100), (8, 256), (2, 512), (4, 5)):
            for mode in ("rand", "ties", "pipe"):
                save_balltree(f"balltree_{k:02d}", d, nt, G, 45.0 if k % 4 else 0.0, mode, 100 + k)
                k += 1
    print("balltree goldens:", k)
