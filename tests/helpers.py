"""Shared helpers for the test-suite (golden loading, trained-like weights, error metrics)."""
from __future__ import annotations

import glob
import json
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_layer_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    cfg = json.loads(bytes(z["cfg"]).decode())
    sd = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd/")}
    return cfg, sd, torch.from_numpy(z["x"]), torch.from_numpy(z["out"])


def balltree_goldens():
    return sorted(glob.glob(os.path.join(GOLDEN, "balltree_*.npz")))


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    """||a-b||_2 / ||b||_2 in fp64 -- THE relative error used for the 1e-3 tolerance."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def rel_max(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-300))


def oracle_kwargs(cfg):
    kw = cfg["kwargs"]
    cls = cfg["cls"]
    out = dict(heads=kw["heads"], dim_head=kw["dim_head"], slice_num=kw["slice_num"],
               rotate=kw.get("rotate", 1), mlp_ratio=kw.get("mlp_ratio", 4))
    if cls.endswith("2D"):
        out.update(dimensionality=2, grid=(kw["H"], kw["W"]))
    elif cls.endswith("3D"):
        out.update(dimensionality=3, grid=(kw["H"], kw["W"], kw["D"]))
    else:
        out.update(dimensionality=3)
    return out


def trained_like_(module, seed):
    """Same recipe as tests/golden/make_golden.py::trained_like_ (peaky softmax, tau in both clamps)."""
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


def load_model_golden(name):
    """(cfg dict(variant, kwargs), state_dict, x, fx | None, T | None, out) of a whole-model golden (make_golden.save_model)."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    cfg = json.loads(bytes(z["cfg"]).decode())
    sd = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd/")}
    t = lambda k: torch.from_numpy(z[k]) if k in z.files else None
    return cfg, sd, t("x"), t("fx"), t("T"), t("out")


def model_oracle_kwargs(cfg):
    kw = cfg["kwargs"]
    out = dict(variant=cfg["variant"], n_layers=kw["n_layers"], n_head=kw["n_head"], slice_num=kw["slice_num"],
               mlp_ratio=kw.get("mlp_ratio", 1), ref=kw.get("ref", 8), unified_pos=kw.get("unified_pos", False),
               rotate=kw.get("rotate", 45))
    if cfg["variant"] == "2d":
        out["grid"] = (kw["H"], kw["W"])
    elif cfg["variant"] == "3d":
        out["grid"] = (kw["H"], kw["W"], kw["D"])
    return out
