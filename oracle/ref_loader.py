"""Import the UNMODIFIED reference (``/root/reference/models``) on CPU.  TEST INFRASTRUCTURE ONLY.

Only usable where ``/root/reference`` exists (the build container).  It is used by
``tests/golden/make_golden.py`` to generate the committed golden vectors and by
``tests/test_oracle_vs_reference.py`` (skipped when the reference is absent) to pin
``oracle/haet_oracle.py``.  Nothing on the GPU box and nothing in ``haet_b200/`` imports it.

Shims, each one only replaces a dependency that is absent or
GPU-only, never the reference's own arithmetic:
  * ``timm.models.layers.trunc_normal_``  -> ``torch.nn.init.trunc_normal_``
    (models/HAETransolver_Irregular_Mesh.py:10)
  * ``torch_cluster`` / ``torch_scatter``  -> empty modules (imported at
    models/components/erwinflash/erwin_flash.py:7 and components/utils/scatter_ops.py:2,
    never executed with embed=False, mp_steps=0)
  * ``flash_attn.flash_attn_qkvpacked_func`` -> fp32 ``softmax(QK^T*scale)V``
    (flash-attn is third-party CUDA, fp16-only, unpinned; call site
    models/components/erwinflash/components/attention.py:9,82-84)
  * ``BallMSA.forward`` (attention.py:42-91) re-stated without the ``.to(torch.float16)``
    weight casts, which make the shipped code unrunnable in fp32.
  * top-level module ``balltree`` -> ``oracle/_ref/balltree*.so`` (the reference's own
    Cython/C++ extension compiled by ``oracle/build_ref.py``).
"""
from __future__ import annotations

import importlib
import importlib.util
import math
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "models"))


def _sdpa_qkvpacked(qkv, dropout_p=0.0, softmax_scale=None, causal=False, **_):
    """qkv: [b, n, 3, h, d] -> [b, n, h, d]; plain softmax attention in the input dtype."""
    import torch

    q, k, v = qkv.unbind(dim=2)  # [b, n, h, d]
    q, k, v = (t.permute(0, 2, 1, 3) for t in (q, k, v))  # [b, h, n, d]
    if softmax_scale is None:
        softmax_scale = 1.0 / math.sqrt(q.shape[-1])
    att = torch.softmax(torch.matmul(q, k.transpose(-1, -2)) * softmax_scale, dim=-1)
    return torch.matmul(att, v).permute(0, 2, 1, 3)


def _install_stubs():
    import torch

    if "timm" not in sys.modules:
        timm = types.ModuleType("timm")
        timm_models = types.ModuleType("timm.models")
        timm_layers = types.ModuleType("timm.models.layers")
        timm_layers.trunc_normal_ = torch.nn.init.trunc_normal_
        timm.models = timm_models
        timm_models.layers = timm_layers
        sys.modules.update({"timm": timm, "timm.models": timm_models, "timm.models.layers": timm_layers})
    for name in ("torch_cluster", "torch_scatter"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    fa = types.ModuleType("flash_attn")
    fa.flash_attn_qkvpacked_func = _sdpa_qkvpacked
    sys.modules["flash_attn"] = fa


def load_reference_balltree():
    """Import the reference's compiled ``balltree`` extension from oracle/_ref/ (or None)."""
    from . import build_ref

    if "balltree" in sys.modules:
        return sys.modules["balltree"]
    so = build_ref.ref_so_path()
    if so is None:
        so = build_ref.build_reference_balltree()
    if so is None:
        return None
    spec = importlib.util.spec_from_file_location("balltree", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules["balltree"] = mod
    return mod


_REF = None


def load_reference():
    """Returns the imported reference ``models`` package (as module ``haet_reference_models``)."""
    global _REF
    if _REF is not None:
        return _REF
    if not reference_available():
        raise RuntimeError("reference sources are not present on this machine")
    _install_stubs()
    if load_reference_balltree() is None:
        raise RuntimeError("reference balltree extension could not be built")
    # import the package under a private name so that it cannot shadow anything of ours
    spec = importlib.util.spec_from_file_location(
        "haet_reference_models", os.path.join(REFERENCE_ROOT, "models", "__init__.py"),
        submodule_search_locations=[os.path.join(REFERENCE_ROOT, "models")])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["haet_reference_models"] = mod
    spec.loader.exec_module(mod)

    # BallMSA.forward without fp16 casts (attention.py:42-91)
    att = importlib.import_module("haet_reference_models.components.erwinflash.components.attention")
    import torch.nn.functional as F
    from einops import rearrange

    def forward_fp32(self, x, pos):
        x = x + self.pe_proj(self.compute_rel_pos(pos))
        qkv = self.qkv(x)
        nb = x.shape[0] // self.ball_size
        qkv = rearrange(qkv, "(b n) (three h d) -> b n three h d", b=nb, n=self.ball_size,
                        three=3, h=self.num_heads, d=self.head_dim)
        o = att.flash_attn_qkvpacked_func(qkv, dropout_p=0.0,
                                          softmax_scale=1.0 / math.sqrt(self.head_dim), causal=False)
        return self.proj(rearrange(o, "b n h d -> (b n) (h d)"))

    att.BallMSA.forward = forward_fp32
    _REF = mod
    return mod


_REF_ERWIN = None
NONFLASH_ROOT = os.path.join(REFERENCE_ROOT, "benchmarks", "02-Car-Design-ShapeNetCar", "models", "components", "erwin")


def load_reference_erwin_nonflash():
    """The reference's NON-flash Erwin copy (benchmarks/02-Car-Design-ShapeNetCar/models/components/erwin/), unmodified:
    SDPA attention with the distance bias sigma_att * cdist (components/attention.py:28-32,49-51) and (H E K) qkv columns.
    Runs on the CPU as shipped once ``torch_cluster`` (imported, never called with mp_steps=0) and ``balltree`` resolve."""
    global _REF_ERWIN
    if _REF_ERWIN is not None:
        return _REF_ERWIN
    if not os.path.isdir(NONFLASH_ROOT):
        raise RuntimeError("reference sources are not present on this machine")
    _install_stubs()
    if load_reference_balltree() is None:
        raise RuntimeError("reference balltree extension could not be built")
    spec = importlib.util.spec_from_file_location(
        "haet_reference_erwin_nonflash", os.path.join(NONFLASH_ROOT, "__init__.py"),
        submodule_search_locations=[NONFLASH_ROOT])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["haet_reference_erwin_nonflash"] = mod
    spec.loader.exec_module(mod)
    _REF_ERWIN = mod
    return mod
