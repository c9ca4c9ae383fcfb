"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol include/haet_b200.h
declares, the ctypes structs have the C layout, the host-side sizing logic is consistent, and the drop-in
modules are state_dict-compatible with the reference.  No compute calls (there is no GPU here)."""
import ctypes as C
import os
import re
import subprocess
import sys
import tempfile

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "haet_b200.h")


@pytest.fixture(scope="module")
def lib():
    from haet_b200 import _lib, build
    build.build()
    return _lib.lib()


def test_library_exports_every_declared_symbol(lib):
    from haet_b200 import _lib
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    declared = set(re.findall(r"\b(haet_[a-z0-9_]+)\s*\(", src))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.SYMBOLS), "ctypes binding table and header disagree"
    assert lib.haet_version() == 3


def test_struct_layout_matches_c_header():
    from haet_b200 import _lib
    prog = r'''
#include <stdio.h>
#include <stddef.h>
#include "haet_b200.h"
int main(void) {
  printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(haet_layer_shape), offsetof(haet_layer_shape, grid),
         offsetof(haet_layer_shape, base_temp), sizeof(haet_erwin_cfg), offsetof(haet_erwin_cfg, rotate),
         sizeof(haet_layer_weights), offsetof(haet_layer_weights, bt));
  return 0; }'''
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "t.c")
        open(c, "w").write(prog)
        exe = os.path.join(d, "t")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        got = [int(v) for v in subprocess.check_output([exe]).split()]
    want = [C.sizeof(_lib.LayerShape), _lib.LayerShape.grid.offset, _lib.LayerShape.base_temp.offset,
            C.sizeof(_lib.ErwinCfg), _lib.ErwinCfg.rotate.offset, C.sizeof(_lib.LayerWeights),
            _lib.LayerWeights.bt.offset]
    assert got == want


def test_sizing_functions(lib):
    from haet_b200 import _lib
    sh = _lib.make_shape(1, 1_000_000, 256, 8, 32, 256)
    assert lib.haet_partial_floats(C.byref(sh)) == 8 * 256 * 34
    assert lib.haet_spill_bytes(C.byref(sh)) >= 1_000_000 * 256 * 2
    assert lib.haet_balltree_slots(8, 256) == 8 * 256
    assert lib.haet_balltree_slots(3, 100) == 3 * 128
    assert lib.haet_balltree_slots(1, 1) == -1
    cfg = _lib.make_erwin_cfg([32, 64], [32, 16], [4, 8], [2, 2], [4], [2], [2], 45, True, 2, 3)
    assert lib.haet_erwin_weight_floats(C.byref(cfg)) > 0
    assert lib.haet_layer_scratch_bytes(C.byref(sh), C.byref(cfg)) > lib.haet_spill_bytes(C.byref(sh))
    with pytest.raises(AssertionError):
        _lib.make_erwin_cfg([32, 64], [32, 16], [4, 8], [2, 2], [4], [2], [2, 2], 45)


def test_no_gpu_means_loud_failure(lib):
    import haet_b200 as hb
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    assert lib.haet_device_check() == -3
    m = hb.Physics_Attention_Irregular_Mesh(64, heads=4, dim_head=16, slice_num=32).eval()
    with pytest.raises(hb.HaetError):
        m(torch.randn(1, 8, 64))


@pytest.mark.parametrize("name", ["layer_irregular_small", "layer_2d_small", "layer_3d_small"])
def test_state_dict_roundtrip_with_reference_names(name):
    import haet_b200 as hb
    from tests.helpers import load_layer_golden
    cfg, sd, x, out = load_layer_golden(name)
    m = getattr(hb, cfg["cls"])(**cfg["kwargs"])
    res = m.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    mine = m.state_dict()
    assert list(mine.keys()) == list(sd.keys())
    for k in sd:
        assert mine[k].shape == sd[k].shape and mine[k].dtype == sd[k].dtype, k


def test_from_reference_and_patch_model():
    from oracle import ref_loader
    if not ref_loader.reference_available():
        pytest.skip("reference sources absent")
    import haet_b200 as hb
    ref = ref_loader.load_reference()
    torch.manual_seed(0)
    model = ref.HAETransolver_Irregular_Mesh(space_dim=3, n_layers=2, n_hidden=64, n_head=4, fun_dim=0, out_dim=1,
                                                   slice_num=32, mlp_ratio=1, unified_pos=False)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    hb.patch_model(model)
    assert all(isinstance(b.Attn, hb.Physics_Attention_Irregular_Mesh) for b in model.blocks)
    assert list(model.state_dict().keys()) == list(sd.keys())
    for k, v in model.state_dict().items():
        assert torch.equal(v, sd[k]), k
    a = model.blocks[0].Attn
    assert a.erwin.rotate == 45 and a.erwin.mlp_ratio == 1 and a.heads == 4 and a.dim_head == 16
