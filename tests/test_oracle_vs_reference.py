"""Where /root/reference exists (the build container) run the UNMODIFIED reference beside the
oracle.  Skipped on the GPU box, where only the golden vectors pin the oracle."""
import itertools

import pytest
import torch

from oracle import haet_oracle as O
from oracle import ref_loader
from tests.helpers import rel_l2, trained_like_

pytestmark = pytest.mark.skipif(not ref_loader.reference_available(), reason="reference sources absent")


@pytest.mark.parametrize("d,nt,G", list(itertools.product([2, 3], [1, 8], [2, 3, 5, 32, 100, 256, 512])))
def test_balltree_oracle_vs_reference_extension(d, nt, G):
    bt = ref_loader.load_reference_balltree()
    g = torch.Generator().manual_seed(d * 1000 + nt * 10 + G)
    for angle, mode in itertools.product([0.0, 45.0, 30.0], ["rand", "ties"]):
        T = nt * G
        pos = torch.rand(T, d, generator=g) if mode == "rand" else torch.randint(0, 4, (T, d), generator=g).float() / 4
        bidx = torch.arange(nt).repeat_interleave(G)
        balls = [min(32, G), min(16, max(G // 2, 1))]
        r = bt.build_balltree_with_rotations(pos, bidx, [2], balls, angle)
        o = O.balltree_build_with_rotations(pos, bidx, [2], balls, angle)
        assert torch.equal(r[0], o[0]) and torch.equal(r[1], o[1])
        for a, b in zip(r[2], o[2]):
            assert (a is None and b is None) or torch.equal(a, b)


def test_rotation_matmul_is_forward_fma_chain():
    """The oracle's fp32 rotation (fmaf chain) is bit-identical to the torch CPU matmul the reference runs."""
    import ctypes
    import numpy as np
    bt = ref_loader.load_reference_balltree()
    lib = O._lib()
    for d in (2, 3):
        a = torch.rand(4096, d, generator=torch.Generator().manual_seed(d))
        R = bt.generate_rotation_matrix(45.0, d, torch.device("cpu"))
        want = torch.matmul(a, R).numpy()
        Rm = np.zeros(9, dtype=np.float32)
        lib.haet_oracle_rotation_matrix(ctypes.c_double(45.0), ctypes.c_int(d), O._p(Rm, ctypes.c_float))
        assert np.array_equal(Rm[:d * d].reshape(d, d), R.numpy())
        got = np.empty((4096, d), dtype=np.float32)
        an = np.ascontiguousarray(a.numpy())
        lib.haet_oracle_rotate(O._p(an, ctypes.c_float), ctypes.c_int64(4096), ctypes.c_int(d),
                               O._p(Rm, ctypes.c_float), O._p(got, ctypes.c_float))
        assert np.array_equal(got, want)


@pytest.mark.parametrize("variant", ["irregular", "2d", "3d"])
def test_layer_oracle_vs_reference_module(variant):
    ref = ref_loader.load_reference()
    torch.manual_seed(3)
    if variant == "irregular":
        m = ref.PhysicsAttention.Physics_Attention_Irregular_Mesh(64, heads=4, dim_head=16, slice_num=64,
                                                                  rotate=45, mlp_ratio=2)
        kw = dict(heads=4, dim_head=16, slice_num=64, dimensionality=3, rotate=45, mlp_ratio=2)
        B, N = 2, 777
    elif variant == "2d":
        m = ref.PhysicsAttention.Physics_Attention_Structured_Mesh_2D(32, heads=4, dim_head=8, slice_num=32, H=11,
                                                                      W=9, rotate=45, mlp_ratio=2)
        kw = dict(heads=4, dim_head=8, slice_num=32, dimensionality=2, grid=(11, 9), rotate=45, mlp_ratio=2)
        B, N = 2, 99
    else:
        m = ref.PhysicsAttention.Physics_Attention_Structured_Mesh_3D(32, heads=4, dim_head=8, slice_num=32, H=5,
                                                                      W=4, D=6, rotate=45, mlp_ratio=2)
        kw = dict(heads=4, dim_head=8, slice_num=32, dimensionality=3, grid=(5, 4, 6), rotate=45, mlp_ratio=2)
        B, N = 1, 120
    m.eval()
    trained_like_(m, 5)
    x = torch.randn(B, N, m.in_project_x.weight.shape[1])
    with torch.no_grad():
        want = m(x)
    got = O.physics_attention_forward(m.state_dict(), x, **kw)
    assert rel_l2(got, want) < 2e-6
