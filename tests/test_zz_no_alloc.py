"""haet_erwin_forward_static: the Erwin pass with its scratch in a file-scope device array of the library (no caller
workspace, no allocation; include/haet_b200.h, csrc/erwin.cu).  It has no counterpart in the reference -- it was added
for callers that cannot allocate -- so the check is against the existing entry
point, haet_erwin_forward_trees with a caller workspace: same kernels, same arithmetic, same order => bitwise equal (the
tolerance the fused-vs-per-block Erwin test in test_gpu_parity.py uses).

Collected last on purpose: must pass alone in a fresh process and after the others.
"""
import ctypes as C

import pytest
import torch

from tests.helpers import rel_l2

pytestmark = pytest.mark.gpu


def _erwin(G, variant="flash"):
    import haet_b200 as hb
    torch.manual_seed(G)
    e = hb.ErwinTransformer(c_in=32, c_hidden=[32, 64], ball_sizes=[min(32, G), min(16, G // 2)], enc_num_heads=[4, 8],
                            enc_depths=[2, 2], dec_num_heads=[4], dec_depths=[2], strides=[2], rotate=45, mlp_ratio=2,
                            dimensionality=3, mp_steps=0, embed=False, variant=variant)
    g = torch.Generator().manual_seed(G + 1)
    with torch.no_grad():
        for n, p in e.named_parameters():
            if p.dim() == 2:
                p.copy_(torch.randn(p.shape, generator=g) / p.shape[1] ** 0.5)
            elif n.endswith("bias"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
    return e.cuda().eval(), g


def _case(G, n_trees, variant="flash", scale=1.0):
    """Inputs, tree and the existing entry point's output for n_trees trees of G slice tokens."""
    from haet_b200.balltree import balltree_device
    from haet_b200.erwin import erwin_device
    e, g = _erwin(G, variant)
    x = (scale * torch.randn(n_trees * G, 32, generator=g)).cuda()
    pos = torch.rand(n_trees * G, 3, generator=g).cuda()
    tree_idx, tree_mask, rots = balltree_device(pos, n_trees, G, e.strides, e.ball_sizes, float(e.rotate))
    want = erwin_device(e, x, pos, tree_idx, tree_mask, rots, n_trees)
    return dict(e=e, x=x, pos=pos, tree_idx=tree_idx, tree_mask=tree_mask, rots=rots, n_trees=n_trees, want=want,
                w=e.packed_weights(x.device), out=torch.full_like(x, float("nan")))


def _call_static(c):
    import haet_b200 as hb
    l = hb.lib()
    e = c["e"]
    nl = e.cfg.n_levels
    rot_ptrs = (C.c_void_p * nl)(*[(r.data_ptr() if r is not None else None) for r in c["rots"]])
    st = torch.cuda.current_stream().cuda_stream
    return l.haet_erwin_forward_static(C.byref(e.cfg), c["x"].data_ptr(), c["pos"].data_ptr(), c["x"].shape[0], c["n_trees"],
                                       c["tree_idx"].data_ptr(), c["tree_mask"].data_ptr(), c["tree_idx"].numel(), rot_ptrs,
                                       c["w"].data_ptr(), c["out"].data_ptr(), st)


def test_static_scratch_erwin_equals_workspace_entry_point_twice_in_a_row():
    import haet_b200 as hb
    l = hb.lib()
    # call 1: 8 trees x 256 tokens (the layer's own shape); call 2: FEWER, ragged trees (100 tokens -> padded slots) with the
    # distance-bias variant and large inputs, so every level's buffers sit at other offsets of the array than what call 1
    # left there; call 3: the first case again, on top of call 2's leftovers.
    a = _case(256, 8)
    b = _case(100, 3, variant="erwin", scale=50.0)
    assert hb._lib.HAET_OK == 0
    for c in (a, b):   # warm-up of both shapes through the new entry point: symbol lookup, shared-memory opt-ins, module load
        assert _call_static(c) == 0, l.haet_last_error()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    n0 = l.haet_launch_count()
    for c in (a, b, a):
        c["out"].fill_(float("nan"))   # in place: every call has to produce its result again
        assert _call_static(c) == 0, l.haet_last_error()
        torch.cuda.synchronize()
        assert torch.isfinite(c["out"]).all()
        assert rel_l2(c["out"], c["want"]) == 0.0 and torch.equal(c["out"], c["want"])
    free1, _ = torch.cuda.mem_get_info()
    assert free1 >= free0, f"device memory shrank by {free0 - free1} bytes across calls that must not allocate"
    assert l.haet_launch_count() > n0


def test_static_scratch_erwin_rejects_sizes_beyond_the_array():
    import haet_b200 as hb
    l = hb.lib()
    c = _case(256, 8)
    per_slot = l.haet_erwin_workspace_bytes(C.byref(c["e"].cfg), 1 << 20) / float(1 << 20)
    limit = 8 * 1024 * 1024
    too_many = int(limit / per_slot) + 4096            # slots whose scratch cannot fit
    nl = c["e"].cfg.n_levels
    rot_ptrs = (C.c_void_p * nl)(*[(r.data_ptr() if r is not None else None) for r in c["rots"]])
    n0 = l.haet_launch_count()
    # the size check comes before anything is read or launched: the (too small) buffers are never touched
    rc = l.haet_erwin_forward_static(C.byref(c["e"].cfg), c["x"].data_ptr(), c["pos"].data_ptr(), too_many, too_many // 256,
                                     c["tree_idx"].data_ptr(), c["tree_mask"].data_ptr(), too_many, rot_ptrs,
                                     c["w"].data_ptr(), c["out"].data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == hb._lib.HAET_ENOSUP
    assert b"HAET_ERWIN_STATIC_SCRATCH_BYTES" in l.haet_last_error()
    assert l.haet_launch_count() == n0
    torch.cuda.synchronize()
    # and the array still works afterwards
    assert _call_static(c) == 0, l.haet_last_error()
    torch.cuda.synchronize()
    assert torch.equal(c["out"], c["want"])
