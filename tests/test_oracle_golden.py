"""Pin the oracle (oracle/haet_oracle.py + oracle/balltree_oracle.c) against the golden vectors
generated from the unmodified reference (tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import haet_oracle as O
from tests.helpers import balltree_goldens, load_layer_golden, oracle_kwargs, rel_l2


@pytest.mark.parametrize("path", balltree_goldens())
def test_balltree_oracle_bit_exact(path):
    z = np.load(path)
    nt, G = int(z["n_trees"]), int(z["G"])
    pos = torch.from_numpy(z["pos"])
    bidx = torch.arange(nt).repeat_interleave(G)
    ti, tm, rots = O.balltree_build_with_rotations(pos, bidx, z["strides"].tolist(), z["ball_sizes"].tolist(),
                                                   float(z["angle"]))
    assert np.array_equal(ti.numpy(), z["tree_idx"])
    assert np.array_equal(tm.numpy(), z["tree_mask"])
    for i, r in enumerate(rots):
        if float(z["angle"]) > 0:
            assert np.array_equal(r.numpy(), z[f"rot{i}"]), f"level {i}"
        else:
            assert r is None


@pytest.mark.parametrize("name", ["layer_irregular_small", "layer_irregular_g64_fresh", "layer_2d_small",
                                  "layer_3d_small"])
def test_layer_oracle_matches_reference_golden(name):
    cfg, sd, x, out = load_layer_golden(name)
    got = O.physics_attention_forward(sd, x, **oracle_kwargs(cfg))
    # same fp32 arithmetic on the same machine type: agreement is at rounding level
    assert rel_l2(got, out) < 2e-6


def test_layer_oracle_fp64_close_to_fp32():
    cfg, sd, x, out = load_layer_golden("layer_irregular_small")
    taps = {}
    got32 = O.physics_attention_forward(sd, x, taps=taps, **oracle_kwargs(cfg))
    tree = (taps["tree_idx"], taps["tree_mask"], taps["rot_idx"])
    got64 = O.physics_attention_forward(sd, x, dtype=torch.float64, tree=tree, **oracle_kwargs(cfg))
    assert rel_l2(got64.float(), got32) < 1e-4
