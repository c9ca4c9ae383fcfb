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


def test_layer_oracle_train_mode_matches_reference_golden():
    """SURVEY 8f-4 (forward only): module.train() -- BatchNorm1d of BallPooling / BallUnpooling on batch statistics
    (pooling.py:23,56); 100 slices, so the duplicated leaf slots of the padded tree enter the statistics as in the reference."""
    cfg, sd, x, out = load_layer_golden("layer_irregular_train")
    got = O.physics_attention_forward(sd, x, bn_training=True, **oracle_kwargs(cfg))
    assert rel_l2(got, out) < 2e-6
    assert rel_l2(O.physics_attention_forward(sd, x, **oracle_kwargs(cfg)), out) > 1e-3   # eval statistics give something else


@pytest.mark.parametrize("name", ["layer_irregular_small", "layer_2d_small", "layer_3d_small"])
def test_streamed_oracle_equals_materialising_oracle(name):
    """The streamed restatement used at the BASELINE sizes (no N x G weights in memory; conv front-ends tap by tap) is the
    same function as the golden-pinned one: fp64 against fp64, chunked so that the running-state merge is exercised."""
    cfg, sd, x, out = load_layer_golden(name)
    kw = oracle_kwargs(cfg)
    taps = {}
    want = O.physics_attention_forward(sd, x, taps=taps, dtype=torch.float64, **kw)
    st = O.streamed_slice_state(sd, x, heads=kw["heads"], dim_head=kw["dim_head"], grid=kw.get("grid"), chunk=97)
    assert rel_l2(st["S"], taps["S"]) < 1e-12
    W = taps["W"]                                                    # [B, H, N, G]
    lse = torch.logsumexp(torch.log(W) + 0, dim=2)                   # log sum_n W = 0: the normaliser is consistent
    assert float(lse.abs().max()) < 1e-9
    rows = torch.arange(0, x.shape[1], 5)
    got = O.streamed_output_rows(sd, st, taps["S_out"], rows, heads=kw["heads"], dim_head=kw["dim_head"])
    assert rel_l2(got, want[:, rows]) < 1e-12
    # and its normaliser reproduces W itself on the sampled rows: exp(z - lse)
    own, _ = O.split_state_dict(sd, torch.float64)
    xc = st["x_proj"][:, rows].reshape(x.shape[0], -1, kw["heads"], kw["dim_head"])
    c = torch.log(-torch.log(torch.tensor(1e-6, dtype=torch.float32))).double()
    z = (torch.nn.functional.linear(xc, own["in_project_slice.weight"], own["in_project_slice.bias"]) - c) / \
        st["tau"][:, rows][..., None]
    assert rel_l2(torch.exp(z - st["lse"][:, None]).permute(0, 2, 1, 3), W[:, :, rows]) < 1e-12


def test_layer_oracle_fp64_close_to_fp32():
    cfg, sd, x, out = load_layer_golden("layer_irregular_small")
    taps = {}
    got32 = O.physics_attention_forward(sd, x, taps=taps, **oracle_kwargs(cfg))
    tree = (taps["tree_idx"], taps["tree_mask"], taps["rot_idx"])
    got64 = O.physics_attention_forward(sd, x, dtype=torch.float64, tree=tree, **oracle_kwargs(cfg))
    assert rel_l2(got64.float(), got32) < 1e-4


def _load_erwin_bias_golden():
    import json
    import os
    from tests.helpers import GOLDEN
    z = np.load(os.path.join(GOLDEN, "erwin_bias_small.npz"))
    cfg = json.loads(bytes(z["cfg"]).decode())
    sd = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd/")}
    return cfg, sd, torch.from_numpy(z["x"]), torch.from_numpy(z["pos"]), int(z["pts"]), torch.from_numpy(z["out"])


def test_erwin_distance_bias_oracle_matches_nonflash_reference_golden():
    """the oracle's variant="erwin" (sigma_att * cdist bias, (H E K) qkv columns) against the output of the
    reference's non-flash Erwin copy (golden made by tests/golden/make_golden.py::save_erwin_bias)."""
    cfg, sd, x, pos, pts, out = _load_erwin_bias_golden()
    bidx = torch.arange(x.shape[0] // pts).repeat_interleave(pts)
    got = O.erwin_forward(sd, dict(cfg, variant="erwin"), x, pos, bidx)
    assert rel_l2(got, out) < 2e-6
    # and the bias matters: the flash arithmetic on the same weights is far away
    assert rel_l2(O.erwin_forward(sd, dict(cfg, variant="flash"), x, pos, bidx), out) > 1e-2


MODELS = ["model_irregular_small", "model_irregular_256", "model_irregular_256_upos", "model_2d_small"]


@pytest.mark.parametrize("name", MODELS)
def test_model_oracle_matches_reference_golden(name):
    """the oracle's Model.forward restatement (preprocess MLP, get_grid, placeholder, time embedding, blocks,
    ln_3 + mlp2) against whole-model outputs of the unmodified reference."""
    from tests.helpers import load_model_golden, model_oracle_kwargs
    cfg, sd, x, fx, T, out = load_model_golden(name)
    got = O.model_forward(sd, x, fx, T, **model_oracle_kwargs(cfg))
    assert got.shape == out.shape
    assert rel_l2(got, out) < 5e-6


@pytest.mark.parametrize("name", MODELS)
def test_model_containers_load_reference_state_dict(name):
    """haet_b200.HAETransolver_* take the reference's constructor arguments and its state_dict, strictly."""
    import haet_b200 as hb
    from tests.helpers import load_model_golden
    cfg, sd, *_ = load_model_golden(name)
    cls = {"irregular": hb.HAETransolver_Irregular_Mesh, "2d": hb.HAETransolver_Structured_Mesh_2D,
           "3d": hb.HAETransolver_Structured_Mesh_3D}[cfg["variant"]]
    m = cls(**cfg["kwargs"])
    assert set(m.state_dict()) == set(sd)
    m.load_state_dict(sd, strict=True)
    with pytest.raises(hb.HaetError):
        m.eval()(torch.zeros(1, 10, cfg["kwargs"]["space_dim"]), None)    # CPU tensors: no fallback
    if cfg["variant"] != "irregular" and cfg["kwargs"].get("unified_pos"):
        kw = cfg["kwargs"]
        sizes = (kw["H"], kw["W"]) + ((kw["D"],) if cfg["variant"] == "3d" else ())
        want = O.structured_grid_distances(sizes, kw["ref"])             # pinned by the golden above
        assert torch.equal(m.pos.reshape(want.shape), want)
