"""Parity AT THE STATED CONFIG SIZES (BASELINE.json configs 3, 4 and 5's per-GPU shard) against the streamed fp64 oracle
(oracle/haet_oracle.py: streamed_slice_state / streamed_output_rows, which never materialise the N x G weights and are
pinned to 1e-15 against the golden-pinned materialising oracle in tests/test_oracle_golden.py), plus the adversarial
range cases the small meshes cannot see:

  * a slice column whose first tile holds an outlier 25.5 / 31 log2-units above a bulk of ~55 k points in the same chunk
    (pass 1: fp16 P under the lazy running reference against the fp32 l),
  * Erwin outputs of magnitude 1e5 and 1e-6 (pass 2: S' is a single fp16 term behind a per-(b, h) power-of-two scale),
  * projected points beyond the fp16 range (the operand image saturates instead of producing NaN).

Tolerance: rel-L2 <= 1e-3 (north_star); the observed values are printed and quoted in BASELINE.md.

lse2 (the global softmax normaliser, an INTERNAL quantity) is held to a mean-abs 3e-3 / max-abs 5e-2 bound in log2 units:
with trained-like weights at 1M points the largest logits reach |z2| ~ 150-500 and z = logits / tau amplifies the 2^-17
relative error of the 3-term bf16 in-projection by |z| / tau.  Measured on the oracle itself (x_proj
perturbed by 2^-17, N = 262,144): lse2 max-abs 1.0e-2, mean-abs 1.7e-4, S rel-L2 9e-5; the reference's own fp32
arithmetic moves lse2 by 9e-4.  The weights W = 2^(z2 - lse2) of the same points move together, which is why the
outputs agree to 1e-4.
"""
import ctypes as C

import pytest
import torch

from oracle import haet_oracle as O
from tests.helpers import rel_l2, trained_like_

pytestmark = pytest.mark.gpu
TOL = 1e-3


def _taps_tree(t):
    return (t["tree_idx"].cpu(), t["tree_mask"].cpu(), [r.cpu() if r is not None else None for r in t["rot_idx"]])


def _check_against_streamed(m, sd, x, okw, grid=None, n_rows=4096, tag=""):
    """Run the module on the GPU; compare S / lse2 / pos with the streamed oracle, then the oracle's Erwin (on the
    oracle's S, with the device's positions and tree: SURVEY 7.3 hard part 2) and `n_rows` sampled output rows."""
    m = m.cuda()
    m.keep_taps = True
    got = m(x.cuda())
    t = m.last_taps
    H, Cc, G = okw["heads"], okw["dim_head"], okw["slice_num"]
    st = O.streamed_slice_state(sd, x, heads=H, dim_head=Cc, grid=grid)
    e_S = rel_l2(t["S"], st["S"])
    lse2 = st["lse"] / torch.log(torch.tensor(2.0, dtype=torch.float64))
    d_lse = (t["lse2"].cpu().double() - lse2).abs()
    e_lse, e_lse_mean = float(d_lse.max()), float(d_lse.mean())
    pos = O.slice_positions(st["S"].float().reshape(-1, Cc), okw["dimensionality"])
    e_pos = rel_l2(t["pos"], pos)
    _, erw = O.split_state_dict(sd)
    ecfg = O.default_erwin_cfg(H, Cc, G, okw["dimensionality"], rotate=okw["rotate"], mlp_ratio=okw["mlp_ratio"])
    T = st["S"].shape[0] * H * G
    bidx = torch.arange(T // G).repeat_interleave(G)
    S_out = O.erwin_forward(erw, ecfg, st["S"].float().reshape(T, Cc), t["pos"].cpu(), bidx, tree=_taps_tree(t))
    e_erwin = rel_l2(t["S_out"], S_out)
    g = torch.Generator().manual_seed(1234)
    rows = torch.randperm(x.shape[1], generator=g)[:n_rows].sort().values
    want = O.streamed_output_rows(sd, st, S_out, rows, heads=H, dim_head=Cc)
    e_out = rel_l2(got[:, rows.cuda()], want)
    print(f"\n[fullsize {tag}] N={x.shape[1]} G={G}: S {e_S:.2e}  lse2 max-abs {e_lse:.2e} mean-abs {e_lse_mean:.2e}  pos {e_pos:.2e}  "
          f"erwin {e_erwin:.2e}  out({len(rows)} rows) {e_out:.2e}")
    assert torch.isfinite(got).all()
    assert e_S < TOL and e_lse < 5e-2 and e_lse_mean < 3e-3 and e_pos < TOL and e_erwin < TOL and e_out < TOL
    return dict(S=e_S, lse2=e_lse, pos=e_pos, erwin=e_erwin, out=e_out)


@pytest.mark.parametrize("tag,N,G", [("cfg4", 1_000_000, 256), ("cfg5-shard", 1_000_000, 512)])
def test_irregular_full_size_matches_streamed_oracle(tag, N, G):
    """BASELINE config 4 (1M points, 256 slices) and the per-GPU shard of config 5 (1M of the 8M points, 512 slices)."""
    import haet_b200 as hb
    torch.manual_seed(G)
    kw = dict(dim=256, heads=8, dim_head=32, slice_num=G, rotate=45, mlp_ratio=2)
    m = hb.Physics_Attention_Irregular_Mesh(**kw).eval()
    trained_like_(m, 3)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    x = torch.nn.functional.layer_norm(torch.randn(1, N, 256), (256,))
    _check_against_streamed(m, sd, x, dict(heads=8, dim_head=32, slice_num=G, rotate=45, mlp_ratio=2, dimensionality=3),
                            tag=tag)


def test_structured_3d_full_size_matches_streamed_oracle():
    """BASELINE config 3: 64^3 grid, 256 hidden, 64 slices -- the 128 x 256-tile tcgen05 implicit-GEMM Conv3d route."""
    import haet_b200 as hb
    torch.manual_seed(64)
    kw = dict(dim=256, heads=8, dim_head=32, slice_num=64, H=64, W=64, D=64, rotate=45, mlp_ratio=2)
    m = hb.Physics_Attention_Structured_Mesh_3D(**kw).eval()
    trained_like_(m, 5)
    with torch.no_grad():   # default conv init x 27 taps gives |x_proj| ~ 0.6: keep the softmax peaky but finite
        m.in_project_x.weight.mul_(0.5)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    N = 64 ** 3
    x = torch.nn.functional.layer_norm(torch.randn(1, N, 256), (256,))
    _check_against_streamed(m, sd, x, dict(heads=8, dim_head=32, slice_num=64, rotate=45, mlp_ratio=2, dimensionality=3),
                            grid=(64, 64, 64), tag="cfg3")


def _identity_projection_(m):
    """x_proj = x, tau = 0.5 for every point: the logits are then a known function of the input."""
    with torch.no_grad():
        m.in_project_x.weight.copy_(torch.eye(256))
        m.in_project_x.bias.zero_()
        m.ada_temp_linear.weight.zero_()
        m.ada_temp_linear.bias.zero_()


def test_pass1_outlier_column_at_full_chunk_size():
    """1M points = 18 chunks of ~55 k per head.  Columns g = 0 / 1 see, in the FIRST tile of chunk 0, a point whose logit is
    25.5 / 31 log2-units above the bulk; the bulk has a non-zero mean, so mass that the fp16 P of the accumulation loses
    while the fp32 l keeps it biases S / l ("What's weak" 1a).  A numpy emulation of the two schemes on this
    very column: reference = first-tile max (round 1) 9.0e-4, reference = max - 7 log2 units (now) 1.9e-4, the latter
    being the fp16 rounding floor of the hi operand."""
    import haet_b200 as hb
    N, G = 1_000_000, 256
    torch.manual_seed(7)
    m = hb.Physics_Attention_Irregular_Mesh(256, heads=8, dim_head=32, slice_num=G, rotate=45, mlp_ratio=2).eval()
    trained_like_(m, 3)
    _identity_projection_(m)
    log2e = 1.4426950408889634
    with torch.no_grad():
        w = m.in_project_slice.weight
        w.mul_(0.2)
        w[0].zero_(); w[0, 0] = 1.0
        w[1].zero_(); w[1, 1] = 1.0
        m.in_project_slice.bias.zero_()
    x = 3.0 + 0.1 * torch.randn(1, N, 256)
    # z2 = (x_proj[n, c] - shift) * (log2e / 0.5): an offset d in x is 2 log2e d log2-units
    for h in range(8):
        x[0, 3, h * 32 + 0] = 3.0 + 25.5 / (2 * log2e)
        x[0, 5, h * 32 + 1] = 3.0 + 31.0 / (2 * log2e)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    m = m.cuda()
    m.keep_taps = True
    m(x.cuda())
    t = m.last_taps
    st = O.streamed_slice_state(sd, x, heads=8, dim_head=32)
    lse2 = st["lse"] / torch.log(torch.tensor(2.0, dtype=torch.float64))
    for g in (0, 1):
        e = rel_l2(t["S"][:, :, g], st["S"][:, :, g])
        # the bulk's share of the column mass (head 0), so that the check is known to bite: 1 - W[outlier point, g]
        n0 = 3 if g == 0 else 5
        z_out = (x[0, n0, g].double() - float(torch.log(-torch.log(torch.tensor(1e-6))))) / 0.5
        share = 1.0 - float(torch.exp(z_out - st["lse"][0, 0, g]))
        print(f"\n[outlier column {g}] S rel-L2 {e:.2e}, bulk share of the column mass {share:.2e}")
        assert e < 5e-4, (g, e)
    assert rel_l2(t["S"], st["S"]) < 5e-4
    assert float((t["lse2"].cpu().double() - lse2).abs().max()) < 1e-3


def _passes_by_hand(m, x):
    """slice_reduce + slice_merge through the C-ABI; returns what haet_deslice_out needs to be driven with chosen states."""
    from haet_b200 import _lib
    l = _lib.lib()
    B, N, _ = x.shape
    sh = m._shape(B, N)
    st = torch.cuda.current_stream().cuda_stream
    spill = torch.empty(l.haet_spill_bytes(C.byref(sh)), dtype=torch.uint8, device="cuda")
    work = torch.empty(max(l.haet_slice_workspace_bytes(C.byref(sh)), l.haet_deslice_workspace_bytes(C.byref(sh))),
                       dtype=torch.uint8, device="cuda")
    part = torch.empty(l.haet_partial_floats(C.byref(sh)), device="cuda")
    keep = dict(wx=m._wx().float(), bx=m.in_project_x.bias.detach(), ws=m.in_project_slice.weight.detach().contiguous(),
                bs=m.in_project_slice.bias.detach(), wt=m.ada_temp_linear.weight.detach().view(-1).contiguous(),
                wo=m.to_out[0].weight.detach().contiguous(), bo=m.to_out[0].bias.detach())
    bt = m._bt()
    _lib.check(l.haet_slice_reduce(C.byref(sh), x.data_ptr(), keep["wx"].data_ptr(), keep["bx"].data_ptr(),
                                   keep["ws"].data_ptr(), keep["bs"].data_ptr(), keep["wt"].data_ptr(), bt,
                                   spill.data_ptr(), part.data_ptr(), work.data_ptr(), st))
    T = B * m.heads * m.slice_num
    eid = torch.empty(T, m.dim_head, device="cuda")
    lse = torch.empty(T, device="cuda")
    pos = torch.empty(T, m.dimensionality, device="cuda")
    _lib.check(l.haet_slice_merge(C.byref(sh), part.data_ptr(), 1, eid.data_ptr(), lse.data_ptr(), pos.data_ptr(), st))

    def deslice(states):
        o = torch.empty(B, N, m.dim, device="cuda")
        _lib.check(l.haet_deslice_out(C.byref(sh), spill.data_ptr(), lse.data_ptr(), states.data_ptr(),
                                      keep["ws"].data_ptr(), keep["bs"].data_ptr(), keep["wt"].data_ptr(), bt,
                                      keep["wo"].data_ptr(), keep["bo"].data_ptr(), o.data_ptr(), work.data_ptr(), st))
        return o
    return eid, lse, deslice


@pytest.mark.parametrize("mag", [1e5, 3e7, 1e-6, 1.0])
def test_pass2_state_range(mag):
    """Erwin outputs far outside O(1): S' goes to the tensor core as ONE fp16 term, so without the per-(b, h) power-of-two
    scale 1e5 overflows (65504) and 1e-6 sits in the subnormals ("What's weak" 1b).  One head carries a
    different magnitude than the others and one row is 2^-12 of its head's largest."""
    import haet_b200 as hb
    N, G = 4099, 256
    torch.manual_seed(11)
    m = hb.Physics_Attention_Irregular_Mesh(256, heads=8, dim_head=32, slice_num=G, rotate=45, mlp_ratio=2).eval()
    trained_like_(m, 4)
    with torch.no_grad():
        m.to_out[0].bias.zero_()   # an O(0.1) bias would hide a 1e-6 signal behind the fp32 rounding of the output itself
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    x = torch.nn.functional.layer_norm(torch.randn(1, N, 256), (256,))
    m = m.cuda()
    eid, lse, deslice = _passes_by_hand(m, x.cuda())
    g = torch.Generator().manual_seed(5)
    states = torch.randn(8, G, 32, generator=g) * mag
    states[3] *= 1e-3
    states[5, 17] *= 2.0 ** -12
    got = deslice(states.reshape(8 * G, 32).cuda().contiguous())
    st = O.streamed_slice_state(sd, x, heads=8, dim_head=32)
    want = O.streamed_output_rows(sd, st, states.reshape(8 * G, 32), torch.arange(N), heads=8, dim_head=32)
    e = rel_l2(got, want)
    print(f"\n[state range {mag:g}] out rel-L2 {e:.2e}")
    assert torch.isfinite(got).all()
    assert e < 5e-4


def test_projection_beyond_fp16_range_saturates():
    """|x_proj| > 65504 cannot be stored in the fp16 hi/lo operand image: it saturates (documented in haet_b200.h) and the
    layer output stays finite -- it never turns into (inf, -inf) -> NaN logits."""
    import haet_b200 as hb
    torch.manual_seed(2)
    m = hb.Physics_Attention_Irregular_Mesh(256, heads=8, dim_head=32, slice_num=64, rotate=45, mlp_ratio=2).eval()
    _identity_projection_(m)
    with torch.no_grad():
        m.in_project_slice.weight.mul_(1e-5)
    x = torch.randn(1, 2000, 256)
    x[0, 17, 40] = 3.0e5
    x[0, 900, 7] = -1.0e6
    out = m.cuda()(x.cuda())
    assert torch.isfinite(out).all()
