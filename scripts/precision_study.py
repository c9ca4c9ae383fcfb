"""Which operand precision does each contraction of the layer tolerate?  CPU study with the oracle's
arithmetic: round selected operands to bf16 / tf32 and report rel-L2 of the layer output vs fp32.
(Design input for the tensor-core kernels; numbers quoted in DESIGN.md.)"""
import math, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from oracle import haet_oracle as O
from tests.helpers import trained_like_, rel_l2
import haet_b200 as hb


def rnd(t, mode):
    if mode == "f32":
        return t
    if mode == "bf16":
        return t.bfloat16().float()
    if mode == "tf32":      # round-to-nearest to 10 mantissa bits
        i = t.contiguous().view(torch.int32)
        i = (i + 0x1000) & ~0x1FFF
        return i.view(torch.float32)
    if mode == "tf32t":     # truncation (what the tensor core does with raw fp32 words)
        i = t.contiguous().view(torch.int32) & ~0x1FFF
        return i.view(torch.float32)
    if mode == "bf16x2":    # hi + lo split: ~16 mantissa bits
        hi = t.bfloat16().float()
        return hi + (t - hi).bfloat16().float()
    raise ValueError(mode)


def layer(sd, x, kw, p_in, p_logit, p_w, p_state, tree=None, pos=None, taps=None):
    own, erw = O.split_state_dict(sd)
    B, N, _ = x.shape
    H, C, G = kw["heads"], kw["dim_head"], kw["slice_num"]
    xp = F.linear(rnd(x, p_in), rnd(own["in_project_x.weight"], p_in), own["in_project_x.bias"])
    xp = xp.reshape(B, N, H, C).permute(0, 2, 1, 3).contiguous()
    tau = 0.5 + F.linear(xp, own["ada_temp_linear.weight"], own["ada_temp_linear.bias"]).clamp(-0.4, 0.4)
    c = torch.log(-torch.log(torch.tensor(1e-6)))
    logits = F.linear(rnd(xp, p_logit), rnd(own["in_project_slice.weight"], p_logit)) + own["in_project_slice.bias"] - c
    z = logits / tau
    m = z.amax(2, keepdim=True)
    e = torch.exp(z - m)
    l = e.sum(2, keepdim=True)
    S = torch.einsum("bhnc,bhng->bhgc", rnd(xp, p_w), rnd(e, p_w)) / l.transpose(2, 3) / (1 + 1e-5)
    S_flat = S.reshape(B * H * G, C)
    if pos is None:
        pos = O.slice_positions(S_flat, 3)
    cfg = O.default_erwin_cfg(H, C, G, 3, rotate=kw["rotate"], mlp_ratio=kw["mlp_ratio"])
    bidx = torch.arange(B * H).repeat_interleave(G)
    t2 = {}
    So = O.erwin_forward(erw, cfg, S_flat, pos, bidx, tree=tree, taps=t2)
    W = e / l
    out_x = torch.einsum("bhgc,bhng->bhnc", rnd(So.reshape(B, H, G, C), p_state), rnd(W, p_w))
    out_x = out_x.permute(0, 2, 1, 3).reshape(B, N, H * C)
    out = F.linear(rnd(out_x, p_in), rnd(own["to_out.0.weight"], p_in), own["to_out.0.bias"])
    if taps is not None:
        taps.update(tree=(t2["tree_idx"], t2["tree_mask"], t2["rot_idx"]), pos=pos, S=S)
    return out


if __name__ == "__main__":
    torch.manual_seed(0)
    for (N, G, scale) in ((8192, 256, 1.0), (8192, 32, 1.0), (8192, 256, 2.0)):
        kw = dict(heads=8, dim_head=32, slice_num=G, rotate=45, mlp_ratio=2)
        m = hb.Physics_Attention_Irregular_Mesh(256, **kw).eval()
        trained_like_(m, 1)
        with torch.no_grad():
            m.in_project_slice.weight.mul_(scale)
        sd = m.state_dict()
        x = F.layer_norm(torch.randn(1, N, 256), (256,))
        taps = {}
        ref = layer(sd, x, kw, "f32", "f32", "f32", "f32", taps=taps)
        print(f"N={N} G={G} ws_scale={scale}")
        for cfgp in (("bf16", "f32", "f32", "f32"), ("tf32", "f32", "f32", "f32"), ("tf32t", "f32", "f32", "f32"),
                     ("f32", "bf16", "f32", "f32"), ("f32", "tf32", "f32", "f32"), ("f32", "tf32t", "f32", "f32"),
                     ("f32", "bf16x2", "f32", "f32"),
                     ("f32", "f32", "bf16", "f32"), ("f32", "f32", "tf32", "f32"), ("f32", "f32", "tf32t", "f32"),
                     ("f32", "f32", "f32", "bf16"), ("f32", "f32", "f32", "tf32"),
                     ("bf16", "bf16", "bf16", "bf16"), ("tf32", "tf32", "tf32", "tf32"), ("bf16", "tf32", "tf32", "tf32"),
                     ("bf16", "bf16x2", "bf16", "bf16"), ("bf16", "tf32", "bf16", "bf16")):
            t = {}
            got = layer(sd, x, kw, *cfgp, tree=taps["tree"], pos=taps["pos"], taps=t)
            print(f"  in={cfgp[0]:6s} logit={cfgp[1]:6s} w={cfgp[2]:6s} state={cfgp[3]:6s}  out rel_l2={rel_l2(got, ref):.2e}"
                  f"  S rel_l2={rel_l2(t['S'], taps['S']):.2e}")
