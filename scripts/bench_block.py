"""Attention half of a TransolverErwinBlock, fx = Attn(ln_1(fx)) + fx (HAETransolver_Irregular_Mesh.py:125), at BASELINE
config 4 (1M points, 256 hidden, 256 slices) on one GPU: torch LayerNorm + layer + torch add against the fused
forward_block (haet_slice_reduce_ln / haet_deslice_out_res).  CUDA events, mean of 10 after 3 warm-ups, eager launches.

    python scripts/bench_block.py [points] [slices]   -> one JSON line
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import haet_b200 as hb
from haet_b200 import _lib

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
G = int(sys.argv[2]) if len(sys.argv) > 2 else 256
torch.manual_seed(0)
m = hb.Physics_Attention_Irregular_Mesh(256, heads=8, dim_head=32, slice_num=G, mlp_ratio=2).cuda().eval()
ln = torch.nn.LayerNorm(256).cuda()
fx = torch.randn(1, N, 256, device="cuda") * 1.7 + 0.4


def timed(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


with torch.no_grad():
    t_ln = timed(lambda: ln(fx))
    xn = ln(fx)
    t_layer = timed(lambda: m(xn))
    y = m(xn)
    t_add = timed(lambda: y + fx)
    t_unfused = timed(lambda: m(ln(fx)) + fx)
    t_fused = timed(lambda: m.forward_block(fx, ln))
    a, b = m(ln(fx)), m.forward_block(fx, ln)
    err = float(((b - fx) - a).double().norm() / a.double().norm())
    # the two kernels the fusion touches, timed alone through the stage mask
    l = _lib.lib()
    stages = {}
    for name, mask in (("in_project", 1), ("to_out", 8)):   # each figure includes the same merge / tree / Erwin launches
        l.haet_set_stage_mask(mask)
        stages[name + "_plain"] = timed(lambda: m(xn))
        stages[name + "_fused"] = timed(lambda: m.forward_block(fx, ln))
    l.haet_set_stage_mask(15)

    # the MLP half: fx = mlp(ln_2(fx)) + fx with the reference's MLP layout (mlp.py:67-74), hidden = mlp_ratio * 256
    class RefMLP(torch.nn.Module):
        def __init__(self, dim, hidden):
            super().__init__()
            self.linear_pre = torch.nn.Sequential(torch.nn.Linear(dim, hidden), torch.nn.GELU())
            self.linear_post = torch.nn.Linear(hidden, dim)
            self.linears = torch.nn.ModuleList()

        def forward(self, x):
            return self.linear_post(self.linear_pre(x))

    mlp_res = {}
    for ratio in (1, 2):
        mlp = RefMLP(256, 256 * ratio).cuda()
        ln2 = torch.nn.LayerNorm(256).cuda()
        t_torch = timed(lambda: mlp(ln2(fx)) + fx, n=5, warm=2)
        t_ours = timed(lambda: hb.block_mlp(fx, ln2, mlp))
        ra, rb = mlp(ln2(fx)), hb.block_mlp(fx, ln2, mlp) - fx
        mlp_res[f"mlp_ratio_{ratio}"] = {"ms_torch_fp32": t_torch, "ms_fused": t_ours,
                                         "rel_l2_mlp_term_vs_torch_fp32": float((rb - ra).double().norm() / ra.double().norm())}
        t_block_torch = timed(lambda: (lambda y: mlp(ln2(y)) + y)(m(ln(fx)) + fx), n=5, warm=2)
        t_block = timed(lambda: hb.block_mlp(m.forward_block(fx, ln), ln2, mlp))
        mlp_res[f"mlp_ratio_{ratio}"].update({"ms_whole_block_unfused": t_block_torch, "ms_whole_block_fused": t_block,
                                              "block_points_per_s_fused": N / (t_block * 1e-3)})
    # an 8-block stack (BASELINE config 4 "full model": n_layers 8; without the model's preprocess MLP and output head,
    # HAETransolver_Irregular_Mesh.py:297-336), every block with its own weights, fused path, mlp_ratio 2
    blocks = []
    for i in range(8):
        torch.manual_seed(100 + i)
        blocks.append((hb.Physics_Attention_Irregular_Mesh(256, heads=8, dim_head=32, slice_num=G, mlp_ratio=2).cuda().eval(),
                       torch.nn.LayerNorm(256).cuda(), torch.nn.LayerNorm(256).cuda(), RefMLP(256, 512).cuda()))

    def stack():
        y = fx
        for attn, l1, l2, mm in blocks:
            y = hb.block_mlp(attn.forward_block(y, l1), l2, mm)
        return y

    t_stack = timed(stack, n=5, warm=2)
    stack_res = {"blocks": 8, "ms": t_stack, "points_per_s": N / (t_stack * 1e-3), "finite": bool(torch.isfinite(stack()).all())}

print(json.dumps({"config": f"block attention half, N={N}, G={G}", "ms_layernorm_torch": t_ln, "ms_layer": t_layer,
                  "ms_residual_add_torch": t_add, "ms_unfused_total": t_unfused, "ms_fused_total": t_fused,
                  "points_per_s_unfused": N / (t_unfused * 1e-3), "points_per_s_fused": N / (t_fused * 1e-3),
                  "rel_l2_attention_term_fused_vs_unfused": err, "stages_ms_incl_small_kernels": stages, "mlp_half": mlp_res, "stack_of_8_blocks_fused": stack_res,
                  "torch_matmul_tf32": torch.backends.cuda.matmul.allow_tf32}))
