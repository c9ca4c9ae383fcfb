"""Per-layer forward time of the BASELINE configs' layer shapes on one GPU (eager launches, CUDA events, mean of 10).

    python scripts/bench_configs.py        -> one JSON line per config (written to profiles/ by the caller)
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import haet_b200 as hb

CFGS = [
    ("cfg1 irregular N=32186 G=32", "Physics_Attention_Irregular_Mesh", 1, 32186, dict(dim=256, heads=8, dim_head=32, slice_num=32, mlp_ratio=2)),
    ("cfg2 2D Darcy 85x85 B=4 D=128 G=64", "Physics_Attention_Structured_Mesh_2D", 4, 85 * 85, dict(dim=128, heads=8, dim_head=16, slice_num=64, H=85, W=85, mlp_ratio=2)),
    ("cfg3 3D 64^3 D=256 G=64", "Physics_Attention_Structured_Mesh_3D", 1, 64 ** 3, dict(dim=256, heads=8, dim_head=32, slice_num=64, H=64, W=64, D=64, mlp_ratio=2)),
    ("cfg4 irregular N=1M G=256", "Physics_Attention_Irregular_Mesh", 1, 1_000_000, dict(dim=256, heads=8, dim_head=32, slice_num=256, mlp_ratio=2)),
    ("cfg5 irregular 1M points per GPU G=512", "Physics_Attention_Irregular_Mesh", 1, 1_000_000, dict(dim=256, heads=8, dim_head=32, slice_num=512, mlp_ratio=2)),
]
for name, cls, B, N, kw in CFGS:
    torch.manual_seed(0)
    m = getattr(hb, cls)(**kw).cuda().eval()
    x = torch.nn.functional.layer_norm(torch.randn(B, N, kw["dim"], device="cuda"), (kw["dim"],))
    for _ in range(3):
        m(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        m(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(json.dumps({"config": name, "B": B, "N": N, "ms_per_layer": ms, "points_per_s": B * N / (ms * 1e-3)}), flush=True)
