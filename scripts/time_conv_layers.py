"""Times of the conv front-end layers (cfg 3 and a 512 x 512 2-D grid, 256 hidden): python scripts/time_conv_layers.py"""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, haet_b200 as hb
from tests.helpers import trained_like_
dev = torch.device("cuda")
for name, cls, kw, B, N in (("cfg3", hb.Physics_Attention_Structured_Mesh_3D, dict(dim=256, heads=8, dim_head=32, slice_num=64, H=64, W=64, D=64, rotate=45, mlp_ratio=2), 1, 64**3),
                            ("2d_512x512", hb.Physics_Attention_Structured_Mesh_2D, dict(dim=256, heads=8, dim_head=32, slice_num=64, H=512, W=512, rotate=45, mlp_ratio=2), 1, 512*512),
                            ("cfg2", hb.Physics_Attention_Structured_Mesh_2D, dict(dim=128, heads=8, dim_head=16, slice_num=64, H=85, W=85, rotate=45, mlp_ratio=1), 4, 85*85)):
    torch.manual_seed(0)
    m = cls(**kw).eval(); trained_like_(m, 1); m = m.to(dev)
    x = torch.nn.functional.layer_norm(torch.randn(B, N, kw["dim"], device=dev), (kw["dim"],))
    ms, mode = bench.timed_replay(lambda: m(x), 10, 3, dev)
    print(name, round(ms, 4), "ms", mode)
