"""One cfg-3 layer forward (64^3 grid, 256 hidden, 64 slices) for ncu: python scripts/profile_conv.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import haet_b200 as hb
from tests.helpers import trained_like_
torch.manual_seed(0)
m = hb.Physics_Attention_Structured_Mesh_3D(dim=256, heads=8, dim_head=32, slice_num=64, H=64, W=64, D=64, rotate=45, mlp_ratio=2).eval()
trained_like_(m, 1)
m = m.cuda()
x = torch.nn.functional.layer_norm(torch.randn(1, 64 ** 3, 256, device="cuda"), (256,))
for _ in range(2):
    out = m(x)
torch.cuda.synchronize()
print("ok", float(out.abs().mean()))
