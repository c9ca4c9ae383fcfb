"""The MLP half of a fused block, mlp(ln_2(fx)) + fx with mlp_ratio 2, at the bench shape (for ncu): python scripts/profile_mlp.py [points] [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import haet_b200 as hb
from haet_b200.models import MLP

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
torch.manual_seed(0)
mlp = MLP(256, 512, 256, n_layers=0, res=False).cuda().eval()
ln = torch.nn.LayerNorm(256).cuda()
fx = torch.randn(1, N, 256, device="cuda") * 1.7 + 0.4
for _ in range(iters):
    out = hb.block_mlp(fx, ln, mlp)
torch.cuda.synchronize()
print("ok", float(out.abs().mean()))
