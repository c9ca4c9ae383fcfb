"""One fused Attn(ln_1(fx)) + fx at the bench shape (for ncu): python scripts/profile_block.py [points] [slices] [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
G = int(sys.argv[2]) if len(sys.argv) > 2 else 256
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 2
m = bench.make_module(G, "cuda")
ln = torch.nn.LayerNorm(256).cuda()
fx = torch.randn(1, N, 256, device="cuda") * 1.7 + 0.4
for _ in range(iters):
    out = m.forward_block(fx, ln)
torch.cuda.synchronize()
print("ok", float(out.abs().mean()))
