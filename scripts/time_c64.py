"""Layer time at dim_head 64 (256 hidden, 4 heads, 1M points, 256 slices): tcgen05 pair mode against the fp32 SIMT kernels."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import haet_b200 as hb

torch.manual_seed(0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
m = hb.Physics_Attention_Irregular_Mesh(256, heads=4, dim_head=64, slice_num=256, rotate=45, mlp_ratio=2).eval().cuda()
x = torch.nn.functional.layer_norm(torch.randn(1, N, 256, device="cuda"), (256,))
for path in (0, 1):
    hb.lib().haet_set_path(path)
    for _ in range(3):
        y = m(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        y = m(x)
    e1.record()
    torch.cuda.synchronize()
    print(f"dim_head 64, N={N}, path {path} ({'tcgen05 pair mode' if path == 0 else 'fp32 SIMT'}): {e0.elapsed_time(e1) / 5:.3f} ms / layer", flush=True)
hb.lib().haet_set_path(0)
