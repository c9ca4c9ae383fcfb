import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import haet_b200 as hb
from haet_b200 import _lib
from tests.helpers import trained_like_, rel_l2
l = _lib.lib()
def run(N, G, B=1, scale=1.0, time_it=False):
    torch.manual_seed(N + G)
    m = hb.Physics_Attention_Irregular_Mesh(256, heads=8, dim_head=32, slice_num=G, rotate=45, mlp_ratio=2).eval()
    trained_like_(m, 3)
    with torch.no_grad(): m.in_project_slice.weight.mul_(scale)
    m = m.cuda(); m.keep_taps = True
    x = torch.nn.functional.layer_norm(torch.randn(B, N, 256, device="cuda"), (256,))
    res = {}
    for path in (1, 0):
        _lib.check(l.haet_set_path(path))
        out = m(x); torch.cuda.synchronize()
        res[path] = (m.last_taps["S"].clone(), m.last_taps["lse2"].clone(), out.clone())
        if time_it:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3): m(x)
            e1.record(); torch.cuda.synchronize()
            print(f"   path {path}: {e0.elapsed_time(e1)/3:.3f} ms per layer")
    S1, L1, O1 = res[1]; S0, L0, O0 = res[0]
    print(f"N={N} G={G} B={B} scale={scale}: S rel_l2={rel_l2(S0, S1):.3e} lse maxabs={float((L0-L1).abs().max()):.3e} "
          f"out rel_l2={rel_l2(O0, O1):.3e} finite={bool(torch.isfinite(O0).all())}", flush=True)
for a in [(64, 32), (4099, 32), (5000, 128), (100000, 256), (3000, 512), (20000, 100), (7777, 256, 2), (50000, 256, 1, 3.0)]:
    run(*a)
run(1_000_000, 256, time_it=True)
