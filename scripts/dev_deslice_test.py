import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import haet_b200 as hb
from haet_b200 import _lib
from tests.helpers import trained_like_, rel_l2
l = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
def run(N, G, B=1, time_it=False):
    torch.manual_seed(N + G)
    m = hb.Physics_Attention_Irregular_Mesh(256, heads=8, dim_head=32, slice_num=G, rotate=45, mlp_ratio=2).eval()
    trained_like_(m, 3)
    m = m.cuda()
    x = torch.nn.functional.layer_norm(torch.randn(B, N, 256, device="cuda"), (256,))
    sh = m._shape(B, N); psh = C.byref(sh)
    wx, bx = m._wx().float(), m.in_project_x.bias.detach()
    ws, bs = m.in_project_slice.weight.detach().contiguous(), m.in_project_slice.bias.detach()
    wt, bt = m.ada_temp_linear.weight.detach().view(-1).contiguous(), m._bt()
    wo, bo = m.to_out[0].weight.detach().contiguous(), m.to_out[0].bias.detach()
    spill = torch.zeros(l.haet_spill_bytes(psh), dtype=torch.uint8, device="cuda")
    work = torch.zeros(max(l.haet_slice_workspace_bytes(psh), l.haet_deslice_workspace_bytes(psh)), dtype=torch.uint8, device="cuda")
    part = torch.zeros(l.haet_partial_floats(psh), device="cuda")
    T = B * 8 * G
    eid = torch.empty(T, 32, device="cuda"); lse = torch.empty(T, device="cuda"); pos = torch.empty(T, 3, device="cuda")
    outs = {}
    for path in (1, 0):
        _lib.check(l.haet_set_path(path))
        _lib.check(l.haet_slice_reduce(psh, x.data_ptr(), wx.data_ptr(), bx.data_ptr(), ws.data_ptr(), bs.data_ptr(), wt.data_ptr(), bt,
                                       spill.data_ptr(), part.data_ptr(), work.data_ptr(), st))
        _lib.check(l.haet_slice_merge(psh, part.data_ptr(), 1, eid.data_ptr(), lse.data_ptr(), pos.data_ptr(), st))
        states = torch.randn(T, 32, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5))
        out = torch.zeros(B, N, 256, device="cuda")
        _lib.check(l.haet_deslice_out(psh, spill.data_ptr(), lse.data_ptr(), states.data_ptr(), ws.data_ptr(), bs.data_ptr(), wt.data_ptr(), bt,
                                      wo.data_ptr(), bo.data_ptr(), out.data_ptr(), work.data_ptr(), st))
        torch.cuda.synchronize()
        outs[path] = out.clone()
        if time_it:
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record()
            _lib.check(l.haet_slice_reduce(psh, x.data_ptr(), wx.data_ptr(), bx.data_ptr(), ws.data_ptr(), bs.data_ptr(), wt.data_ptr(), bt,
                                           spill.data_ptr(), part.data_ptr(), work.data_ptr(), st))
            e[1].record()
            _lib.check(l.haet_deslice_out(psh, spill.data_ptr(), lse.data_ptr(), states.data_ptr(), ws.data_ptr(), bs.data_ptr(), wt.data_ptr(), bt,
                                          wo.data_ptr(), bo.data_ptr(), out.data_ptr(), work.data_ptr(), st))
            e[2].record(); torch.cuda.synchronize()
            print(f"   path {path}: slice_reduce {e[0].elapsed_time(e[1]):.3f} ms, deslice_out {e[1].elapsed_time(e[2]):.3f} ms")
    print(f"N={N} G={G} B={B}: out rel_l2 (tc vs simt, same states)={rel_l2(outs[0], outs[1]):.3e} finite={bool(torch.isfinite(outs[0]).all())}", flush=True)
for a in [(128, 32), (4099, 32), (5000, 128), (100000, 256), (3000, 512), (20000, 100), (7777, 256, 2)]:
    run(*a)
run(1_000_000, 256, time_it=True)
