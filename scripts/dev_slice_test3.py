import sys, os, ctypes as C, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import haet_b200 as hb
from haet_b200 import _lib
from tests.helpers import trained_like_, rel_l2
l = _lib.lib()
N, G = 5000, 128
torch.manual_seed(N + G)
m = hb.Physics_Attention_Irregular_Mesh(256, heads=8, dim_head=32, slice_num=G, rotate=45, mlp_ratio=2).eval()
trained_like_(m, 3)
m = m.cuda()
x = torch.nn.functional.layer_norm(torch.randn(1, N, 256, device="cuda"), (256,))
sh = m._shape(1, N); psh = C.byref(sh)
st = torch.cuda.current_stream().cuda_stream
wx, bx = m._wx().float(), m.in_project_x.bias.detach()
ws, bs = m.in_project_slice.weight.detach().contiguous(), m.in_project_slice.bias.detach()
wt, bt = m.ada_temp_linear.weight.detach().view(-1).contiguous(), m._bt()
_lib.check(l.haet_set_path(0))
spill = torch.zeros(l.haet_spill_bytes(psh), dtype=torch.uint8, device="cuda")
work = torch.zeros(l.haet_slice_workspace_bytes(psh), dtype=torch.uint8, device="cuda")
part = torch.zeros(l.haet_partial_floats(psh), device="cuda")
_lib.check(l.haet_slice_reduce(psh, x.data_ptr(), wx.data_ptr(), bx.data_ptr(), ws.data_ptr(), bs.data_ptr(), wt.data_ptr(), bt,
                               spill.data_ptr(), part.data_ptr(), work.data_ptr(), st))
torch.cuda.synchronize()
xp = spill.view(torch.float32)[:N * 256].view(N, 8, 32).double()
chunk = 320; nch = (N + chunk - 1) // chunk
parts = work.view(torch.float32)[:nch * 8 * G * 34].view(nch, 8, G, 34).double()
c = float(torch.log(-torch.log(torch.tensor(1e-6))))
tau = 0.5 + (xp @ wt.double() + bt).clamp(-0.4, 0.4)          # [N, 8]
z2 = ((xp @ ws.double().t() + bs.double() - c) / tau[..., None]) * math.log2(math.e)   # [N, 8, G]
h, g = 3, 89
for ch in range(nch):
    zz = z2[ch * chunk:(ch + 1) * chunk, h, g]
    xx = xp[ch * chunk:(ch + 1) * chunk, h]
    mt, lt = parts[ch, h, g, 0], parts[ch, h, g, 1]
    e = torch.exp2(zz - mt)
    l_ref = e.sum(); S_ref = (e[:, None] * xx).sum(0)
    St = parts[ch, h, g, 2:]
    # per tile max
    tmaxs = [float(zz[i:i + 64].max()) for i in range(0, zz.numel(), 64)]
    print(f"chunk {ch}: m_tc={float(mt):.3f} truemax={float(zz.max()):.3f} l_tc={float(lt):.5e} l_ref={float(l_ref):.5e} "
          f"S relerr={float((St - S_ref).norm() / S_ref.norm()):.2e}  tilemax={['%.1f' % v for v in tmaxs]}")
