import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from haet_b200 import _lib
l = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
M, K, Nn = 1_000_000, 256, 256
x = torch.randn(M, K, device="cuda"); w = torch.randn(Nn, K, device="cuda") * 0.05; b = torch.randn(Nn, device="cuda")
y = torch.zeros(M, Nn, device="cuda"); img = torch.zeros(4 * Nn * K, dtype=torch.uint8, device="cuda")
for _ in range(3):
    _lib.check(l.haet_debug_gemm(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), M, K, Nn, img.data_ptr(), st))
torch.cuda.synchronize()
print("ok")
