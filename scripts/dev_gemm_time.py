"""Timing probe: Y = X W^T through haet_debug_gemm for a few (K, Nn) at 1M rows (diagnostic)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from haet_b200 import _lib
l = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
M = 1_000_000
for (K, Nn) in [(256, 256), (256, 128), (128, 256), (128, 128), (64, 128)]:
    x = torch.randn(M, K, device="cuda"); w = torch.randn(Nn, K, device="cuda") * 0.05; b = torch.randn(Nn, device="cuda")
    y = torch.zeros(M, Nn, device="cuda"); img = torch.zeros(4 * Nn * K, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        _lib.check(l.haet_debug_gemm(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), M, K, Nn, img.data_ptr(), st))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        _lib.check(l.haet_debug_gemm(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), M, K, Nn, img.data_ptr(), st))
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    gb = M * (K + Nn) * 4 / 1e9
    print(f"K={K} Nn={Nn}: {ms:.3f} ms  {gb/ms*1e3:.0f} GB/s of unique HBM traffic", flush=True)
