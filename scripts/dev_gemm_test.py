import ctypes as C, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from haet_b200 import _lib
l = _lib.lib()
torch.manual_seed(0)
st = torch.cuda.current_stream().cuda_stream
def run(M, K, Nn, path):
    x = torch.randn(M, K, device="cuda"); w = torch.randn(Nn, K, device="cuda") * 0.05; b = torch.randn(Nn, device="cuda")
    y = torch.zeros(M, Nn, device="cuda"); img = torch.zeros(4 * Nn * K, dtype=torch.uint8, device="cuda")
    _lib.check(l.haet_set_path(path))
    _lib.check(l.haet_debug_gemm(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), M, K, Nn, img.data_ptr(), st))
    torch.cuda.synchronize()
    ref = (x.double() @ w.double().t() + b.double())
    err = float((y.double() - ref).norm() / ref.norm()); mx = float((y.double() - ref).abs().max())
    return err, mx, (x, w, b, y, img)
for (M, K, Nn) in [(128, 64, 128), (128, 256, 256), (1000, 256, 256), (100000, 256, 256), (777, 128, 128), (4099, 64, 256), (5000, 256, 128)]:
    for path in (1, 0):
        err, mx, _ = run(M, K, Nn, path)
        print(f"M={M} K={K} Nn={Nn} path={path} rel_l2={err:.3e} maxabs={mx:.3e}", flush=True)
# timing at 1M x 256 x 256
for path in (0, 1):
    err, mx, (x, w, b, y, img) = run(1_000_000, 256, 256, path)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        _lib.check(l.haet_debug_gemm(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), 1_000_000, 256, 256, img.data_ptr(), st))
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"1M x256x256 path={path}: {ms:.3f} ms  ({2e6*256*256/ms/1e9:.1f} TFLOP/s algorithmic, {2*1.024/ms*1e3:.0f} GB/s) rel_l2={err:.2e}", flush=True)
