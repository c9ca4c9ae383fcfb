"""Correctness + timing of the Nn = 256 GEMM through both kernels (HAET_GEMM_1CTA selects the single-CTA one)."""
import sys, os, subprocess
if len(sys.argv) == 1:
    for env in ({}, {"HAET_GEMM_DBG": "1"}, {"HAET_GEMM_DBG": "2"}, {"HAET_GEMM_DBG": "3"}, {"HAET_GEMM_1CTA": "1"}):
        e = dict(os.environ); e.update(env)
        print("env", env, flush=True)
        subprocess.run([sys.executable, __file__, "child"], env=e)
    sys.exit(0)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from haet_b200 import _lib
l = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
for (M, K, Nn) in [(1000, 256, 256), (4099, 64, 256), (1_000_000, 256, 256)]:
    x = torch.randn(M, K, device="cuda"); w = torch.randn(Nn, K, device="cuda") * 0.05; b = torch.randn(Nn, device="cuda")
    y = torch.zeros(M, Nn, device="cuda"); img = torch.zeros(4 * Nn * K, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        _lib.check(l.haet_debug_gemm(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), M, K, Nn, img.data_ptr(), st))
    torch.cuda.synchronize()
    ref = x[:4099].double() @ w.double().t() + b.double()
    err = float((y[:4099].double() - ref).norm() / ref.norm())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        _lib.check(l.haet_debug_gemm(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), M, K, Nn, img.data_ptr(), st))
    e1.record(); torch.cuda.synchronize()
    print(f"M={M} K={K} Nn={Nn}: rel_l2={err:.2e}  {e0.elapsed_time(e1)/5:.3f} ms", flush=True)
