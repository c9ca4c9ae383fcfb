"""Cycle accounting of pass 2's warp roles (development): builds libhaet_b200_trace.so with -DHAET_TRACE (clock64 counters in
deslice_tc.cu), runs the config-4 layer and prints the average cycles per item each role spends waiting / working.

    python scripts/trace_pass2.py build      # here (nvcc cross-compiles)
    HAET_B200_LIB=$PWD/haet_b200/libhaet_b200_trace.so python scripts/trace_pass2.py run     # on the GPU box
"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
TRACE_LIB = os.path.join(ROOT, "haet_b200", "libhaet_b200_trace.so")

if sys.argv[1:] == ["build"]:
    from haet_b200 import build as B
    objs = []
    os.makedirs("/tmp/haet_trace", exist_ok=True)
    for src in B.sources():
        obj = os.path.join("/tmp/haet_trace", os.path.basename(src)[:-3] + ".o")
        subprocess.run([B._nvcc(), *B.ARCH, *B.NVCC_FLAGS, "-DHAET_TRACE", "-c", src, "-o", obj], check=True)
        objs.append(obj)
    subprocess.run([B._nvcc(), *B.ARCH, "-shared", "-o", TRACE_LIB, *objs, "-lcudart"], check=True)
    print(TRACE_LIB)
    sys.exit(0)

import torch
import bench
import haet_b200 as hb

l = hb.lib()
l.haet_debug_dt_trace.restype, l.haet_debug_dt_trace.argtypes = C.c_int, [C.c_void_p, C.c_int]
N, G = 1_000_000, int(sys.argv[2]) if len(sys.argv) > 2 else 256
m = bench.make_module(G, "cuda")
x = torch.nn.functional.layer_norm(torch.randn(1, N, 256, device="cuda"), (256,))
for _ in range(2):
    m(x)
torch.cuda.synchronize()
assert l.haet_debug_dt_trace(None, 1) == 0
l.haet_debug_st_trace.restype, l.haet_debug_st_trace.argtypes = C.c_int, [C.c_void_p, C.c_int]
assert l.haet_debug_st_trace(None, 1) == 0
m(x)
torch.cuda.synchronize()
buf = (C.c_ulonglong * 64)()
assert l.haet_debug_dt_trace(buf, 0) == 0
v = list(buf)
names = {0: "softmax: wait a_full", 1: "softmax: wait d_full", 2: "softmax: tcgen05.ld + wait (per half, summed)", 3: "softmax: exponentials (per half, summed)",
         4: "softmax: tcgen05.st issue", 5: "softmax: wait::st + fence + arrive", 6: "softmax: whole item", 16: "MMA1: wait p_free",
         17: "MMA1: wait a_full", 18: "MMA1: issue 7 MMAs + commit", 24: "MMA2: wait o_empty", 25: "MMA2: wait p_full", 26: "MMA2: issue + commits"}
print(f"-- per softmax warp: kernel entry -> loop start {v[8] / max(v[10], 1):.0f} cycles, loop {v[9] / max(v[10], 1):.0f} cycles; per CTA: entry -> all warps done {v[11] / max(v[12], 1):.0f} cycles")
l.haet_debug_st_trace.restype, l.haet_debug_st_trace.argtypes = C.c_int, [C.c_void_p, C.c_int]
b1 = (C.c_ulonglong * 64)()
assert l.haet_debug_st_trace(b1, 0) == 0
u = list(b1)
n = max(u[7], 1)
print(f"== pass 1: per softmax warp loop {u[9] / max(u[10], 1):.0f} cycles; items counted {u[7]}")
for k, nm in ((0, "softmax: wait b_full"), (1, "softmax: wait d_full"), (3, "softmax: both halves (ld, exponentials, st issue)"), (5, "softmax: wait::st + fence + arrive"), (6, "softmax: whole item")):
    print(f"   {nm:52s} {u[k] / n:9.1f} cycles / item")
n = max(u[23], 1)
print(f"   MMA issuer: wait p_full {u[16] / n:.1f}, wait b_full {u[17] / n:.1f}, issue 4 MMA2 + commits {u[18] / n:.1f}, issue 7 MMA1 + commit {u[19] / n:.1f} cycles / item ({u[23]} items)")
print("== pass 2")
for base, cnt in ((0, 7), (16, 19), (24, 27)):
    n = max(v[cnt], 1)
    print(f"-- items counted (warp lane 0): {v[cnt]}")
    for k in range(base, cnt):
        if k in names:
            print(f"   {names[k]:52s} {v[k] / n:9.1f} cycles / item")
