"""PCIe copy rates of the box (pinned host <-> device, 1 GB): one copy, k concurrent chunk copies, both directions at once.
    python scripts/pcie_probe.py"""
import torch, time
dev = torch.device("cuda")
n = 256_000_000  # floats = 1.024 GB
h_in = torch.empty(n, dtype=torch.float32).pin_memory()
h_out = torch.empty(n, dtype=torch.float32).pin_memory()
d_in = torch.empty(n, dtype=torch.float32, device=dev)
d_out = torch.empty(n, dtype=torch.float32, device=dev)
h_in.fill_(1.0)
streams = [torch.cuda.Stream() for _ in range(8)]

def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps

def h2d(k):
    c = n // k
    for i in range(k):
        with torch.cuda.stream(streams[i]):
            d_in[i * c:(i + 1) * c].copy_(h_in[i * c:(i + 1) * c], non_blocking=True)

def d2h(k):
    c = n // k
    for i in range(k):
        with torch.cuda.stream(streams[4 + i % 4]):
            h_out[i * c:(i + 1) * c].copy_(d_out[i * c:(i + 1) * c], non_blocking=True)

gb = n * 4 / 1e9
for k in (1, 2, 4):
    print(f"H2D x{k}: {gb / timed(lambda: h2d(k)):.1f} GB/s   D2H x{k}: {gb / timed(lambda: d2h(k)):.1f} GB/s   "
          f"both x{k}: {gb / timed(lambda: (h2d(k), d2h(k))):.1f} GB/s per direction")
