"""Streaming HOST point clouds through a layer.

The reference keeps every mesh on the device for the whole model; a caller that holds meshes in host memory pays one
host->device copy of the points and one device->host copy of the result per layer call, and for the 1M-point layer
those two copies (1 GB each over PCIe) are 10x the kernel time.  A single call cannot overlap them with its own
compute -- the slice softmax runs over ALL points, so the first output row needs the last input row -- but
consecutive calls are independent: `HostStream` runs request i + 1's upload, request i's kernels and request i - 1's
download concurrently on `depth` CUDA streams (PCIe is full duplex, the two copy engines are separate).

    hs = HostStream(layer, depth=2)
    for xh, oh in batches:            # pinned host tensors [B, N, dim]
        hs.submit(xh, oh)             # asynchronous; returns the event that marks `oh` complete
    hs.drain()
"""
import os

import torch


def bind_host_to_gpu_numa(device=None) -> dict:
    """Restrict this process to the CPUs of the NUMA node its GPU hangs off, so that the pinned host buffers it allocates
    afterwards (first-touch placement) and the threads that fill them are local to the GPU's PCIe root.  With one process per
    GPU this spreads the host-memory traffic of N ranks over the sockets instead of piling it on the node the launcher
    started on -- the end-to-end (host-resident) path is bound by PCIe and host-memory bandwidth, not by the kernels.
    A no-op (reported as such) when the platform exposes no NUMA information or the node's CPUs are outside the
    process' allowed set (a container pinned to one node)."""
    info = {"bound": False}
    try:
        idx = torch.device(device).index if device is not None else torch.cuda.current_device()
        if idx is None:
            idx = torch.cuda.current_device()
        p = torch.cuda.get_device_properties(idx)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as fh:
            node = int(fh.read().strip())
        info.update(gpu=idx, pci=bdf, numa_node=node)
        if node < 0:
            return info
        with open(f"/sys/devices/system/node/node{node}/cpulist") as fh:
            cpus = set()
            for part in fh.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        info.update(node_cpus=len(cpus), allowed_cpus=len(allowed), usable_cpus=len(use))
        if use and use != allowed:
            os.sched_setaffinity(0, use)
            info["bound"] = True
    except Exception as ex:   # no sysfs / no such attribute: leave the affinity alone
        info["error"] = f"{type(ex).__name__}: {ex}"
    return info


class HostStream:
    def __init__(self, module, depth: int = 2, device=None):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.module = module
        self.device = torch.device(device) if device is not None else next(module.parameters()).device
        if self.device.type != "cuda":
            raise RuntimeError("HostStream needs the module on a CUDA device (no CPU fallback)")
        self.streams = [torch.cuda.Stream(self.device) for _ in range(depth)]
        self.done = [None] * depth
        self.count = 0

    def submit(self, x_host: torch.Tensor, out_host: torch.Tensor) -> torch.cuda.Event:
        """Queue one forward of `x_host` (pinned, [B, N, dim]) into `out_host` (pinned, same shape)."""
        if x_host.is_cuda or out_host.is_cuda:
            raise ValueError("HostStream.submit takes host tensors")
        if not (x_host.is_pinned() and out_host.is_pinned()):
            raise ValueError("HostStream.submit needs pinned host tensors (tensor.pin_memory())")
        if out_host.shape != x_host.shape:
            raise ValueError(f"out_host {tuple(out_host.shape)} must match x_host {tuple(x_host.shape)}")
        slot = self.count % len(self.streams)
        s = self.streams[slot]
        s.wait_stream(torch.cuda.current_stream(self.device))   # whatever produced / last used the host buffers
        with torch.cuda.stream(s):
            xd = x_host.to(self.device, non_blocking=True)
            out_host.copy_(self.module(xd), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(s)
        self.done[slot] = ev
        self.count += 1
        return ev

    def drain(self, join_current: bool = True) -> None:
        """Wait (host side) until every submitted request has landed in its host buffer."""
        for s in self.streams:
            if join_current:
                torch.cuda.current_stream(self.device).wait_stream(s)
        for ev in self.done:
            if ev is not None:
                ev.synchronize()
