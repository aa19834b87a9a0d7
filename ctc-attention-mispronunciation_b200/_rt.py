"""Small runtime helpers shared by the host-side drop-ins: device selection, workspace cache, stream handle."""
from __future__ import annotations

import torch

from . import _lib

_ws_cache = {}


def device(t=None) -> torch.device:
    """The CUDA device the call runs on.  Raises (no CPU fallback) when CUDA is unavailable."""
    if t is not None and isinstance(t, torch.Tensor) and t.is_cuda:
        # the C ABI launches on the CURRENT device: make the tensor's device current (one process may hold several)
        if t.device.index != torch.cuda.current_device():
            torch.cuda.set_device(t.device)
        return t.device
    if not torch.cuda.is_available():
        raise _lib.MddCtcLibraryError("mddctc needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def workspace(nbytes: int, dev: torch.device, tag: str = "ws") -> torch.Tensor:
    """Grow-only per-(device, tag) scratch buffer; the C ABI never allocates."""
    key = (dev.index, tag)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
        _ws_cache[key] = buf
    return buf


def stream_ptr(dev: torch.device) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


_copy_streams = {}


def copy_stream(dev: torch.device) -> "torch.cuda.Stream":
    """A per-device side stream for host -> device transfers that overlap the kernels of the current stream."""
    st = _copy_streams.get(dev.index)
    if st is None:
        st = torch.cuda.Stream(device=dev)
        _copy_streams[dev.index] = st
    return st


def bind_to_gpu_numa(index: int):
    """Best effort: restrict this process to the CPUs of the NUMA node GPU `index` hangs off, so that page-locked
    staging buffers allocated afterwards (first touch) sit next to the GPU's PCIe root.  Returns a short description
    of what was done (None when the topology is not visible)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        bus = bus.lower()
        if len(bus.split(":")[0]) == 8:                 # nvml prints an 8-digit PCI domain, sysfs a 4-digit one
            bus = bus[4:]
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & set(os.sched_getaffinity(0))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return "numa node %d (%d cpus)" % (node, len(allowed))
    except Exception:
        return None
