"""Host <-> device staging for the data-model level wrappers.

The reference's API hands over ordinary (pageable) numpy arrays of several GB.  Copying those to the
GPU directly goes through the driver's small bounce buffer at a few GB/s and, at the data-model
level, used to cost 50x the gridding itself.  Here every chunk is first copied -- in parallel, by
torch's multi-threaded CPU copy, converting the element type on the way when asked (int64 flags ->
uint8, complex128 -> complex64 for single-precision plans) -- into a pinned buffer from a small
ring, and goes from there to the device asynchronously at PCIe speed while the host fills the next
buffer.  Downloads run the same ring backwards.  Pinned buffers come from torch's caching host
allocator, so repeated calls do not pay for page-locking again.
"""
import numpy


class Stager:
    def __init__(self, torch, dev, slots=6):
        self.torch = torch
        self.dev = dev
        self.slots = [None] * slots
        self.events = [None] * slots
        self.pending = [None] * slots   # (pinned view, numpy destination) of downloads in flight
        self.i = 0
        self.stream = torch.cuda.current_stream(dev)

    def _slot(self, nbytes):
        torch = self.torch
        k = self.i % len(self.slots)
        self.i += 1
        if self.events[k] is not None:
            self.events[k].synchronize()
        self._flush(k)
        buf = self.slots[k]
        if buf is None or buf.numel() < nbytes:
            self.slots[k] = buf = torch.empty(int(nbytes), dtype=torch.uint8, pin_memory=True)
        if self.events[k] is None:
            self.events[k] = torch.cuda.Event()
        return k, buf

    def _flush(self, k):
        if self.pending[k] is not None:
            pin, dst = self.pending[k]
            self.torch.from_numpy(dst).copy_(pin)     # parallel host copy out of the pinned buffer
            self.pending[k] = None

    def put(self, arr, dtype=None):
        """numpy array -> device tensor of ``dtype`` (default: the array's own), asynchronously."""
        torch = self.torch
        src = torch.from_numpy(arr if arr.flags.c_contiguous else numpy.ascontiguousarray(arr))
        dtype = dtype or src.dtype
        n = src.numel()
        itemsize = torch.empty((), dtype=dtype).element_size()
        k, buf = self._slot(max(1, n * itemsize))
        pin = buf[: n * itemsize].view(dtype).reshape(src.shape)
        pin.copy_(src)                                 # multi-threaded, converts the element type
        out = torch.empty(src.shape, device=self.dev, dtype=dtype)
        out.copy_(pin, non_blocking=True)
        self.events[k].record(self.stream)
        return out

    def get(self, dev_tensor, dst):
        """device tensor -> numpy array ``dst`` (same shape / dtype, C-contiguous view), completed
        lazily: by the time the ring slot is reused or ``finish()`` is called."""
        n = dev_tensor.numel() * dev_tensor.element_size()
        k, buf = self._slot(max(1, n))
        pin = buf[:n].view(dev_tensor.dtype).reshape(dev_tensor.shape)
        pin.copy_(dev_tensor, non_blocking=True)
        self.events[k].record(self.stream)
        self.pending[k] = (pin, dst)

    def finish(self):
        for k in range(len(self.slots)):
            if self.events[k] is not None:
                self.events[k].synchronize()
            self._flush(k)


def bind_host_to_gpu(torch, index):
    """Pin the calling process to the CPUs next to GPU ``index`` (its PCIe root's NUMA node, from
    sysfs), so that pinned staging buffers are allocated in -- and host copies run from -- the
    memory the GPU's DMA engines reach without crossing the socket interconnect.  With several
    ranks per node this keeps each rank's host <-> device traffic on its own socket.  Returns a
    short description; does nothing when the platform does not say (containers, one node)."""
    import os
    try:
        props = torch.cuda.get_device_properties(index)
        bus = "%04x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
        base = "/sys/bus/pci/devices/" + bus
        with open(base + "/numa_node") as fh:
            node = int(fh.read().strip())
        if node < 0:
            return "unchanged (no NUMA node reported for %s)" % bus
        with open(base + "/local_cpulist") as fh:
            text = fh.read().strip()
        cpus = set()
        for part in text.split(","):
            if "-" in part:
                lo, hi = part.split("-")
                cpus.update(range(int(lo), int(hi) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return "unchanged (no allowed CPU on NUMA node %d)" % node
        os.sched_setaffinity(0, cpus)
        return "NUMA node %d of GPU %s, %d CPUs" % (node, bus, len(cpus))
    except (OSError, AttributeError, ValueError) as exc:
        return "unchanged (%s)" % type(exc).__name__
