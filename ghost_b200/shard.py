"""Frame / crop sharder: independent frames are split into contiguous blocks, one per rank;
weights are replicated; there is no collective inside the forward.  The only exchange is the final
gather of finished uint8 frames on the consumer rank.

One process per GPU under torch.distributed (NCCL on the GPUs, gloo in the CPU tests)."""
import os

import torch
import torch.distributed as dist


def shard_range(n_items, rank, world):
    """Contiguous block [lo, hi) of rank `rank`: the first n_items % world ranks get one extra item."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class FrameSharder(object):
    """Owns the rank's frame range and gathers finished frames in frame order."""

    def __init__(self, n_frames, rank=None, world=None, group=None):
        if rank is None:
            rank = dist.get_rank(group) if dist.is_initialized() else 0
        if world is None:
            world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.n_frames, self.rank, self.world, self.group = n_frames, rank, world, group
        self.lo, self.hi = shard_range(n_frames, rank, world)

    @property
    def local_count(self):
        return self.hi - self.lo

    def local_indices(self):
        return range(self.lo, self.hi)

    def batches(self, batch_size, align=1):
        """Yield (lo, hi) sub-ranges of the local block; batch boundaries are multiples of `align`
        (faces of one frame stay together: align = faces per frame)."""
        step = max(align, batch_size // align * align)
        i = self.lo
        while i < self.hi:
            j = min(self.hi, i + step)
            yield i, j
            i = j

    def gather(self, local_frames, dst=0):
        """Gather [local_count, ...] uint8 tensors into [n_frames, ...] on rank `dst` (None elsewhere).
        Ranks may own different counts; each sends exactly its own frames."""
        if self.world == 1:
            return local_frames
        self.gather_async(local_frames, dst)
        return self.wait()

    def gather_async(self, local_frames, dst=0):
        """Start the gather of one step's finished frames and return at once: the transfer is ordered after the work
        already queued on the current stream and runs beside whatever is queued next (NCCL's own stream on the GPUs),
        so step k's frames travel while step k+1 computes.  At most two gathers are in flight: starting gather k+1
        first orders the current stream after gather k-1, so the caller may reuse a `local_frames` buffer two steps
        later (double-buffer it), and rank `dst` alternates between two frame-ordered receive buffers.
        wait() returns the most recently started gather's result."""
        if self.world == 1:
            self._last = local_frames
            return
        if local_frames.shape[0] != self.local_count:
            raise ValueError("expected %d local frames, got %d" % (self.local_count, local_frames.shape[0]))
        pending = getattr(self, "_pending", [])
        while len(pending) > 1:              # keep one older gather in flight, retire the rest
            self._retire(pending.pop(0))
        tail = tuple(local_frames.shape[1:])
        works, out = [], None
        if self.rank == dst:
            bufs = getattr(self, "_bufs", None)
            if bufs is None or bufs[0].shape[1:] != tail or bufs[0].dtype != local_frames.dtype:
                bufs = self._bufs = [torch.empty((self.n_frames,) + tail, dtype=local_frames.dtype,
                                                 device=local_frames.device) for _ in range(2)]
                self._turn = 0
            out = bufs[self._turn % 2]
            self._turn += 1
            ops = []
            for r in range(self.world):
                lo, hi = shard_range(self.n_frames, r, self.world)
                if r == dst:
                    out[lo:hi].copy_(local_frames)
                elif hi > lo:
                    ops.append(dist.P2POp(dist.irecv, out[lo:hi], r, self.group))
            if ops:
                works = dist.batch_isend_irecv(ops)      # one grouped launch for all the peers
        elif self.local_count > 0:
            works = [dist.isend(local_frames.contiguous(), dst=dst, group=self.group)]
        pending.append((works, out))
        self._pending = pending

    def _retire(self, item):
        works, out = item
        for w in works:
            w.wait()        # CUDA: the current stream waits for the transfer; gloo: the host does
        self._last = out

    def wait(self):
        """Order the current stream (gloo: the host) after every gather started so far; returns the last one's
        frame-ordered tensor on the destination rank, None elsewhere."""
        pending = getattr(self, "_pending", [])
        while pending:
            self._retire(pending.pop(0))
        return getattr(self, "_last", None)


def bind_to_gpu_numa(device_index):
    """Pin this process (one per GPU) to the CPUs of the NUMA node its GPU hangs off, so that the pinned host buffers
    allocated afterwards (first touch) and the copy threads sit next to the GPU's PCIe root port.  Reads the PCI
    bus id from torch and the node from sysfs; returns the node, or None when the platform does not say (single
    socket, container without sysfs) - then nothing is changed."""
    import torch
    try:
        bdf = torch.cuda.get_device_properties(device_index).pci_bus_id
        dom = torch.cuda.get_device_properties(device_index).pci_domain_id
        dev = torch.cuda.get_device_properties(device_index).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bdf, dev)
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = []
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.extend(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except (OSError, ValueError, AttributeError, RuntimeError):
        return None
