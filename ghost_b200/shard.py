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
        if local_frames.shape[0] != self.local_count:
            raise ValueError("expected %d local frames, got %d" % (self.local_count, local_frames.shape[0]))
        tail = tuple(local_frames.shape[1:])
        if self.rank == dst:
            out = torch.empty((self.n_frames,) + tail, dtype=local_frames.dtype, device=local_frames.device)
            reqs = []
            for r in range(self.world):
                lo, hi = shard_range(self.n_frames, r, self.world)
                if r == dst:
                    out[lo:hi].copy_(local_frames)
                elif hi > lo:
                    reqs.append(dist.irecv(out[lo:hi], src=r, group=self.group))
            for q in reqs:
                q.wait()
            return out
        if self.local_count > 0:
            dist.send(local_frames.contiguous(), dst=dst, group=self.group)
        return None


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
