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


class _DevMem(object):
    """A raw device allocation as a torch-importable buffer (__cuda_array_interface__)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 3, "strides": None}


class PeerGather(object):
    """The final frame gather over peer-mapped memory - GPUs of one box, one process per GPU.

    The consumer rank `dst` owns two frame-ordered buffers [n_frames, ...] (cudaMalloc + CUDA IPC handle,
    ghost_peer_alloc); every producer maps them once and moves its finished frames with ONE cudaMemcpyAsync per step
    on its own copy stream: a copy-engine transfer over NVLink that takes no SM on either side, so step k's frames
    travel while step k+1 computes at full speed (an NCCL send / recv pair runs a kernel on both GPUs).  A 32-bit flag
    per (buffer, producer) in the consumer's memory carries the step number behind the copy; the consumer's stream
    waits for it with cuStreamWaitValue32, and the consumer returns credits the same way, so nothing on the host ever
    blocks.  Same interface and frame order as FrameSharder.gather_async / wait: the result of wait() is valid until
    the second next gather_async() on the consumer.  Construction is collective (handles travel through
    torch.distributed)."""

    def __init__(self, n_frames, frame_shape, dtype=torch.uint8, dst=0, group=None, device=None):
        import ctypes as C
        from . import _lib
        self._C, self._lib = C, _lib.load()
        self._check = _lib.check
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.n_frames, self.dst, self.group = int(n_frames), int(dst), group
        self.lo, self.hi = shard_range(self.n_frames, self.rank, self.world)
        self.dev = device or torch.device("cuda", torch.cuda.current_device())
        self.frame_shape = tuple(int(v) for v in frame_shape)
        self.dtype = dtype
        self.frame_bytes = int(torch.empty((), dtype=dtype).element_size())
        for v in self.frame_shape:
            self.frame_bytes *= v
        self.k = 0                                   # gathers started
        self.s_copy = torch.cuda.Stream(device=self.dev)
        self.ev_ready = torch.cuda.Event()
        self.ev_copied = [torch.cuda.Event(), torch.cuda.Event()]
        lib = self._lib
        # per rank: a flag block of 2 * world + 2 words.  consumer: words [b * world + r] = "producer r's frames of
        # a step with buffer b have landed" (step number + 1); producers: word [2 * world + b] = credit for buffer b
        handles = [None] * self.world
        mine = {}
        with torch.cuda.device(self.dev):
            fp, fh = C.c_void_p(), C.create_string_buffer(64)
            self._check(lib.ghost_peer_alloc(4 * (2 * self.world + 2), C.byref(fp), fh))
            self._flags = fp.value
            mine["flags"] = fh.raw
            self._bufs = []
            if self.rank == self.dst:
                for _ in range(2):
                    bp, bh = C.c_void_p(), C.create_string_buffer(64)
                    self._check(lib.ghost_peer_alloc(self.n_frames * self.frame_bytes, C.byref(bp), bh))
                    self._bufs.append(bp.value)
                    mine.setdefault("bufs", []).append(bh.raw)
        dist.all_gather_object(handles, mine, group=group)
        self._peer_flags = {}
        self._remote_bufs = []
        with torch.cuda.device(self.dev):
            if self.rank == self.dst:
                for r in range(self.world):
                    if r != self.dst:
                        p = C.c_void_p()
                        self._check(lib.ghost_peer_open(handles[r]["flags"], C.byref(p)))
                        self._peer_flags[r] = p.value
                self._views = [torch.as_tensor(_DevMem(b, self.n_frames * self.frame_bytes), device=self.dev)
                               .view(dtype).view((self.n_frames,) + self.frame_shape) for b in self._bufs]
            else:
                p = C.c_void_p()
                self._check(lib.ghost_peer_open(handles[self.dst]["flags"], C.byref(p)))
                self._peer_flags[self.dst] = p.value
                for h in handles[self.dst]["bufs"]:
                    q = C.c_void_p()
                    self._check(lib.ghost_peer_open(h, C.byref(q)))
                    self._remote_bufs.append(q.value)
        dist.barrier(group=group)

    def _sp(self, stream):
        return self._C.c_void_p(stream.cuda_stream)

    def local_out(self):
        """Consumer rank: its own slot of the buffer the NEXT gather_async() fills - a step that writes its frames
        straight into this view saves the local copy (pass the same tensor to gather_async).  None on producers."""
        if self.rank != self.dst or self.hi <= self.lo:
            return None
        return self._views[self.k & 1][self.lo:self.hi]

    def gather_async(self, local_frames):
        """Start the gather of one step: local_frames [local_count, ...] on this rank's device, finished on the
        current stream.  Returns at once."""
        C, lib = self._C, self._lib
        if local_frames.shape[0] != self.hi - self.lo or tuple(local_frames.shape[1:]) != self.frame_shape:
            raise ValueError("expected %d local frames of shape %s" % (self.hi - self.lo, self.frame_shape))
        k, b = self.k, self.k & 1
        self.k += 1
        main = torch.cuda.current_stream(self.dev)
        with torch.cuda.device(self.dev):
            if self.rank == self.dst:
                if k >= 2:
                    # the caller is done with the result of step k - 2: its buffer may be written again
                    for r, fl in self._peer_flags.items():
                        self._check(lib.ghost_peer_signal(C.c_void_p(fl + 4 * (2 * self.world + b)), k, self._sp(main)))
                if self.hi > self.lo and local_frames.data_ptr() != self._views[b][self.lo:self.hi].data_ptr():
                    self._views[b][self.lo:self.hi].copy_(local_frames)
            elif self.hi > self.lo:
                src = local_frames.contiguous()
                self.ev_ready.record(main)
                self.s_copy.wait_event(self.ev_ready)
                sc = self._sp(self.s_copy)
                if k >= 2:
                    self._check(lib.ghost_peer_wait(C.c_void_p(self._flags + 4 * (2 * self.world + b)), k, sc))
                self._check(lib.ghost_peer_copy(C.c_void_p(self._remote_bufs[b] + self.lo * self.frame_bytes),
                                                C.c_void_p(src.data_ptr()), (self.hi - self.lo) * self.frame_bytes, sc))
                self._check(lib.ghost_peer_signal(C.c_void_p(self._peer_flags[self.dst] + 4 * (b * self.world + self.rank)),
                                                  k + 1, sc))
                self.ev_copied[b].record(self.s_copy)
                self._keep = getattr(self, "_keep", [None, None])
                self._keep[b] = src          # the source stays alive until its copy has been ordered behind
                # the caller may overwrite a frame buffer two steps later: order the main stream behind copy k - 1
                if k >= 1:
                    main.wait_event(self.ev_copied[b ^ 1])

    def wait(self):
        """Consumer: order the current stream behind the arrival of every producer's frames of the last started
        gather and return the frame-ordered tensor; producers: order the current stream behind their own copy."""
        C, lib = self._C, self._lib
        if self.k == 0:
            return None
        k, b = self.k - 1, (self.k - 1) & 1
        main = torch.cuda.current_stream(self.dev)
        with torch.cuda.device(self.dev):
            if self.rank == self.dst:
                for r in range(self.world):
                    lo, hi = shard_range(self.n_frames, r, self.world)
                    if r != self.dst and hi > lo:
                        self._check(lib.ghost_peer_wait(C.c_void_p(self._flags + 4 * (b * self.world + r)), k + 1,
                                                        self._sp(main)))
                return self._views[b]
            if self.hi > self.lo:
                main.wait_event(self.ev_copied[b])
        return None

    def close(self):
        """Collective: unmap / free the peer buffers (after a device synchronize on every rank)."""
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
        lib = self._lib
        with torch.cuda.device(self.dev):
            for p in list(self._peer_flags.values()) + self._remote_bufs:
                lib.ghost_peer_close(self._C.c_void_p(p))
            dist.barrier(group=self.group)
            self._views = []
            for p in self._bufs + [self._flags]:
                lib.ghost_peer_free(self._C.c_void_p(p))
        self._peer_flags, self._remote_bufs, self._bufs = {}, [], []


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
