"""CPU, world_size 2 (gloo): the frame sharder's final gather returns every frame once, in frame order,
with uneven shards and with a rank that owns nothing."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ghost_b200 import FrameSharder, shard_range


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_frames, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = FrameSharder(n_frames)
        assert (sh.rank, sh.world) == (rank, world) and (sh.lo, sh.hi) == shard_range(n_frames, rank, world)
        # every "frame" carries its global index, as a finished uint8 frame would
        local = torch.stack([torch.full((4, 6, 3), i % 251, dtype=torch.uint8) for i in sh.local_indices()]) \
            if sh.local_count else torch.empty((0, 4, 6, 3), dtype=torch.uint8)
        out = sh.gather(local, dst=0)
        if rank == 0:
            ok = out.shape[0] == n_frames and all(int(out[i, 0, 0, 0]) == i % 251 for i in range(n_frames))
            q.put(bool(ok))
        else:
            assert out is None
        # pipelined form: three steps in flight two deep, every step's frames arrive in order and intact
        for step in range(3):
            loc = (local.to(torch.int32) + step).to(torch.uint8) if sh.local_count else local
            sh.gather_async(loc, dst=0)     # the caller computes step + 1 while this travels
        last = sh.wait()
        if rank == 0:
            ok2 = all(int(last[i, 0, 0, 0]) == (i % 251 + 2) % 256 for i in range(n_frames))
            q.put(bool(ok2))
        else:
            assert last is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_frames", [7, 1, 300])
def test_gather_in_frame_order_world2(n_frames):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_frames, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
    assert q.get(timeout=5) is True


def test_numa_binding_is_a_noop_without_topology():
    """bind_to_gpu_numa: no CUDA device / no NUMA information -> None, CPU affinity untouched."""
    from ghost_b200.shard import bind_to_gpu_numa

    before = os.sched_getaffinity(0)
    assert bind_to_gpu_numa(0) is None or isinstance(bind_to_gpu_numa(0), int)
    if not torch.cuda.is_available():
        assert os.sched_getaffinity(0) == before
