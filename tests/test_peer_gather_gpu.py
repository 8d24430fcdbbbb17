"""GPU, two processes on two GPUs of one box: ghost_b200.PeerGather (the final frame gather over CUDA-IPC peer
memory) delivers every rank's frames once, in frame order, over several pipelined steps, with uneven
shards - the same properties tests/test_shard_gloo.py checks for the NCCL / gloo path.  Skipped on one-GPU boxes."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_frames, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import ghost_b200
        shape = (6, 10, 3)
        pg = ghost_b200.PeerGather(n_frames, shape, dst=0, device=dev)
        lo, hi = ghost_b200.shard_range(n_frames, rank, world)
        bufs = [torch.empty((hi - lo,) + shape, dtype=torch.uint8, device=dev) for _ in range(2)]
        ok = True
        steps = 5
        for step in range(steps):
            own = pg.local_out() if step % 2 else None          # consumer: alternate zero-copy slot / own buffer
            t = own if own is not None else bufs[step % 2]
            idx = torch.arange(lo, hi, device=dev, dtype=torch.int32).view(-1, 1, 1, 1)
            t.copy_(((idx * 7 + step * 31) % 251).to(torch.uint8).expand(hi - lo, *shape))
            pg.gather_async(t)
            res = pg.wait()
            if rank == 0:
                torch.cuda.synchronize()
                want = ((torch.arange(n_frames, dtype=torch.int32) * 7 + step * 31) % 251).to(torch.uint8)
                got = res[:, 0, 0, 0].cpu()
                ok = ok and bool(torch.equal(got, want)) and bool((res == res[:, :1, :1, :1]).all().item())
            else:
                assert res is None
        pg.close()
        if rank == 0:
            q.put(ok)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one box")
@pytest.mark.parametrize("n_frames", [7, 64])
def test_peer_gather_in_frame_order_two_gpus(n_frames):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_frames, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
