"""In-graph timeline of the igemm_kernel launches of one ArcFace and one AEI_Net forward (B=64, bf16, CUDA-graph
replay): first-CTA-entry / last-CTA-exit of every launch by %globaltimer, so per-launch duration and the gap to
the next igemm launch (launch latency + whatever non-igemm kernels sit between) can be read without a profiler.
Needs the profiling build (`make -C ghost_b200/csrc PROF=1`)."""
import ctypes as C
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("GHOST_B200_LIB", os.path.join(_ROOT, "ghost_b200", "libghost_b200_prof.so"))

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import ghost_b200 as gb
from ghost_b200 import _lib

lib = _lib.load()
lib.ghost_prof_timeline_reset.restype = C.c_int
lib.ghost_prof_timeline_read.restype = C.c_int
lib.ghost_prof_timeline_read.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
SLOTS = 8192
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64


def timeline(label, fn):
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    assert lib.ghost_prof_timeline_reset() == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    t0 = np.zeros(SLOTS, dtype=np.uint64)
    t1 = np.zeros(SLOTS, dtype=np.uint64)
    meta = np.zeros((SLOTS, 12), dtype=np.int32)
    n = lib.ghost_prof_timeline_read(C.c_void_p(t0.ctypes.data), C.c_void_p(t1.ctypes.data), C.c_void_p(meta.ctypes.data))
    t0 = t0[:n].astype(np.int64)
    t1 = t1[:n].astype(np.int64)
    dur = (t1 - t0) / 1e3
    gap = np.append((t0[1:] - t1[:-1]) / 1e3, 0.0)
    print("== %s: %d igemm launches, forward %.3f ms by events, first entry -> last exit %.3f ms, sum of igemm durations "
          "%.3f ms, sum of gaps %.3f ms" % (label, n, e0.elapsed_time(e1), (t1[-1] - t0[0]) / 1e6, dur.sum() / 1e3,
                                             gap.sum() / 1e3))
    print("   idx  mode  BN  tiles  num_kb  grid win   dur_us  gap_after_us | CTA 0 cycles: prologue, issuer done, epilogue done, exit; last tile: accumulator ready, staged")
    for i in range(n):
        m = meta[i]
        print("   %3d  %4d %3d %6d %6d %5d %3d %8.2f %8.2f | %6d %7d %7d %7d ; %7d %7d %7d %7d" % (
            i, m[0] // 1000, m[0] % 1000, m[1], m[2], m[3] // 10, m[3] % 10, dur[i], gap[i], m[4], m[5], m[6], m[7], m[8],
            m[9], m[10], m[11]))


arc = gb.iresnet100().cuda().half().eval()
arc.enable_graph(True)
x = torch.randn(B, 3, 112, 112, device="cuda")
timeline("iresnet100 B=%d" % B, lambda: arc(x))

G = gb.AEI_Net("unet", num_blocks=2, c_id=512).cuda().half().eval()
G.enable_graph(True)
Xt = torch.randn(B, 3, 256, 256, device="cuda")
z = torch.nn.functional.normalize(torch.randn(B, 512, device="cuda"))
timeline("AEI_Net unet/2 B=%d" % B, lambda: G(Xt, z))
