"""Per-role stall counters of igemm_kernel for a few conv shapes.  Needs the profiling build of the library
(`make -C ghost_b200/csrc PROF=1`, which adds clock64() stamps around the barrier waits and the
ghost_prof_igemm_read symbol); rebuild without PROF afterwards - the counters cost a few percent."""
import ctypes as C
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("GHOST_B200_LIB", os.path.join(_ROOT, "ghost_b200", "libghost_b200_prof.so"))

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
from ghost_b200 import _lib
from util import fptr, nhwc

lib = _lib.load()
read = lib.ghost_prof_igemm_read
read.argtypes = [C.c_void_p]
read.restype = C.c_int
g = torch.Generator().manual_seed(0)
shapes = [(64, 64, 56, 64), (64, 128, 28, 128), (64, 256, 14, 256), (64, 512, 7, 512), (64, 512, 32, 512),
          (32, 64, 256, 64), (64, 128, 128, 64), (64, 256, 64, 256)]
if len(sys.argv) > 1:
    shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
for (N, Cin, H, Cout) in shapes:
    x = nhwc(torch.randn(N, Cin, H, H, generator=g).cuda(), torch.bfloat16)
    w = np.ascontiguousarray((torch.randn(Cout, Cin, 3, 3, generator=g) * 0.05).numpy().astype(np.float32))
    y = torch.empty((N, H, H, Cout), dtype=torch.bfloat16, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for bn in (0, 64, 128, 256):
        if bn > Cout:
            continue

        def run():
            rc = lib.ghost_test_conv(C.c_void_p(x.data_ptr()), N, H, H, Cin, fptr(w), None, Cout, 0, 3, 1, 1, 0, 0.0,
                                     C.c_void_p(y.data_ptr()), _lib.BF16, bn, st)
            _lib.check(rc, test_api=True)

        for _ in range(3):
            run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        buf = np.zeros((160, 12), dtype=np.uint64)
        assert read(C.c_void_p(buf.ctypes.data)) == 0
        cfg = buf[159].astype(np.int64)
        grid = int(cfg[7])
        b = buf[:grid].astype(np.float64)
        fl = 2.0 * N * H * H * Cout * Cin * 9
        us = e0.elapsed_time(e1) * 1e3
        kbs = cfg[3] * cfg[4]
        med = np.median(b, axis=0)
        print("N=%d Cin=%d H=%d Cout=%d hint=%3d | BN=%d stages=%d dual=%d num_kb=%d tiles/cta(0)=%d win=%d bres=%d grid=%d"
              " | %.1f us (host events incl. weight pack) %.0f TF/s" % (N, Cin, H, Cout, bn, *cfg[:8], us, fl / us / 1e6))
        print("   issuer: total %.0f cyc, wait full %.0f (%.0f%%), wait tmem-empty %.0f (%.0f%%); per K block %.0f cyc"
              % (med[2], med[0], 100 * med[0] / med[2], med[1], 100 * med[1] / med[2], med[2] / max(kbs, 1)))
        print("   A producer: total %.0f cyc, wait empty %.0f (%.0f%%), wait h %.0f; epilogue warp: wait tfull %.0f over %d tiles"
              % (med[5], med[3], 100 * med[3] / max(med[5], 1), med[4], med[6], int(med[7])))
        print("   CTA: prologue %.0f cyc, epilogue done at %.0f, exit at %.0f cyc (max %.0f) = %.2f us by globaltimer (max %.2f)"
              % (med[9], med[10], med[8], b[:, 8].max(), med[11] / 1e3, b[:, 11].max() / 1e3))
    del x, y
