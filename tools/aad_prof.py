"""Stall accounting of the fused AAD launch (igemm_kernel, AAD mode) for the decoder block shapes: where the first
epilogue warp of every CTA and the producer / issuer spend their cycles.  Needs the profiling build (make PROF=1)."""
import ctypes as C
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("GHOST_B200_LIB", os.path.join(_ROOT, "ghost_b200", "libghost_b200_prof.so"))
sys.path.insert(0, _ROOT)
sys.path.insert(0, os.path.join(_ROOT, "tests"))
import numpy as np
import torch
from ghost_b200 import _lib
from util import fptr

lib = _lib.load()
for nm in ("ghost_prof_aad_read", "ghost_prof_igemm_read"):
    getattr(lib, nm).argtypes = [C.c_void_p]
    getattr(lib, nm).restype = C.c_int
g = torch.Generator().manual_seed(0)
c_id = 512
shapes = [(64, 256, 64, 64), (64, 128, 128, 128), (64, 64, 256, 256), (64, 32, 512, 512), (64, 16, 1024, 1024)]
for (N, H, c_x, c_attr) in shapes:
    h = torch.randn(N, H, H, c_x, generator=g).cuda().bfloat16()
    za = torch.randn(N, H, H, c_attr, generator=g).cuda().bfloat16()
    out = torch.empty_like(h)
    r = lambda *s: np.ascontiguousarray((torch.randn(*s, generator=g) * 0.05).numpy().astype(np.float32))
    args = [fptr(r(N, c_id)), N, H, H, c_x, c_attr, c_id, fptr(r(c_x, c_attr)), fptr(r(c_x)), fptr(r(c_x, c_attr)),
            fptr(r(c_x)), fptr(r(c_x, c_id)), fptr(r(c_x)), fptr(r(c_x, c_id)), fptr(r(c_x)), fptr(r(c_x)), 0.1]
    for _ in range(2):
        rc = lib.ghost_test_aad(C.c_void_p(h.data_ptr()), C.c_void_p(za.data_ptr()), *args, C.c_void_p(out.data_ptr()),
                                _lib.BF16, C.c_void_p(torch.cuda.current_stream().cuda_stream))
        _lib.check(rc, test_api=True)
    a = np.zeros((160, 8), dtype=np.uint64)
    b = np.zeros((160, 12), dtype=np.uint64)
    assert lib.ghost_prof_aad_read(C.c_void_p(a.ctypes.data)) == 0
    assert lib.ghost_prof_igemm_read(C.c_void_p(b.ctypes.data)) == 0
    cfg = b[159].astype(np.int64)
    grid = int(cfg[7])
    am = np.median(a[:grid].astype(np.float64), axis=0)
    bm = np.median(b[:grid].astype(np.float64), axis=0)
    by = N * H * H * (2 * c_x + c_attr) * 2
    print("AAD N=%d %dx%d c_x=%d c_attr=%d | BN=%d stages=%d num_kb=%d bres=%d grid=%d | CTA %.1f us by globaltimer = %.0f GB/s"
          % (N, H, H, c_x, c_attr, cfg[0], cfg[1], cfg[3], cfg[6], grid, bm[11] / 1e3, by / bm[11]))
    print("   epilogue warp 0: %d pieces, %.0f cyc each: wait h %.0f, pair barrier %.0f, wait accumulator %.0f, "
          "mask+blend %.0f, store+slot release %.0f" % (am[6], am[5] / max(am[6], 1), am[0] / max(am[6], 1),
                                                         am[1] / max(am[6], 1), am[2] / max(am[6], 1),
                                                         am[3] / max(am[6], 1), am[4] / max(am[6], 1)))
    print("   issuer: total %.0f, wait full %.0f%%, wait acc-empty %.0f%% | A producer: total %.0f, wait empty %.0f%%, "
          "wait h slot %.0f%%" % (bm[2], 100 * bm[0] / bm[2], 100 * bm[1] / bm[2], bm[5], 100 * bm[3] / max(bm[5], 1),
                                  100 * bm[4] / max(bm[5], 1)))
    del h, za, out
