"""One igemm conv per (shape, N-tile hint, one-CTA / CTA-pair kernel) through the test API, timed by CUDA events
inside ghost_test_conv (GHOST_TEST_CONV_REPS); the pair kernel's output is compared with the one-CTA kernel's."""
import ctypes as C
import os
import sys

os.environ["GHOST_TEST_CONV_REPS"] = "20"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
from ghost_b200 import _lib
from util import fptr, nhwc

lib = _lib.load()
g = torch.Generator().manual_seed(0)
# (N, Cin, H, Cout, k, stride): iresnet layers 2-4, AADBlk4-6 convs, encoder convs
shapes = [(64, 256, 14, 256, 3, 1), (64, 128, 28, 128, 3, 1), (64, 512, 7, 512, 3, 1), (64, 64, 56, 64, 3, 1),
          (64, 512, 32, 512, 3, 1), (64, 1024, 16, 512, 3, 1), (64, 512, 32, 256, 3, 1), (64, 256, 64, 256, 3, 1),
          (64, 128, 64, 256, 4, 2), (64, 1024, 4, 1024, 3, 1), (64, 128, 28, 256, 3, 2)]
for (N, Cin, H, Cout, k, stride) in shapes:
    x = nhwc(torch.randn(N, Cin, H, H, generator=g).cuda(), torch.bfloat16)
    w = np.ascontiguousarray((torch.randn(Cout, Cin, k, k, generator=g) * 0.05).numpy().astype(np.float32))
    Ho = (H + 2 - k) // stride + 1
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    fl = 2.0 * N * Ho * Ho * Cout * Cin * k * k
    ref = None
    for force, name in ((2, "one-CTA"), (1, "pair   ")):
        for bn in (0, 64, 128, 256):
            if bn > Cout:
                continue
            y = torch.zeros((N, Ho, Ho, Cout), dtype=torch.bfloat16, device="cuda")
            rc = lib.ghost_test_conv(C.c_void_p(x.data_ptr()), N, H, H, Cin, fptr(w), None, Cout, 0, k, stride, 1, 0, 0.0,
                                     C.c_void_p(y.data_ptr()), _lib.BF16, bn | (force << 10), st)
            if rc != 0:
                print("N=%d Cin=%d H=%d Cout=%d k%d s%d %s bn=%3d: ERROR %s" % (N, Cin, H, Cout, k, stride, name, bn,
                                                                               lib.ghost_test_last_error()))
                continue
            ms = lib.ghost_test_last_ms()
            if ref is None:
                ref = y.clone()
            same = torch.equal(y, ref)
            print("N=%d Cin=%4d H=%2d Cout=%4d k%d s%d %s bn=%3d: %7.2f us %7.1f TFLOP/s  %s" % (
                N, Cin, H, Cout, k, stride, name, bn, ms * 1e3, fl / ms / 1e9, "==" if same else
                "DIFF max %.3e" % (y.float() - ref.float()).abs().max().item()))
