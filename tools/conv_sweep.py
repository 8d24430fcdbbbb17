"""One igemm conv launch per (shape, N-tile hint) through the test API, in a fixed order: run it under
`ncu --metrics gpu__time_duration.sum -k regex:igemm_kernel --csv` and match the durations to the printed list."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
from ghost_b200 import _lib
from util import fptr, nhwc

lib = _lib.load()
g = torch.Generator().manual_seed(0)
shapes = [(64, 256, 14, 256), (64, 128, 28, 128), (64, 64, 56, 64), (64, 512, 32, 512)]
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
        run()
        torch.cuda.synchronize()
        fl = 2.0 * N * H * H * Cout * Cin * 9
        print("launch: N=%d Cin=%d H=%d Cout=%d bn_hint=%3d  GFLOP=%.2f" % (N, Cin, H, Cout, bn, fl / 1e9))
