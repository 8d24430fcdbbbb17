"""Three representative igemm_kernel launches for an `ncu --set full -k regex:igemm_kernel -c 3` capture:
   1. AADBlk8 fused AAD layer   (B=32, 64ch @256x256, c_attr=64)        - HBM-bound class
   2. AADBlk8 3x3 conv 64->64   (B=32, 256x256)                         - small-N conv
   3. AADBlk5 3x3 conv 512->512 (B=64, 32x32)                           - large-N conv (tensor-bound class)"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
from ghost_b200 import _lib
from util import fptr, run_conv

lib = _lib.load()
g = torch.Generator().manual_seed(0)
N, H, c_x, c_attr, c_id = 32, 256, 64, 64, 512
h = torch.randn(N, H, H, c_x, generator=g).cuda().bfloat16()
za = torch.randn(N, H, H, c_attr, generator=g).cuda().bfloat16()
out = torch.empty_like(h)
r = lambda *s: np.ascontiguousarray((torch.randn(*s, generator=g) * 0.05).numpy().astype(np.float32))
zid = r(N, c_id)
rc = lib.ghost_test_aad(C.c_void_p(h.data_ptr()), C.c_void_p(za.data_ptr()), fptr(zid), N, H, H, c_x, c_attr, c_id,
                        fptr(r(c_x, c_attr)), fptr(r(c_x)), fptr(r(c_x, c_attr)), fptr(r(c_x)), fptr(r(c_x, c_id)),
                        fptr(r(c_x)), fptr(r(c_x, c_id)), fptr(r(c_x)), fptr(r(c_x)), 0.1, C.c_void_p(out.data_ptr()),
                        _lib.BF16, C.c_void_p(torch.cuda.current_stream().cuda_stream))
_lib.check(rc, test_api=True)
x = torch.randn(N, 64, 256, 256, generator=g).cuda()
w = torch.randn(64, 64, 3, 3, generator=g) * 0.05
run_conv(x, w, None, 0, 3, 1, 1, 0, 0.0, _lib.BF16)
del x
x = torch.randn(64, 512, 32, 32, generator=g).cuda()
w = torch.randn(512, 512, 3, 3, generator=g) * 0.02
run_conv(x, w, None, 0, 3, 1, 1, 0, 0.0, _lib.BF16)
torch.cuda.synchronize()
print("ok")
