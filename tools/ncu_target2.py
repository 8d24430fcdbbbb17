"""Four more igemm_kernel launches for an `ncu --set full -k regex:igemm_kernel -c 4` capture (ArcFace / decoder shapes):
   1. iresnet100 layer2 3x3 conv 128->128 @28x28, B=64      2. iresnet100 layer3 3x3 conv 256->256 @14x14, B=64
   3. AADBlk8 3x3 conv 64->64 @256x256, B=32 (row-window)   4. AADBlk7 3x3 conv 128->64 @128x128, B=64 (row-window)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from ghost_b200 import _lib
from util import run_conv

g = torch.Generator().manual_seed(0)
for (N, Cin, H, Cout) in ((64, 128, 28, 128), (64, 256, 14, 256), (32, 64, 256, 64), (64, 128, 128, 64)):
    x = torch.randn(N, Cin, H, H, generator=g).cuda()
    w = torch.randn(Cout, Cin, 3, 3, generator=g) * 0.05
    run_conv(x, w, None, 0, 3, 1, 1, 0, 0.0, _lib.BF16)
    del x
torch.cuda.synchronize()
print("ok")
