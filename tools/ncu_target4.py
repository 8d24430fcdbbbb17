"""Two AEI_Net(unet/2) forwards at B=64 for an ncu capture of tapn_kernel (the two 64->3 convs of AADBlk8);
capture with -k regex:tapn_kernel -s 2 -c 2 (the warm forward)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ghost_b200
from ghost_b200 import synth

B = 64
G = ghost_b200.AEI_Net("unet", 2, 512).eval().bfloat16().cuda()
crops = torch.from_numpy(synth.crops_u8(B)).cuda()
z = torch.nn.functional.normalize(torch.randn(B, 512, device="cuda")).bfloat16()
for _ in range(2):
    G.swap_u8(crops, z)
    torch.cuda.synchronize()
print("ok")
