"""Two un-graphed super-resolution forwards at B=20 (bf16) for ncu captures (the second one is the warm one).
igemm_kernel launches per forward: 54, in op order - 0 seg.actv256 (thin-K conv through the TMA-store epilogue),
47 ups.3.norm_s.gemm_modulate (fused SPADE, 256x256), 51 ups.3.conv_0 (3x3 128->64 row-window), 53 out.to_rgb."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ghost_b200
from ghost_b200 import synth

B = int(sys.argv[1]) if len(sys.argv) > 1 else 20
G = ghost_b200.LIPSPADEGenerator().bfloat16().cuda()
crops = torch.from_numpy(synth.crops_u8(B)).cuda()
out = torch.empty_like(crops)
for _ in range(2):
    G.enhance_u8(crops, out)
    torch.cuda.synchronize()
print("ok", G.stats(B))
