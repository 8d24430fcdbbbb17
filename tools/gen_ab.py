"""AEI_Net unet/2 B=64 fp16, CUDA-graph replay: ms per forward and an output checksum, for A/B runs of a lowering
switch or a kernel change.  usage: python tools/gen_ab.py [out.npy] [B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import ghost_b200 as gb

B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
torch.manual_seed(3)
G = gb.AEI_Net("unet", num_blocks=2, c_id=512).cuda().half().eval()
G.enable_graph(True)
Xt = torch.randn(B, 3, 256, 256, device="cuda").clamp_(-1, 1)
z = torch.nn.functional.normalize(torch.randn(B, 512, device="cuda"))
for _ in range(4):
    y = G(Xt, z)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for rep in range(5):
    e0.record()
    for _ in range(10):
        y = G(Xt, z)
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / 10)
y0 = y[0] if isinstance(y, (tuple, list)) else y
print("AEI_Net unet/2 B=%d: %.4f ms per forward (best of 5 x 10), mean |Y| %.6f" % (B, best, float(y0.float().abs().mean())))
if len(sys.argv) > 1:
    np.save(sys.argv[1], y0[:4].float().cpu().numpy())
