"""iresnet100 B=64 fp16, CUDA-graph replay: ms per forward and the embedding bits, for A/B runs of a lowering switch
(GHOST_CHAIN=0/1, ...).  usage: python tools/arc_chain_ab.py out.npy [B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import ghost_b200 as gb

B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
torch.manual_seed(3)
arc = gb.iresnet100().cuda().half().eval()
arc.enable_graph(True)
x = torch.randn(B, 3, 112, 112, device="cuda")
for _ in range(5):
    z = arc(x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for rep in range(5):
    e0.record()
    for _ in range(20):
        z = arc(x)
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / 20)
print("GHOST_CHAIN=%s B=%d: %.4f ms per forward (best of 5 x 20), |z| mean %.4f" % (
    os.environ.get("GHOST_CHAIN", "default"), B, best, float(z.float().norm(dim=1).mean())))
np.save(sys.argv[1], z.float().cpu().numpy())
