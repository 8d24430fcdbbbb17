"""Graph-replay time of one ArcFace and one AEI_Net forward (bf16, B=64 by default), CUDA events over 20 replays.
For A/B runs of the environment switches of the lowering (GHOST_PDL, GHOST_STG, GHOST_AWIN, ...)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ghost_b200 as gb

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = 20


def timed(fn):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    best, tot = 1e9, 0.0
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = min(best, ms)
        tot += ms
    return best, tot / reps


arc = gb.iresnet100().cuda().bfloat16().eval()
arc.enable_graph(True)
x = torch.randn(B, 3, 112, 112, device="cuda")
G = gb.AEI_Net("unet", num_blocks=2, c_id=512).cuda().bfloat16().eval()
G.enable_graph(True)
Xt = torch.randn(B, 3, 256, 256, device="cuda")
z = torch.nn.functional.normalize(torch.randn(B, 512, device="cuda"))
a = timed(lambda: arc(x))
g = timed(lambda: G(Xt, z))
print("switches: %s" % " ".join("%s=%s" % (k, v) for k, v in sorted(os.environ.items()) if k.startswith("GHOST_")))
print("iresnet100 B=%d: best %.3f ms, mean %.3f ms | AEI_Net unet/2: best %.3f ms, mean %.3f ms | sum of means %.3f ms"
      % (B, a[0], a[1], g[0], g[1], a[1] + g[1]))
