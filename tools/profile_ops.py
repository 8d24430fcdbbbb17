"""Per-op CUDA-event profile of one generator / ArcFace forward at batch B (GPU box).  Writes a table."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ghost_b200
from ghost_b200 import synth

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
bb = sys.argv[2] if len(sys.argv) > 2 else "unet"
nb = int(sys.argv[3]) if len(sys.argv) > 3 else 2
CLS = ["conv_tc", "aad_tc", "conv_simt", "stats", "upsample", "misc"]
G = ghost_b200.AEI_Net(bb, nb, 512).eval().half().cuda()
arc = ghost_b200.iresnet100().eval().half().cuda()
crops = torch.from_numpy(synth.crops_u8(B)).cuda()
z = arc.embed_hwc(torch.from_numpy(synth.sources_u8(B)).cuda(), out_dtype=torch.float16)
G.swap_u8(crops, z)
torch.cuda.synchronize()
for name, net in (("AEI_Net %s/%d" % (bb, nb), G), ("iresnet100", arc)):
    prof = net.profile(B, reps=5)
    tot = sum(o["ms"] for o in prof)
    print("== %s  B=%d  %d ops  %.3f ms (un-graphed, per-op events)" % (name, B, len(prof), tot))
    for c in range(6):
        sel = [o for o in prof if o["cls"] == c]
        if not sel:
            continue
        ms = sum(o["ms"] for o in sel)
        fl = sum(o["flops"] for o in sel)
        by = sum(o["bytes"] for o in sel)
        print("   class %-9s n=%3d  %8.3f ms (%4.1f%%)  %8.1f TFLOP/s  %8.1f GB/s" % (
            CLS[c], len(sel), ms, 100 * ms / tot, fl / ms / 1e9 if ms else 0, by / ms / 1e6 if ms else 0))
    print("   %-28s %-9s %9s %10s %10s" % ("op", "class", "ms", "TFLOP/s", "GB/s"))
    for o in prof:
        print("   %-28s %-9s %9.4f %10.1f %10.1f" % (o["name"], CLS[o["cls"]], o["ms"],
                                                     o["flops"] / o["ms"] / 1e9 if o["ms"] else 0,
                                                     o["bytes"] / o["ms"] / 1e6 if o["ms"] else 0))
