"""Per-op CUDA-event profile of one super-resolution forward (LIPSPADEGenerator, bf16) at batch B (GPU box) plus
graph-replayed throughput.  python tools/sr_profile.py [B=20] [reps=5]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ghost_b200
from ghost_b200 import synth

B = int(sys.argv[1]) if len(sys.argv) > 1 else 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
CLS = ["conv_tc", "aad_tc", "conv_simt", "stats", "upsample", "misc"]
G = ghost_b200.LIPSPADEGenerator().bfloat16().cuda()
crops = torch.from_numpy(synth.crops_u8(B)).cuda()
G.enhance_u8(crops)
torch.cuda.synchronize()
prof = G.profile(B, reps=reps)
tot = sum(o["ms"] for o in prof)
print("== LIPSPADEGenerator bf16  B=%d  %d ops  %.3f ms (un-graphed, per-op events) = %.1f crops/s" % (
    B, len(prof), tot, B / tot * 1e3))
for c in range(6):
    sel = [o for o in prof if o["cls"] == c]
    if not sel:
        continue
    ms = sum(o["ms"] for o in sel)
    fl = sum(o["flops"] for o in sel)
    by = sum(o["bytes"] for o in sel)
    print("   class %-9s n=%3d  %8.3f ms (%4.1f%%)  %8.1f TFLOP/s  %8.1f GB/s" % (
        CLS[c], len(sel), ms, 100 * ms / tot, fl / ms / 1e9 if ms else 0, by / ms / 1e6 if ms else 0))
groups = {}
for o in prof:
    k = o["name"].split(".")[-1] if o["name"].startswith(("head", "G_mid", "ups")) else o["name"].split(".")[0]
    g = groups.setdefault(k, [0.0, 0.0, 0.0, 0])
    g[0] += o["ms"]; g[1] += o["flops"]; g[2] += o["bytes"]; g[3] += 1
print("   by role:")
for k, g in sorted(groups.items(), key=lambda kv: -kv[1][0]):
    print("   %-14s n=%3d %8.3f ms (%4.1f%%) %8.1f TFLOP/s %8.1f GB/s" % (k, g[3], g[0], 100 * g[0] / tot,
                                                                           g[1] / g[0] / 1e9, g[2] / g[0] / 1e6))
print("   %-28s %-9s %9s %10s %10s" % ("op", "class", "ms", "TFLOP/s", "GB/s"))
for o in prof:
    print("   %-28s %-9s %9.4f %10.1f %10.1f" % (o["name"], CLS[o["cls"]], o["ms"],
                                                 o["flops"] / o["ms"] / 1e9 if o["ms"] else 0,
                                                 o["bytes"] / o["ms"] / 1e6 if o["ms"] else 0))
G.enable_graph(True)
out = torch.empty_like(crops)
for _ in range(3):
    G.enhance_u8(crops, out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    G.enhance_u8(crops, out)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
n, fl = G.stats(B)
print("graph replay: %.3f ms / batch of %d = %.1f crops/s, %.1f TFLOP/s over %d launches" % (ms, B, B / ms * 1e3,
                                                                                          fl / ms / 1e9, n))
