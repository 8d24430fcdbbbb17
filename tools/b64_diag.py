"""Diagnostic: where does the fp16 path differ from the fp32 oracle at the bench's B = 64 inputs, and how does the
reference's own G.half() (torch eager on the GPU) do on the same inputs?"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, ghost_b200
from ghost_b200 import weights as W
from oracle import aei_ref, iresnet_ref

B = int(os.environ.get("B", 64))
d = bench.make_inputs(0, "default")
sd = W.init_state_dict(W.aei_spec("unet", 2, 512), seed=0)
arc_sd = W.unit_embedding_(W.init_state_dict(W.iresnet100_spec(), seed=0))
torch.set_num_threads(os.cpu_count() or 1)
z = iresnet_ref.embed_hwc_ref(arc_sd, d["sources"][:B]).half().float()
Xt = aei_ref.transform_target_ref(d["crops"][:B], half=True).float()
Yr, _ = aei_ref.aei_forward(sd, Xt, z, "unet", 2)

def rep(name, Y):
    dd = (Y.float().cpu() - Yr).abs()
    per = dd.reshape(B, -1)
    print("%-28s max %.3e mean %.3e  frac>1e-2 %.2e  frac>3e-2 %.2e  worst crop %d (max %.3e)" % (
        name, dd.max().item(), dd.mean().item(), (dd > 1e-2).float().mean().item(), (dd > 3e-2).float().mean().item(),
        per.max(1).values.argmax().item(), per.max(1).values.max().item()))
    return dd

G = ghost_b200.AEI_Net("unet", 2, 512, seed=0).eval().half().cuda()
G.return_attr = False
Y16, _ = G(Xt.cuda().half(), z.cuda().half())
d16 = rep("ours fp16 B=%d" % B, Y16)
Gb = ghost_b200.AEI_Net("unet", 2, 512, seed=0).eval().bfloat16().cuda(); Gb.return_attr = False
Yb, _ = Gb(Xt.cuda().bfloat16(), z.cuda().bfloat16())
rep("ours bf16", Yb)
Gf = ghost_b200.AEI_Net("unet", 2, 512, seed=0).eval().float().cuda(); Gf.return_attr = False
nf = min(B, 8)
Yf, _ = Gf(Xt[:nf].cuda(), z[:nf].cuda())
dd = (Yf.cpu() - Yr[:nf]).abs(); print("ours fp32 first %d: max %.3e mean %.3e" % (nf, dd.max().item(), dd.mean().item()))
with torch.no_grad():
    sdh = {k: v.cuda().half() for k, v in sd.items() if v.dtype.is_floating_point}
    attrs = aei_ref.encode(sdh, Xt.cuda().half(), "unet")
    Yh = aei_ref.generate(sdh, attrs, z.cuda().half(), 2)
dh = rep("torch eager G.half() cuda", Yh)
# the worst crop alone
w = d16.reshape(B, -1).max(1).values.argmax().item()
Y1, _ = G(Xt[w:w+1].cuda().half(), z[w:w+1].cuda().half())
print("worst crop %d alone vs in batch: max %.3e" % (w, (Y1.float() - Y16[w:w+1].float()).abs().max().item()))
idx = d16[w].reshape(-1).argmax().item(); c, r = divmod(idx, 65536); y, x = divmod(r, 256)
print("worst pixel: c %d y %d x %d: ref %.4f ours %.4f eager-half %.4f" % (c, y, x, Yr[w, c, y, x], Y16[w, c, y, x].float(), Yh[w, c, y, x].float()))
# are outliers co-located between ours and torch half?
m16, mh = (d16 > 1e-2), (dh > 1e-2)
print("pixels >1e-2: ours %d, eager-half %d, both %d" % (m16.sum().item(), mh.sum().item(), (m16 & mh).sum().item()))
