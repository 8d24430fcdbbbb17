"""How far is the device ArcFace z_id (fp16 / bf16 storage) from the fp32 oracle, and what does that distance alone do
to Y (fp32 generator on both sides)?"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, ghost_b200
from ghost_b200 import weights as W
from oracle import aei_ref, iresnet_ref
torch.set_num_threads(os.cpu_count() or 1)
n = 8
d = bench.make_inputs(0, "default")
arc_sd = W.unit_embedding_(W.init_state_dict(W.iresnet100_spec(), seed=0))
z = iresnet_ref.embed_hwc_ref(arc_sd, d["sources"][:n])
src = torch.from_numpy(d["sources"][:n]).cuda()
sd = W.init_state_dict(W.aei_spec("unet", 2, 512), seed=0)
Xt = aei_ref.transform_target_ref(d["crops"][:n], half=True).float()
Y0, _ = aei_ref.aei_forward(sd, Xt, z, "unet", 2)
for name, dt in (("fp32", torch.float32), ("fp16", torch.float16), ("bf16", torch.bfloat16)):
    arc = bench.make_arcface(dt)
    zc = arc.embed_hwc(src, out_dtype=torch.float32).cpu()
    cos = torch.nn.functional.cosine_similarity(zc, z, dim=1)
    rel = ((zc - z).norm(dim=1) / z.norm(dim=1))
    Y1, _ = aei_ref.aei_forward(sd, Xt, zc, "unet", 2)
    dy = (Y1 - Y0).abs()
    print("arcface %s: min cos %.7f  rel |dz|/|z| max %.3e  max|dz| %.3e  -> Y (fp32 generator): max %.3e mean %.3e" % (
        name, cos.min().item(), rel.max().item(), (zc - z).abs().max().item(), dy.max().item(), dy.mean().item()))
Y2, _ = aei_ref.aei_forward(sd, Xt, z.half().float(), "unet", 2)
dy = (Y2 - Y0).abs(); print("z.half() alone -> Y: max %.3e mean %.3e" % (dy.max().item(), dy.mean().item()))
