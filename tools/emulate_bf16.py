"""Scratch: emulate the bf16 storage points of the CUDA path in torch CPU to separate precision from bugs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from ghost_b200 import synth, weights as W
from oracle import aei_ref

def r(t, on=True):
    return t.to(torch.bfloat16).float() if on else t

def emu(sd, Xt, z, bb, nb, round_act=True, round_w=True, round_m=True, round_attr=True, round_z=True):
    sd = {k: v.float() for k, v in sd.items() if v.dtype.is_floating_point}
    # encoder in fp32 then round attrs (isolates the generator)
    attrs = aei_ref.encode(sd, Xt, bb)
    attrs = [r(a, round_attr) for a in attrs]
    zz = r(z, round_z)
    def aad(p, h, za):
        hn = F.instance_norm(h, eps=1e-5)
        c_x = h.shape[1]
        ga = F.conv2d(za, r(sd[p+".conv1.weight"], round_w), sd[p+".conv1.bias"])
        ba = F.conv2d(za, r(sd[p+".conv2.weight"], round_w), sd[p+".conv2.bias"])
        gi = F.linear(zz, r(sd[p+".fc1.weight"], round_w), sd[p+".fc1.bias"]).reshape(-1, c_x, 1, 1)
        bi = F.linear(zz, r(sd[p+".fc2.weight"], round_w), sd[p+".fc2.bias"]).reshape(-1, c_x, 1, 1)
        M = torch.sigmoid(F.conv2d(hn, sd[p+".conv_h.weight"], sd[p+".conv_h.bias"]))
        return r(F.relu((1-M)*(ga*hn+ba) + M*(gi*hn+bi)), round_act)
    m = F.conv_transpose2d(zz.reshape(zz.shape[0], -1, 1, 1), r(sd["generator.up1.weight"], round_w), sd["generator.up1.bias"])
    m = r(m, round_m)
    for k in range(1, 9):
        p = "generator.AADBlk%d" % k
        x = m
        for i in range(nb):
            x = aad("%s.add_blocks.%d" % (p, 3*i), x, attrs[k-1])
            x = F.conv2d(x, r(sd["%s.add_blocks.%d.weight" % (p, 3*i+2)], round_w), None, padding=1)
            if i < nb-1: x = r(x, round_act)
        if (p+".last_add_block.0.conv1.weight") in sd:
            h = aad(p+".last_add_block.0", m, attrs[k-1])
            h = r(F.conv2d(h, r(sd[p+".last_add_block.2.weight"], round_w), None, padding=1), round_act)
        else:
            h = m
        y = x + h
        if k < 8:
            y = r(y, round_m)
            m = r(F.interpolate(y, scale_factor=2, mode="bilinear", align_corners=True), round_m)
    return torch.tanh(y)

bb, nb = "unet", 2
sd = W.perturb_(W.init_state_dict(W.aei_spec(bb, nb, 512), seed=0), seed=1)
Xt, z = synth.crops_pm1(1, seed=7), synth.z_ids(1, seed=7)
with torch.no_grad():
    Yr, _ = aei_ref.aei_forward(sd, Xt, z, bb, nb)
    for name, kw in [("all", {}), ("weights only", dict(round_act=False, round_m=False, round_attr=False, round_z=False)),
                     ("acts only", dict(round_w=False, round_m=False, round_attr=False, round_z=False)),
                     ("m only", dict(round_w=False, round_act=False, round_attr=False, round_z=False)),
                     ("attr only", dict(round_w=False, round_act=False, round_m=False, round_z=False)),
                     ("z only", dict(round_w=False, round_act=False, round_m=False, round_attr=False)),
                     ("none", dict(round_w=False, round_act=False, round_m=False, round_attr=False, round_z=False))]:
        Y = emu(sd, Xt, z, bb, nb, **kw)
        d = (Y - Yr).abs()
        print("%-14s max %.3e mean %.3e" % (name, d.max(), d.mean()))
