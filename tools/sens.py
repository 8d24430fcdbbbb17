import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from ghost_b200 import synth, weights as W
from oracle import aei_ref

def rnd(t, dt):
    return t.to(dt).float() if dt is not None else t

def emu(sd, Xt, z, bb, nb, wdt, adt, zdt, attrdt, blocks):
    """round weights/acts only inside generator blocks listed in `blocks`"""
    sd = {k: v.float() for k, v in sd.items() if v.dtype.is_floating_point}
    attrs = [rnd(a, attrdt) for a in aei_ref.encode(sd, Xt, bb)]
    zz = rnd(z, zdt)
    def aad(p, h, za, on):
        w_ = (lambda t: rnd(t, wdt)) if on else (lambda t: t)
        a_ = (lambda t: rnd(t, adt)) if on else (lambda t: t)
        hn = F.instance_norm(h, eps=1e-5)
        c_x = h.shape[1]
        ga = F.conv2d(za, w_(sd[p+".conv1.weight"]), sd[p+".conv1.bias"])
        ba = F.conv2d(za, w_(sd[p+".conv2.weight"]), sd[p+".conv2.bias"])
        gi = F.linear(zz, sd[p+".fc1.weight"], sd[p+".fc1.bias"]).reshape(-1, c_x, 1, 1)
        bi = F.linear(zz, sd[p+".fc2.weight"], sd[p+".fc2.bias"]).reshape(-1, c_x, 1, 1)
        M = torch.sigmoid(F.conv2d(hn, sd[p+".conv_h.weight"], sd[p+".conv_h.bias"]))
        return a_(F.relu((1-M)*(ga*hn+ba) + M*(gi*hn+bi)))
    m = F.conv_transpose2d(zz.reshape(zz.shape[0], -1, 1, 1), sd["generator.up1.weight"], sd["generator.up1.bias"])
    for k in range(1, 9):
        on = k in blocks
        w_ = (lambda t: rnd(t, wdt)) if on else (lambda t: t)
        a_ = (lambda t: rnd(t, adt)) if on else (lambda t: t)
        p = "generator.AADBlk%d" % k
        m = a_(m)
        x = m
        for i in range(nb):
            x = aad("%s.add_blocks.%d" % (p, 3*i), x, attrs[k-1], on)
            x = F.conv2d(x, w_(sd["%s.add_blocks.%d.weight" % (p, 3*i+2)]), None, padding=1)
            if i < nb-1: x = a_(x)
        if (p+".last_add_block.0.conv1.weight") in sd:
            h = aad(p+".last_add_block.0", m, attrs[k-1], on)
            h = a_(F.conv2d(h, w_(sd[p+".last_add_block.2.weight"]), None, padding=1))
        else:
            h = m
        y = x + h
        if k < 8:
            y = a_(y)
            m = F.interpolate(y, scale_factor=2, mode="bilinear", align_corners=True)
    return torch.tanh(y)

bb, nb = sys.argv[1], int(sys.argv[2])
pert = sys.argv[3] == "1"
sd = W.init_state_dict(W.aei_spec(bb, nb, 512), seed=0)
if pert: W.perturb_(sd, seed=1)
Xt, z = synth.crops_pm1(1, seed=7), synth.z_ids(1, seed=7)
bf, hf = torch.bfloat16, torch.float16
with torch.no_grad():
    Yr, _ = aei_ref.aei_forward(sd, Xt, z, bb, nb)
    def rep(name, **kw):
        args = dict(wdt=None, adt=None, zdt=None, attrdt=None, blocks=())
        args.update(kw)
        d = (emu(sd, Xt, z, bb, nb, **args) - Yr).abs()
        print("%-34s max %.3e mean %.3e" % (name, d.max(), d.mean()), flush=True)
    for k in range(1, 9):
        rep("bf16 weights blk%d" % k, wdt=bf, blocks=(k,))
    for k in range(1, 9):
        rep("bf16 acts blk%d" % k, adt=bf, blocks=(k,))
    rep("bf16 z", zdt=bf)
    rep("fp16 z", zdt=hf)
    rep("bf16 attr", attrdt=bf)
    rep("fp16 attr", attrdt=hf)
    rep("fp16 all (w,a,z,attr)", wdt=hf, adt=hf, zdt=hf, attrdt=hf, blocks=range(1, 9))
    rep("bf16 all", wdt=bf, adt=bf, zdt=bf, attrdt=bf, blocks=range(1, 9))
    rep("bf16 blk4-8 only (w,a,attr)", wdt=bf, adt=bf, attrdt=bf, blocks=range(4, 9))
    rep("bf16 blk5-8 only (w,a,attr)", wdt=bf, adt=bf, attrdt=bf, blocks=range(5, 9))
    rep("bf16 blk6-8 only (w,a,attr)", wdt=bf, adt=bf, attrdt=bf, blocks=range(6, 9))
