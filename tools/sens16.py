"""Which 16-bit storage points move Y?  Reference arithmetic in fp32 on the CPU (oracle/aei_ref.py's functions) with
rounding to `dt` switched on per GROUP of storage points:

  enc_w   BN-folded encoder conv / deconv weights        attr   the eight attribute maps (z_attr1..8)
  z       z_id                                            zfc_w  up1 / fc1 / fc2 weights
  gb_w    1x1 gamma/beta conv weights                     c3_w   3x3 conv weights
  a       AADLayer + ReLU outputs (3x3 conv inputs)       x      conv outputs inside a block (next layer's h)
  h       shortcut conv output                            y      block outputs          m   upsampled block inputs

usage: python tools/sens16.py <backbone> <num_blocks> <perturbed 0|1> [fp16|bf16]
Prints, per group: error when ONLY that group is rounded, and when everything BUT that group is rounded."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from ghost_b200 import synth, weights as W
from oracle import aei_ref

GROUPS = ["enc_w", "attr", "z", "zfc_w", "gb_w", "c3_w", "a", "x", "h", "y", "m"]


def emu(sd, Xt, z, bb, nb, dt, on):
    def r(t, g):
        return t.to(dt).float() if g in on else t

    sd = {k: v.float() for k, v in sd.items() if v.dtype.is_floating_point}
    if "enc_w" in on:
        sd = dict(sd)
        for k in list(sd):
            if k.startswith("encoder.") and (k.endswith(".0.weight") or k.endswith("deconv.weight")) and "conv1.0" not in k:
                sd[k] = sd[k].to(dt).float()   # (BN scale is folded before rounding in the kernels; close enough here)
    attrs = [r(a, "attr") for a in aei_ref.encode(sd, Xt, bb)]
    zz = r(z, "z")

    def aad(p, hh, za):
        hn = F.instance_norm(hh, eps=1e-5)
        c_x = hh.shape[1]
        ga = F.conv2d(za, r(sd[p + ".conv1.weight"], "gb_w"), sd[p + ".conv1.bias"])
        ba = F.conv2d(za, r(sd[p + ".conv2.weight"], "gb_w"), sd[p + ".conv2.bias"])
        gi = F.linear(zz, r(sd[p + ".fc1.weight"], "zfc_w"), sd[p + ".fc1.bias"]).reshape(-1, c_x, 1, 1)
        bi = F.linear(zz, r(sd[p + ".fc2.weight"], "zfc_w"), sd[p + ".fc2.bias"]).reshape(-1, c_x, 1, 1)
        M = torch.sigmoid(F.conv2d(hn, sd[p + ".conv_h.weight"], sd[p + ".conv_h.bias"]))
        return r(F.relu((1 - M) * (ga * hn + ba) + M * (gi * hn + bi)), "a")

    m = F.conv_transpose2d(zz.reshape(zz.shape[0], -1, 1, 1), r(sd["generator.up1.weight"], "zfc_w"),
                           sd["generator.up1.bias"])
    m = r(m, "m")
    for k in range(1, 9):
        p = "generator.AADBlk%d" % k
        x = m
        for i in range(nb):
            x = aad("%s.add_blocks.%d" % (p, 3 * i), x, attrs[k - 1])
            x = F.conv2d(x, r(sd["%s.add_blocks.%d.weight" % (p, 3 * i + 2)], "c3_w"), None, padding=1)
            if i < nb - 1:
                x = r(x, "x")
        if (p + ".last_add_block.0.conv1.weight") in sd:
            h = aad(p + ".last_add_block.0", m, attrs[k - 1])
            h = r(F.conv2d(h, r(sd[p + ".last_add_block.2.weight"], "c3_w"), None, padding=1), "h")
        else:
            h = m
        y = x + h
        if k < 8:
            y = r(y, "y")
            m = r(F.interpolate(y, scale_factor=2, mode="bilinear", align_corners=True), "m")
    return torch.tanh(y)


def main():
    bb, nb, pert = sys.argv[1], int(sys.argv[2]), sys.argv[3] == "1"
    dt = torch.bfloat16 if len(sys.argv) > 4 and sys.argv[4] == "bf16" else torch.float16
    sd = W.init_state_dict(W.aei_spec(bb, nb, 512), seed=0)
    if pert:
        W.perturb_(sd, seed=1)
    Xt, z = synth.crops_pm1(1, seed=7), synth.z_ids(1, seed=7)
    with torch.no_grad():
        Yr, _ = aei_ref.aei_forward(sd, Xt, z, bb, nb)

        def rep(name, on):
            d = (emu(sd, Xt, z, bb, nb, dt, set(on)) - Yr).abs()
            print("%-28s max %.3e mean %.3e" % (name, d.max(), d.mean()), flush=True)

        rep("all", GROUPS)
        for g in GROUPS:
            rep("only " + g, [g])
        for g in GROUPS:
            rep("all but " + g, [x for x in GROUPS if x != g])
        for combo in sys.argv[5:]:
            off = combo.split(",")
            rep("all but " + combo, [x for x in GROUPS if x not in off])


if __name__ == "__main__":
    main()
