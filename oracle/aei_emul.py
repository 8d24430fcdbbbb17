"""Storage-rounding emulation of the CUDA data path - TEST INFRASTRUCTURE ONLY.

Same arithmetic as oracle/aei_ref.py (the reference's AEI_Net.forward, network/AEI_Net.py:72-156 and
network/AADLayer.py:20-80) evaluated in fp32 on the CPU, but every tensor that the bf16 mode of the CUDA
path stores in bf16 is rounded to `dt` at the same point: BN-folded conv / deconv weights, the 1x1 gamma/beta
weights, the 3x3 weights, the z_id FC weights and z_id itself, and every activation written to HBM
(encoder features, AAD+ReLU outputs, conv outputs, block outputs, upsampled maps).  Statistics, the
normalise / modulate / mask / blend math and all accumulations stay in fp32, as in the kernels.

It separates "the kernels compute something else" (a bug) from "bf16 storage moves the result" (precision):
the CUDA path must agree with this emulation far more tightly than either agrees with the fp32 reference.
"""
import torch
import torch.nn.functional as F


def _r(t, dt):
    return t.to(dt).float() if dt is not None else t


def _fold(sd, w, p, dt, transposed=False):
    s = sd[p + ".weight"] / torch.sqrt(sd[p + ".running_var"] + 1e-5)
    shift = sd[p + ".bias"] - sd[p + ".running_mean"] * s
    w = w * (s.reshape(1, -1, 1, 1) if transposed else s.reshape(-1, 1, 1, 1))
    return w, shift


def _up2(x):
    return F.interpolate(x, scale_factor=2, mode="bilinear", align_corners=True)


def aei_forward_emul(sd, Xt, z_id, backbone, num_blocks, dt=torch.bfloat16, trace=None):
    with torch.no_grad():
        sd = {k: v.float() for k, v in sd.items() if v.dtype.is_floating_point}
        x = _r(Xt.float(), dt)
        feats = []
        for i in range(1, 8):
            w, b = _fold(sd, sd["encoder.conv%d.0.weight" % i], "encoder.conv%d.1" % i, dt)
            if i > 1:                      # conv1 (Cin = 3) runs on CUDA cores with fp32 weights
                w = _r(w, dt)
            x = _r(F.leaky_relu(F.conv2d(x, w, b, stride=2, padding=1), 0.1), dt)
            feats.append(x)
        attrs = [feats[-1]]
        zc = feats[-1]
        for j in range(1, 7):
            w, b = _fold(sd, sd["encoder.deconv%d.deconv.weight" % j], "encoder.deconv%d.bn" % j, dt, transposed=True)
            y = F.leaky_relu(F.conv_transpose2d(zc, _r(w, dt), b, stride=2, padding=1), 0.1)
            skip = feats[6 - j]
            zc = _r(y + skip, dt) if backbone == "linknet" else torch.cat((_r(y, dt), skip), dim=1)
            attrs.append(zc)
        attrs.append(_r(_up2(attrs[-1]), dt))

        z = _r(z_id.float().reshape(z_id.shape[0], -1), dt)

        def aad(p, h, za):
            hn = F.instance_norm(h, eps=1e-5)
            c_x = h.shape[1]
            ga = F.conv2d(za, _r(sd[p + ".conv1.weight"], dt), sd[p + ".conv1.bias"])
            ba = F.conv2d(za, _r(sd[p + ".conv2.weight"], dt), sd[p + ".conv2.bias"])
            gi = F.linear(z, _r(sd[p + ".fc1.weight"], dt), sd[p + ".fc1.bias"]).reshape(-1, c_x, 1, 1)
            bi = F.linear(z, _r(sd[p + ".fc2.weight"], dt), sd[p + ".fc2.bias"]).reshape(-1, c_x, 1, 1)
            M = torch.sigmoid(F.conv2d(hn, sd[p + ".conv_h.weight"], sd[p + ".conv_h.bias"]))
            return _r(F.relu((1.0 - M) * (ga * hn + ba) + M * (gi * hn + bi)), dt)

        m = F.conv_transpose2d(z.reshape(z.shape[0], -1, 1, 1), _r(sd["generator.up1.weight"], dt),
                               sd["generator.up1.bias"])
        m = _r(m, dt)
        for k in range(1, 9):
            p = "generator.AADBlk%d" % k
            if trace is not None:
                trace["m%d" % k] = m
            x = m
            for i in range(num_blocks):
                a = aad("%s.add_blocks.%d" % (p, 3 * i), x, attrs[k - 1])
                x = F.conv2d(a, _r(sd["%s.add_blocks.%d.weight" % (p, 3 * i + 2)], dt), None, padding=1)
                if i < num_blocks - 1:
                    x = _r(x, dt)
            if (p + ".last_add_block.0.conv1.weight") in sd:
                h = aad(p + ".last_add_block.0", m, attrs[k - 1])
                h = _r(F.conv2d(h, _r(sd[p + ".last_add_block.2.weight"], dt), None, padding=1), dt)
            else:
                h = m
            y = x + h
            if k < 8:
                y = _r(y, dt)
                m = _r(_up2(y), dt)
            if trace is not None:
                trace["a%d" % k] = a
                trace["y%d" % k] = torch.tanh(y) if k == 8 else y
        return torch.tanh(y), tuple(attrs)
