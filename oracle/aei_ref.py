"""CPU oracle for AEI_Net.forward - TEST INFRASTRUCTURE ONLY (never imported by ghost_b200/).

A functional, state_dict-driven restatement (plain torch fp32 on the CPU) of what the reference computes in
    network/AEI_Net.py:19-41   conv4x4 / deconv4x4 (Conv / ConvTranspose k4 s2 p1, no bias -> BN(eval) -> LeakyReLU 0.1)
    network/AEI_Net.py:72-95   MLAttrEncoder.forward (7 convs, 6 deconvs with cat [unet] / add [linknet], bilinear x2)
    network/AEI_Net.py:122-139 AADGenerator.forward (up1, 8 x [AAD_ResBlk, bilinear x2], tanh)
    network/AADLayer.py:20-38  AADLayer.forward
    network/AADLayer.py:40-80  AddBlocksSequential / AAD_ResBlk.forward
    network/AEI_Net.py:153-156 AEI_Net.forward

Pinning: the reference has no golden vectors for this path, so the oracle is pinned against the
reference's own modules run in the build container (oracle/gen_golden.py imports /root/reference, loads the
same state_dict into network.AEI_Net.AEI_Net and writes tests/golden/aei_*.npz); tests/test_oracle_aei.py
checks this restatement against those files everywhere, and against the live reference where it is mounted.
"""
import torch
import torch.nn.functional as F

_ENC = 7
_DEC = 6


def _bn_eval(x, sd, p):
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"], sd[p + ".bias"],
                        False, 0.0, 1e-5)


def _up2(x):
    return F.interpolate(x, scale_factor=2, mode="bilinear", align_corners=True)


def encode(sd, Xt, backbone):
    """MLAttrEncoder.forward (AEI_Net.py:72-95): returns (z_attr1, ..., z_attr8)."""
    feats = []
    x = Xt
    for i in range(1, _ENC + 1):
        x = F.conv2d(x, sd["encoder.conv%d.0.weight" % i], None, stride=2, padding=1)
        x = F.leaky_relu(_bn_eval(x, sd, "encoder.conv%d.1" % i), 0.1)
        feats.append(x)
    attrs = [feats[-1]]
    z = feats[-1]
    for j in range(1, _DEC + 1):
        y = F.conv_transpose2d(z, sd["encoder.deconv%d.deconv.weight" % j], None, stride=2, padding=1)
        y = F.leaky_relu(_bn_eval(y, sd, "encoder.deconv%d.bn" % j), 0.1)
        skip = feats[_ENC - 1 - j]
        z = y + skip if backbone == "linknet" else torch.cat((y, skip), dim=1)
        attrs.append(z)
    attrs.append(_up2(attrs[-1]))
    return tuple(attrs)


def aad_layer(sd, p, h_in, z_attr, z_id):
    """AADLayer.forward (AADLayer.py:20-38)."""
    h = F.instance_norm(h_in, eps=1e-5)
    c_x = h.shape[1]
    g_a = F.conv2d(z_attr, sd[p + ".conv1.weight"], sd[p + ".conv1.bias"])
    b_a = F.conv2d(z_attr, sd[p + ".conv2.weight"], sd[p + ".conv2.bias"])
    g_i = F.linear(z_id, sd[p + ".fc1.weight"], sd[p + ".fc1.bias"]).reshape(-1, c_x, 1, 1)
    b_i = F.linear(z_id, sd[p + ".fc2.weight"], sd[p + ".fc2.bias"]).reshape(-1, c_x, 1, 1)
    A = g_a * h + b_a
    I = g_i * h + b_i
    M = torch.sigmoid(F.conv2d(h, sd[p + ".conv_h.weight"], sd[p + ".conv_h.bias"]))
    return (1.0 - M) * A + M * I


def _aad_chain(sd, p, idxs, h, z_attr, z_id):
    x = h
    for i in idxs:
        x = F.relu(aad_layer(sd, "%s.%d" % (p, 3 * i), x, z_attr, z_id))
        x = F.conv2d(x, sd["%s.%d.weight" % (p, 3 * i + 2)], None, padding=1)
    return x


def aad_resblk(sd, p, h, z_attr, z_id, num_blocks):
    """AAD_ResBlk.forward (AADLayer.py:75-80)."""
    x = _aad_chain(sd, p + ".add_blocks", range(num_blocks), h, z_attr, z_id)
    if (p + ".last_add_block.0.conv1.weight") in sd:      # cin != cout
        h = _aad_chain(sd, p + ".last_add_block", range(1), h, z_attr, z_id)
    return x + h


def generate(sd, attrs, z_id, num_blocks, trace=None):
    """AADGenerator.forward (AEI_Net.py:122-139).  trace (optional dict) receives the block inputs m1..m8 and
    block outputs y1..y8 (y8 after tanh) for the block-level parity tests."""
    z = z_id.reshape(z_id.shape[0], -1)
    m = F.conv_transpose2d(z.reshape(z.shape[0], -1, 1, 1), sd["generator.up1.weight"], sd["generator.up1.bias"])
    for k in range(1, 9):
        if trace is not None:
            trace["m%d" % k] = m
        y = aad_resblk(sd, "generator.AADBlk%d" % k, m, attrs[k - 1], z, num_blocks)
        if trace is not None:
            trace["y%d" % k] = torch.tanh(y) if k == 8 else y
        if k < 8:
            m = _up2(y)
    return torch.tanh(y)


def aei_forward(sd, Xt, z_id, backbone, num_blocks, trace=None):
    """AEI_Net.forward (AEI_Net.py:153-156): returns (Y, attrs)."""
    with torch.no_grad():
        sd = {k: v.float() for k, v in sd.items() if v.dtype.is_floating_point}
        attrs = encode(sd, Xt.float(), backbone)
        return generate(sd, attrs, z_id.float(), num_blocks, trace), attrs


# ------------------------------------------------------------------ boundary conversions
def transform_target_ref(resized_frs, half=True):
    """utils/inference/core.py:13-26 on the CPU: uint8 BGR NHWC -> normalised RGB NCHW.
    half=True keeps the reference's rounding order (fp16 after /255, then (x-0.5)/0.5 in fp16)."""
    x = torch.from_numpy(resized_frs.copy())[:, :, :, [2, 1, 0]] / 255.
    if half:
        x = x.half()
    x = (x - 0.5) / 0.5
    return x.permute(0, 3, 1, 2)


def postprocess_ref(Y):
    """utils/inference/faceshifter_run.py:20-22 in fp32: Y [B,3,H,W] -> uint8 BGR [B,H,W,3] (truncating cast)."""
    y = (Y.float().permute(0, 2, 3, 1) * 0.5 + 0.5) * 255
    return y[:, :, :, [2, 1, 0]].type(torch.uint8).numpy()


def postprocess_ref_half(Y_half):
    """utils/inference/faceshifter_run.py:20-22 as the reference evaluates it with G.half() (inference.py:30): Y is
    an fp16 tensor, so (Y*0.5 + 0.5)*255 runs in fp16 with one rounding per operation, then the truncating uint8 cast.
    torch's CPU half kernels compute each elementwise op in fp32 and round the result to fp16 = the same roundings."""
    assert Y_half.dtype == torch.float16
    y = (Y_half.permute(0, 2, 3, 1) * 0.5 + 0.5) * 255
    return y[:, :, :, [2, 1, 0]].type(torch.uint8).numpy()
