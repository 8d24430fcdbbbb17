"""state_dict layouts (reference key names and shapes) and random initialisation.

AEI_Net keys follow network/AEI_Net.py:44-151 and network/AADLayer.py:5-74 of the reference
(`encoder.conv1.0.weight`, `generator.AADBlk4.add_blocks.3.fc1.bias`, ...); the init rules
follow weight_init (network/AEI_Net.py:8-16): Linear W ~ N(0, 0.001) and zero bias, Conv2d /
ConvTranspose2d W ~ xavier_normal, conv biases keep PyTorch's default U(+-1/sqrt(fan_in)),
BatchNorm gamma=1, beta=0, running mean 0 / var 1.

iresnet100 keys follow the public insightface arcface_torch backbone that the reference
downloads as arcface_model/iresnet.py (download_models.sh:2-3; the file is not in its tree).
"""
import math
from collections import OrderedDict

import torch

ENC_C = [3, 32, 64, 128, 256, 512, 1024, 1024]
DEC_OUT = [1024, 512, 256, 128, 64, 32]
DEC_IN = {"unet": [1024, 2048, 1024, 512, 256, 128], "linknet": [1024, 1024, 512, 256, 128, 64]}
# (cin, cout, c_attr) of AADBlk1..8
BLOCKS = {
    "unet": [(1024, 1024, 1024), (1024, 1024, 2048), (1024, 1024, 1024), (1024, 512, 512),
             (512, 256, 256), (256, 128, 128), (128, 64, 64), (64, 3, 64)],
    "linknet": [(1024, 1024, 1024), (1024, 1024, 1024), (1024, 1024, 512), (1024, 512, 256),
                (512, 256, 128), (256, 128, 64), (128, 64, 32), (64, 3, 32)],
}
IRESNET100_LAYERS = [3, 13, 30, 3]
IRESNET_PLANES = [64, 128, 256, 512]


def _bn(spec, prefix, c):
    spec[prefix + ".weight"] = ((c,), "ones")
    spec[prefix + ".bias"] = ((c,), "zeros")
    spec[prefix + ".running_mean"] = ((c,), "zeros")
    spec[prefix + ".running_var"] = ((c,), "ones")
    spec[prefix + ".num_batches_tracked"] = ((), "count")


def aei_spec(backbone, num_blocks, c_id):
    """OrderedDict key -> (shape, init kind) in the reference's state_dict order."""
    if backbone not in BLOCKS:
        raise NotImplementedError(
            "ghost_b200.AEI_Net supports backbone 'unet' and 'linknet' (got %r; the reference's 'resnet' "
            "attribute encoder, network/resnet.py, is outside the accelerated path)" % (backbone,))
    spec = OrderedDict()
    for i in range(1, 8):
        spec["encoder.conv%d.0.weight" % i] = ((ENC_C[i], ENC_C[i - 1], 4, 4), "xavier_conv")
        _bn(spec, "encoder.conv%d.1" % i, ENC_C[i])
    for i in range(1, 7):
        spec["encoder.deconv%d.deconv.weight" % i] = ((DEC_IN[backbone][i - 1], DEC_OUT[i - 1], 4, 4), "xavier_deconv")
        _bn(spec, "encoder.deconv%d.bn" % i, DEC_OUT[i - 1])
    spec["generator.up1.weight"] = ((c_id, 1024, 2, 2), "xavier_deconv")
    spec["generator.up1.bias"] = ((1024,), "bias_deconv:%d" % (1024 * 4))

    def aad(prefix, c_x, c_attr):
        spec[prefix + ".conv1.weight"] = ((c_x, c_attr, 1, 1), "xavier_conv")
        spec[prefix + ".conv1.bias"] = ((c_x,), "bias:%d" % c_attr)
        spec[prefix + ".conv2.weight"] = ((c_x, c_attr, 1, 1), "xavier_conv")
        spec[prefix + ".conv2.bias"] = ((c_x,), "bias:%d" % c_attr)
        spec[prefix + ".fc1.weight"] = ((c_x, c_id), "linear")
        spec[prefix + ".fc1.bias"] = ((c_x,), "zeros")
        spec[prefix + ".fc2.weight"] = ((c_x, c_id), "linear")
        spec[prefix + ".fc2.bias"] = ((c_x,), "zeros")
        spec[prefix + ".conv_h.weight"] = ((1, c_x, 1, 1), "xavier_conv")
        spec[prefix + ".conv_h.bias"] = ((1,), "bias:%d" % c_x)

    for k, (cin, cout, cattr) in enumerate(BLOCKS[backbone], start=1):
        for i in range(num_blocks):
            out = cin if i < num_blocks - 1 else cout
            aad("generator.AADBlk%d.add_blocks.%d" % (k, 3 * i), cin, cattr)
            spec["generator.AADBlk%d.add_blocks.%d.weight" % (k, 3 * i + 2)] = ((out, cin, 3, 3), "xavier_conv")
        if cin != cout:
            aad("generator.AADBlk%d.last_add_block.0" % k, cin, cattr)
            spec["generator.AADBlk%d.last_add_block.2.weight" % k] = ((cout, cin, 3, 3), "xavier_conv")
    return spec


def iresnet100_spec():
    spec = OrderedDict()
    spec["conv1.weight"] = ((64, 3, 3, 3), "kaiming_conv")
    _bn(spec, "bn1", 64)
    spec["prelu.weight"] = ((64,), "prelu")
    inpl = 64
    for l, (n, pl) in enumerate(zip(IRESNET100_LAYERS, IRESNET_PLANES), start=1):
        for i in range(n):
            p = "layer%d.%d" % (l, i)
            ci = inpl if i == 0 else pl
            _bn(spec, p + ".bn1", ci)
            spec[p + ".conv1.weight"] = ((pl, ci, 3, 3), "kaiming_conv")
            _bn(spec, p + ".bn2", pl)
            spec[p + ".prelu.weight"] = ((pl,), "prelu")
            spec[p + ".conv2.weight"] = ((pl, pl, 3, 3), "kaiming_conv_res")
            _bn(spec, p + ".bn3", pl)
            if i == 0:
                spec[p + ".downsample.0.weight"] = ((pl, ci, 1, 1), "kaiming_conv")
                _bn(spec, p + ".downsample.1", pl)
        inpl = pl
    _bn(spec, "bn2", 512)
    spec["fc.weight"] = ((512, 25088), "fc")
    spec["fc.bias"] = ((512,), "zeros")
    _bn(spec, "features", 512)
    return spec


def _fans(shape, transposed=False):
    rf = 1
    for s in shape[2:]:
        rf *= s
    # torch: fan_in = size(1)*rf, fan_out = size(0)*rf (ConvTranspose2d weights included)
    return shape[1] * rf, shape[0] * rf


def init_tensor(shape, kind, gen):
    if kind == "ones":
        return torch.ones(shape)
    if kind == "zeros":
        return torch.zeros(shape)
    if kind == "count":
        return torch.zeros((), dtype=torch.int64)
    if kind in ("xavier_conv", "xavier_deconv"):
        fi, fo = _fans(shape)
        return torch.randn(shape, generator=gen) * math.sqrt(2.0 / (fi + fo))
    if kind == "linear":
        return torch.randn(shape, generator=gen) * 0.001
    if kind.startswith("bias"):
        fan_in = int(kind.split(":")[1])
        b = 1.0 / math.sqrt(fan_in)
        return (torch.rand(shape, generator=gen) * 2 - 1) * b
    if kind == "kaiming_conv":  # variance-preserving stand-in for the upstream N(0, 0.1)
        fi, _ = _fans(shape)
        return torch.randn(shape, generator=gen) * math.sqrt(2.0 / fi)
    if kind == "kaiming_conv_res":  # last conv of a residual branch: damped so 49 blocks do not blow up
        fi, _ = _fans(shape)
        return torch.randn(shape, generator=gen) * (0.5 * math.sqrt(1.0 / fi))
    if kind == "prelu":
        return torch.full(shape, 0.25)
    if kind == "fc":
        return torch.randn(shape, generator=gen) * math.sqrt(1.0 / shape[1])
    if kind == "kaiming_uniform":  # nn.Conv2d default (kaiming_uniform_, a=sqrt(5)) = U(+-1/sqrt(fan_in))
        fi, _ = _fans(shape)
        return (torch.rand(shape, generator=gen) * 2 - 1) / math.sqrt(fi)
    if kind == "unit":  # torch spectral_norm: normalize(randn)
        v = torch.randn(shape, generator=gen)
        return v / v.norm().clamp_min(1e-12)
    raise ValueError(kind)


def init_state_dict(spec, seed=0):
    gen = torch.Generator().manual_seed(seed)
    return OrderedDict((k, init_tensor(shape, kind, gen)) for k, (shape, kind) in spec.items())


def perturb_(sd, seed=1):
    """The "perturbed" weight set of SURVEY 8(d): randomised BatchNorm parameters / running statistics and
    id-branch FCs large enough to matter, so BN folding and the identity path are numerically visible."""
    gen = torch.Generator().manual_seed(seed)
    for k, v in sd.items():
        if k.endswith("running_mean"):
            v.copy_(torch.randn(v.shape, generator=gen) * 0.1)
            base = k[: -len("running_mean")]
            sd[base + "running_var"].copy_(torch.rand(v.shape, generator=gen) + 0.5)
            sd[base + "weight"].copy_(torch.rand(v.shape, generator=gen) + 0.5)
            sd[base + "bias"].copy_(torch.randn(v.shape, generator=gen) * 0.1)
        elif ".fc1." in k or ".fc2." in k:
            if k.endswith("weight"):
                v.copy_(torch.randn(v.shape, generator=gen) * 0.05)
            else:
                v.copy_(torch.randn(v.shape, generator=gen) * 0.1)
    return sd


def unit_embedding_(arc_sd, pre_bn_std=281.0):
    """Synthetic stand-in for what training gives the ArcFace head: with random-init convs the pre-BN embedding of
    iresnet100 has a per-dimension std of ~280 (|z_id| ~ 6000), while a trained network's `features` BatchNorm1d
    leaves z_id ~ N(0,1) per dimension (|z_id| ~ 22, what synth.z_ids models and what the generator's id branch is
    scaled for).  Sets `features.running_var` so that the random-init network does the same.  Used by bench.py and
    the bench-shape tests; a z_id of magnitude 6000 drives every AAD layer into saturation, where no 16-bit
    arithmetic (the reference's own G.half() included) tracks the fp32 forward."""
    arc_sd["features.running_var"] = torch.full_like(arc_sd["features.running_var"], float(pre_bn_std) ** 2)
    return arc_sd


# ---------------------------------------------------------------------------------------------------------------
# pix2pix super-resolution generator: LIPSPADEGenerator with TestOptions (models/config_sr.py:3-29: ngf=48,
# norm_G='spectralspadesyncbatch3x3', num_upsampling_layers='normal', semantic_nc=3).  Key names follow
# models/networks/generator.py:27-63 (SPADEGenerator), :353-400 (LIPEncoder, LIPSPADEGenerator),
# models/networks/architecture.py:21-50 (SPADEResnetBlock, torch spectral_norm -> weight_orig / weight_u / weight_v),
# models/networks/normalization.py:63-92 (SPADE).
SR_NGF = 48
SR_ENC_C = [48, 96, 192, 384, 768, 768]           # LIPEncoder widths: ngf * min(2^i, 16)
# name, fin, fout of the seven SPADEResnetBlocks in forward order (generator.py:40-55)
SR_BLOCKS = [("head_0", 768, 768), ("G_middle_0", 768, 768), ("G_middle_1", 768, 768),
             ("ups.0", 768, 384), ("ups.1", 384, 192), ("ups.2", 192, 96), ("ups.3", 96, 48)]


def sr_spec():
    """OrderedDict key -> (shape, init kind) of LIPSPADEGenerator(TestOptions) in the reference's state_dict order."""
    spec = OrderedDict()
    nf = SR_NGF
    spec["fc.weight"] = ((16 * nf, 3, 3, 3), "xavier_conv")       # unused by LIPSPADEGenerator.forward (encode is overridden)
    spec["fc.bias"] = ((16 * nf,), "bias:27")

    def spade(prefix, norm_nc):
        nh = 128 if norm_nc > 128 else norm_nc
        spec[prefix + ".param_free_norm.running_mean"] = ((norm_nc,), "zeros")
        spec[prefix + ".param_free_norm.running_var"] = ((norm_nc,), "ones")
        spec[prefix + ".param_free_norm.num_batches_tracked"] = ((), "count")
        spec[prefix + ".mlp_shared.0.weight"] = ((nh, 3, 3, 3), "xavier_conv")
        spec[prefix + ".mlp_shared.0.bias"] = ((nh,), "bias:27")
        spec[prefix + ".mlp_gamma.weight"] = ((norm_nc, nh, 3, 3), "xavier_conv")
        spec[prefix + ".mlp_beta.weight"] = ((norm_nc, nh, 3, 3), "xavier_conv")

    def sn_conv(prefix, co, ci, k, bias):
        if bias:
            spec[prefix + ".bias"] = ((co,), "bias:%d" % (ci * k * k))
        spec[prefix + ".weight_orig"] = ((co, ci, k, k), "kaiming_uniform")
        spec[prefix + ".weight_u"] = ((co,), "unit")
        spec[prefix + ".weight_v"] = ((ci * k * k,), "unit")

    for name, fin, fout in SR_BLOCKS:
        fmid = min(fin, fout)
        sn_conv(name + ".conv_0", fmid, fin, 3, True)
        sn_conv(name + ".conv_1", fout, fmid, 3, True)
        if fin != fout:
            sn_conv(name + ".conv_s", fout, fin, 1, False)
        spade(name + ".norm_0", fin)
        spade(name + ".norm_1", fmid)
        if fin != fout:
            spade(name + ".norm_s", fin)
    for i, c in enumerate([8 * nf, 4 * nf, 2 * nf, nf]):
        spec["to_rgbs.%d.weight" % i] = ((3, c, 3, 3), "xavier_conv")
        spec["to_rgbs.%d.bias" % i] = ((3,), "bias:%d" % (9 * c))
    spec["lip_encoder.model.0.weight"] = ((nf, 3, 3, 3), "xavier_conv")
    for i in range(5):
        c, cn = SR_ENC_C[i], SR_ENC_C[i + 1]
        spec["lip_encoder.model.%d.logit.0.weight" % (3 + 4 * i)] = ((c, c, 3, 3), "xavier_conv")
        spec["lip_encoder.model.%d.logit.1.weight" % (3 + 4 * i)] = ((c,), "ones")
        spec["lip_encoder.model.%d.logit.1.bias" % (3 + 4 * i)] = ((c,), "zeros")
        spec["lip_encoder.model.%d.weight" % (4 + 4 * i)] = ((cn, c, 3, 3), "xavier_conv")
        spec["lip_encoder.model.%d.bias" % (4 + 4 * i)] = ((cn,), "bias:%d" % (9 * c))
    return spec


def perturb_sr_(sd, seed=1):
    """Parity-test weight set for the SR generator: non-trivial InstanceNorm affine parameters of the LIP gates
    (identity at init, so the gate would otherwise be numerically invisible)."""
    gen = torch.Generator().manual_seed(seed)
    for k, v in sd.items():
        if ".logit.1.weight" in k:
            v.copy_(torch.rand(v.shape, generator=gen) + 0.5)
        elif ".logit.1.bias" in k:
            v.copy_(torch.randn(v.shape, generator=gen) * 0.3)
    return sd
