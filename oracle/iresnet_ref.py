"""CPU oracle for the ArcFace iresnet100 embedding - TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the reference imports arcface_model/iresnet.py, which is not in its tree
(download_models.sh:2-3 fetches it from a GitHub release; it is a copy of insightface
recognition/arcface_torch/backbones/iresnet.py, no version pin) and no reference test or fixture
pins its output.  This file restates the published architecture as a functional torch fp32 forward:

    conv3x3(3->64, s1, no bias) -> BN -> PReLU(64)
    layers [3, 13, 30, 3] of IBasicBlock(planes 64/128/256/512), first block of each layer stride 2 with a
        conv1x1(s2, no bias)+BN downsample on the identity:
        BN1 -> conv3x3(s1) -> BN2 -> PReLU -> conv3x3(stride) -> BN3 -> + identity
    BN2d(512) -> flatten -> (dropout p=0) -> Linear(25088 -> 512) -> BN1d(512);   eps = 1e-5 everywhere.
    The output is NOT L2-normalised (utils/inference/core.py:54).

Call sites in the reference: inference.py:15,33-36; utils/inference/core.py:45,54;
utils/inference/video_processing.py:139.  Pre-processing: utils/inference/image_processing.py:23-48 and
F.interpolate(scale_factor=0.5, mode='bilinear', align_corners=True) at core.py:45,54.
"""
import torch
import torch.nn.functional as F

LAYERS = [3, 13, 30, 3]


def _bn(x, sd, p):
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"], sd[p + ".bias"],
                        False, 0.0, 1e-5)


def iresnet100_forward(sd, x):
    with torch.no_grad():
        sd = {k: v.float() for k, v in sd.items() if v.dtype.is_floating_point}
        x = F.conv2d(x.float(), sd["conv1.weight"], None, stride=1, padding=1)
        x = F.prelu(_bn(x, sd, "bn1"), sd["prelu.weight"])
        for l, n in enumerate(LAYERS, start=1):
            for i in range(n):
                p = "layer%d.%d" % (l, i)
                stride = 2 if i == 0 else 1
                out = _bn(x, sd, p + ".bn1")
                out = F.conv2d(out, sd[p + ".conv1.weight"], None, stride=1, padding=1)
                out = F.prelu(_bn(out, sd, p + ".bn2"), sd[p + ".prelu.weight"])
                out = F.conv2d(out, sd[p + ".conv2.weight"], None, stride=stride, padding=1)
                out = _bn(out, sd, p + ".bn3")
                if i == 0:
                    x = _bn(F.conv2d(x, sd[p + ".downsample.0.weight"], None, stride=stride), sd, p + ".downsample.1")
                x = out + x
        x = _bn(x, sd, "bn2")
        x = torch.flatten(x, 1)
        x = F.linear(x, sd["fc.weight"], sd["fc.bias"])
        return F.batch_norm(x, sd["features.running_mean"], sd["features.running_var"], sd["features.weight"],
                            sd["features.bias"], False, 0.0, 1e-5)


def embed_hwc_ref(sd, images):
    """normalize_and_torch(_batch) (image_processing.py:23-48) + the x0.5 bilinear resample (core.py:45,54) +
    netArc.  images: numpy [B,224,224,3] RGB, uint8 or float in [0,1]."""
    x = torch.from_numpy(images.copy()).float()
    if x.max() > 1.:
        x = x / 255.
    x = (x.permute(0, 3, 1, 2) - 0.5) / 0.5
    x = F.interpolate(x, scale_factor=0.5, mode="bilinear", align_corners=True)
    return iresnet100_forward(sd, x)
