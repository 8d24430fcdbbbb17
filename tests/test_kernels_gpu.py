"""Kernel-level parity (GPU): every conv family of the hot path at its real layer shapes, the fused AAD
layer and the bilinear x2, each against plain torch fp32 ops on the same (bf16-rounded) operands."""
import ctypes as C
import zlib

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from ghost_b200 import _lib
from util import fptr, nhwc, round16, run_conv, torch_dtype

pytestmark = pytest.mark.gpu

# (name, N, Cin, H, W, Cout, kind, k, stride, pad, act)   kind 0 = Conv2d, 1 = ConvTranspose2d k4 s2 p1
CONV_CASES = [
    ("g7_3x3_64_256px", 1, 64, 64, 64, 64, 0, 3, 1, 1, 0),
    ("g7_3x3_128_to_64", 2, 128, 32, 32, 64, 0, 3, 1, 1, 0),
    ("g7_3x3_1024_2px", 3, 1024, 2, 2, 1024, 0, 3, 1, 1, 0),
    ("g7_3x3_1024_8px", 2, 1024, 8, 8, 512, 0, 3, 1, 1, 0),
    ("g7_3x3_64_to_3", 1, 64, 32, 32, 3, 0, 3, 1, 1, 0),
    ("g7_3x3_ragged_w", 1, 64, 20, 12, 32, 0, 3, 1, 1, 0),
    ("g1_4x4s2_32_to_64", 1, 32, 64, 64, 64, 0, 4, 2, 1, 1),
    ("g1_4x4s2_512_to_1024", 2, 512, 8, 8, 1024, 0, 4, 2, 1, 1),
    ("g1_4x4s2_1024_4px", 2, 1024, 4, 4, 1024, 0, 4, 2, 1, 1),
    ("g1_4x4s2_cin3", 1, 3, 64, 64, 32, 0, 4, 2, 1, 1),
    ("g2_deconv_1024_2px", 2, 1024, 2, 2, 1024, 1, 4, 2, 1, 1),
    ("g2_deconv_256_32px", 1, 256, 32, 32, 64, 1, 4, 2, 1, 1),
    ("g2_deconv_64_to_32", 1, 64, 16, 16, 32, 1, 4, 2, 1, 1),
    ("g6_1x1_64", 1, 64, 64, 64, 128, 0, 1, 1, 0, 0),
    ("g6_1x1_2048_4px", 2, 2048, 4, 4, 256, 0, 1, 1, 0, 0),
    ("a1_3x3s2_64_112px", 1, 64, 56, 56, 64, 0, 3, 2, 1, 0),
    ("a1_1x1s2_128", 1, 128, 28, 28, 256, 0, 1, 2, 0, 0),
    ("a1_3x3_cin3_prelu_like", 1, 3, 56, 56, 64, 0, 3, 1, 1, 3),
    ("k32_3x3_32ch", 1, 32, 16, 16, 32, 0, 3, 1, 1, 0),
    # more tiles than SMs: persistent CTAs loop over several tiles (TMEM double buffer, smem ring wrap-around)
    ("multi_tile_3x3_64_256px", 2, 64, 256, 256, 64, 0, 3, 1, 1, 0),
    ("multi_tile_3x3_128_to_64", 3, 128, 128, 128, 64, 0, 3, 1, 1, 0),
    ("multi_tile_4x4s2_32_to_64", 4, 32, 128, 128, 64, 0, 4, 2, 1, 1),
    ("multi_tile_deconv_128_to_32", 3, 128, 64, 64, 32, 1, 4, 2, 1, 1),
    ("multi_tile_3x3_64_to_3", 2, 64, 256, 256, 3, 0, 3, 1, 1, 0),
    # row-window mode (wide 3x3 stride-1 convs): streamed weights (128 -> 128), 112-pixel rows, ragged split rows
    ("window_3x3_128_to_128", 2, 128, 128, 128, 128, 0, 3, 1, 1, 3),
    ("window_3x3_64_112px", 2, 64, 112, 112, 64, 0, 3, 1, 1, 0),
    ("window_3x3_64_150px_ragged", 1, 64, 37, 150, 48, 0, 3, 1, 1, 1),
    ("window_3x3_192_to_96", 1, 192, 40, 128, 96, 0, 3, 1, 1, 0),
    # <= 3 output channels: the nine taps ride in N (tapn.cu); ragged tiles, several K chunks, the smallest image
    ("tapn_3x3_128_to_2_ragged", 2, 128, 23, 45, 2, 0, 3, 1, 1, 0),
    ("tapn_3x3_128_to_1", 1, 128, 16, 16, 1, 0, 3, 1, 1, 0),
    ("tapn_3x3_256_to_1_igemm", 1, 256, 16, 16, 1, 0, 3, 1, 1, 0),
    ("tapn_3x3_64_to_3_8x16", 3, 64, 8, 16, 3, 0, 3, 1, 1, 0),
]


def torch_conv(x, w, b, kind, k, stride, pad, act, slope):
    if kind == 0:
        y = F.conv2d(x, w, b, stride=stride, padding=pad)
    else:
        y = F.conv_transpose2d(x, w, b, stride=2, padding=1)
    if act == 1:
        y = F.leaky_relu(y, slope)
    elif act == 3:
        y = F.relu(y)
    return y


@pytest.mark.parametrize("mode", [_lib.F16, _lib.BF16, _lib.F32], ids=["fp16", "bf16", "fp32"])
@pytest.mark.parametrize("case", CONV_CASES, ids=[c[0] for c in CONV_CASES])
def test_conv(case, mode):
    name, N, Cin, H, W, Cout, kind, k, stride, pad, act = case
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.Generator(device="cpu").manual_seed(zlib.crc32(name.encode()))
    x = (torch.rand(N, Cin, H, W, generator=g) * 2 - 1).cuda()
    wshape = (Cout, Cin, k, k) if kind == 0 else (Cin, Cout, 4, 4)
    w = (torch.randn(wshape, generator=g) / (Cin * k * k / (4 if kind else 1)) ** 0.5).cuda()
    b = (torch.randn(Cout, generator=g) * 0.1).cuda()
    x, w = round16(x, mode), round16(w, mode)
    y = run_conv(x, w, b, kind, k, stride, pad, act, 0.1, mode)
    ref = torch_conv(x.double(), w.double(), b.double(), kind, k, stride, pad, act, 0.1).float()
    assert y.shape == ref.shape
    assert torch.isfinite(y).all(), "output has NaN/Inf (unwritten pixels?)"
    err = (y - ref).abs().max().item()
    scale = ref.abs().max().item()
    tol = {_lib.F32: 2e-5, _lib.BF16: 6e-3, _lib.F16: 8e-4}[mode] * max(scale, 1.0)
    assert err <= tol, "%s: max err %.3e (scale %.3f)" % (name, err, scale)


AAD_CASES = [  # N, H, c_x, c_attr
    (2, 2, 1024, 1024), (1, 4, 1024, 2048), (2, 16, 1024, 512), (1, 32, 512, 256), (1, 64, 256, 128),
    (1, 128, 128, 64), (1, 64, 64, 64), (2, 32, 64, 32), (3, 256, 64, 64), (2, 128, 128, 64),
]


def torch_aad(h, za, zid, p):
    hn = F.instance_norm(h, eps=1e-5)
    ga = F.conv2d(za, p["c1w"], p["c1b"])
    ba = F.conv2d(za, p["c2w"], p["c2b"])
    gi = F.linear(zid, p["f1w"], p["f1b"])[:, :, None, None]
    bi = F.linear(zid, p["f2w"], p["f2b"])[:, :, None, None]
    M = torch.sigmoid(F.conv2d(hn, p["chw"], p["chb"]))
    return F.relu((1 - M) * (ga * hn + ba) + M * (gi * hn + bi))


@pytest.mark.parametrize("mode", [_lib.F16, _lib.BF16, _lib.F32], ids=["fp16", "bf16", "fp32"])
@pytest.mark.parametrize("case", AAD_CASES, ids=["n%d_hw%d_cx%d_ca%d" % c for c in AAD_CASES])
def test_aad_layer(case, mode):
    N, H, c_x, c_attr = case
    c_id = 512
    g = torch.Generator().manual_seed(c_x * 7 + H)
    dt = torch_dtype(mode)
    h = (torch.randn(N, c_x, H, H, generator=g) * 1.5 + 0.3).cuda()
    za = torch.randn(N, c_attr, H, H, generator=g).cuda()
    zid = torch.randn(N, c_id, generator=g)
    p = {
        "c1w": torch.randn(c_x, c_attr, 1, 1, generator=g) / c_attr ** 0.5, "c1b": torch.randn(c_x, generator=g) * 0.1,
        "c2w": torch.randn(c_x, c_attr, 1, 1, generator=g) / c_attr ** 0.5, "c2b": torch.randn(c_x, generator=g) * 0.1,
        "f1w": torch.randn(c_x, c_id, generator=g) * 0.04, "f1b": torch.randn(c_x, generator=g) * 0.1,
        "f2w": torch.randn(c_x, c_id, generator=g) * 0.04, "f2b": torch.randn(c_x, generator=g) * 0.1,
        "chw": torch.randn(1, c_x, 1, 1, generator=g) / c_x ** 0.5, "chb": torch.randn(1, generator=g) * 0.1,
    }
    h, za = round16(h, mode), round16(za, mode)
    p["c1w"], p["c2w"] = round16(p["c1w"], mode), round16(p["c2w"], mode)
    hn, zan = nhwc(h, dt), nhwc(za, dt)
    out = torch.full((N, H, H, c_x), float("nan"), dtype=dt, device="cuda")
    hp = {k: np.ascontiguousarray(v.numpy().astype(np.float32)) for k, v in p.items()}
    zh = np.ascontiguousarray(zid.numpy().astype(np.float32))
    lib = _lib.load()
    rc = lib.ghost_test_aad(C.c_void_p(hn.data_ptr()), C.c_void_p(zan.data_ptr()), fptr(zh), N, H, H, c_x, c_attr, c_id,
                            fptr(hp["c1w"]), fptr(hp["c1b"]), fptr(hp["c2w"]), fptr(hp["c2b"]), fptr(hp["f1w"]),
                            fptr(hp["f1b"]), fptr(hp["f2w"]), fptr(hp["f2b"]), fptr(hp["chw"]), float(hp["chb"][0]),
                            C.c_void_p(out.data_ptr()), mode, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, test_api=True)
    pc = {k: v.cuda().double() for k, v in p.items()}
    ref = torch_aad(h.double(), za.double(), zid.cuda().double(), pc).float()
    got = out.float().permute(0, 3, 1, 2)
    assert torch.isfinite(got).all()
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item()
    tol = {_lib.F32: 5e-5, _lib.BF16: 1.2e-2, _lib.F16: 1.6e-3}[mode] * max(scale, 1.0)
    assert err <= tol, "max err %.3e (scale %.3f)" % (err, scale)


@pytest.mark.parametrize("mode", [_lib.F16, _lib.BF16, _lib.F32], ids=["fp16", "bf16", "fp32"])
@pytest.mark.parametrize("shape", [(2, 1024, 2, 2), (1, 64, 128, 128), (3, 128, 7, 5)], ids=str)
def test_upsample2x(shape, mode):
    N, Cc, H, W = shape
    dt = torch_dtype(mode)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(N, Cc, H, W, generator=g).cuda()
    x = round16(x, mode)
    xn = nhwc(x, dt)
    y = torch.full((N, 2 * H, 2 * W, Cc), float("nan"), dtype=dt, device="cuda")
    lib = _lib.load()
    rc = lib.ghost_test_upsample2x(C.c_void_p(xn.data_ptr()), C.c_void_p(y.data_ptr()), mode, N, H, W, Cc,
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, test_api=True)
    ref = F.interpolate(x, scale_factor=2, mode="bilinear", align_corners=True)
    got = y.float().permute(0, 3, 1, 2)
    tol = {_lib.F32: 1e-5, _lib.BF16: 2e-2, _lib.F16: 2.5e-3}[mode]
    assert (got - ref).abs().max().item() <= tol


@pytest.mark.parametrize("mode", [_lib.F16, _lib.BF16, _lib.F32], ids=["fp16", "bf16", "fp32"])
@pytest.mark.parametrize("shape", [(2, 64, 16, 16), (1, 1024, 2, 2), (3, 128, 32, 32), (2, 64, 128, 128)])
def test_upsample2x_fused_stats(shape, mode):
    """The fused kernel writes the same pixels as the plain upsample and the InstanceNorm statistics of exactly
    those (rounded) pixels."""
    N, Cc, H, W = shape
    dt = torch_dtype(mode)
    g = torch.Generator().manual_seed(11)
    x = (torch.randn(N, Cc, H, W, generator=g) * 2.0 + 0.7).cuda()
    xn = nhwc(x, dt)
    lib = _lib.load()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    y0 = torch.empty((N, 2 * H, 2 * W, Cc), dtype=dt, device="cuda")
    _lib.check(lib.ghost_test_upsample2x(C.c_void_p(xn.data_ptr()), C.c_void_p(y0.data_ptr()), mode, N, H, W, Cc, st),
               test_api=True)
    y1 = torch.full_like(y0, float("nan"))
    mean = np.zeros((N, Cc), np.float32)
    var = np.zeros((N, Cc), np.float32)
    _lib.check(lib.ghost_test_upsample2x_stats(C.c_void_p(xn.data_ptr()), C.c_void_p(y1.data_ptr()), mode, N, H, W, Cc,
                                               mean.ctypes.data_as(C.POINTER(C.c_float)),
                                               var.ctypes.data_as(C.POINTER(C.c_float)), st), test_api=True)
    assert torch.equal(y0, y1)
    yd = y1.double()
    ref_mean = yd.mean(dim=(1, 2)).cpu().numpy()
    ref_var = yd.var(dim=(1, 2), unbiased=False).cpu().numpy()
    assert np.abs(mean - ref_mean).max() <= 1e-5 * max(1.0, np.abs(ref_mean).max())
    assert np.abs(var - ref_var).max() <= 2e-5 * max(1.0, ref_var.max())
