"""Helpers shared by the GPU parity tests."""
import ctypes as C

import numpy as np
import torch

from ghost_b200 import _lib


def nhwc(x_nchw, dtype):
    return x_nchw.permute(0, 2, 3, 1).contiguous().to(dtype)


def bf16_round(t):
    return t.to(torch.bfloat16).to(torch.float32)


def f16_round(t):
    return t.to(torch.float16).to(torch.float32)


def round16(t, mode):
    """t rounded through the 16-bit storage type of a tensor-core mode (unchanged in fp32 mode)"""
    return bf16_round(t) if mode == _lib.BF16 else f16_round(t) if mode == _lib.F16 else t


def torch_dtype(mode):
    return {_lib.F32: torch.float32, _lib.BF16: torch.bfloat16, _lib.F16: torch.float16}[mode]


def fptr(a):
    """host float32 numpy array -> void*"""
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return C.c_void_p(a.ctypes.data)


def run_conv(x_nchw, w, bias, kind, k, stride, pad, act, slope, mode, bn_hint=0):
    """ghost_test_conv on a [N,C,H,W] fp32 cuda tensor; returns the NCHW fp32 result."""
    lib = _lib.load()
    dt = torch_dtype(mode)
    N, Cin, H, W = x_nchw.shape
    x = nhwc(x_nchw, dt)
    wh = np.ascontiguousarray(w.detach().cpu().numpy().astype(np.float32))
    bh = None if bias is None else np.ascontiguousarray(bias.detach().cpu().numpy().astype(np.float32))
    if kind == 0:
        Cout = w.shape[0]
        Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    else:
        Cout = w.shape[1]
        Ho, Wo = 2 * H, 2 * W
    y = torch.full((N, Ho, Wo, Cout), float("nan"), dtype=dt, device=x.device)
    rc = lib.ghost_test_conv(C.c_void_p(x.data_ptr()), N, H, W, Cin, fptr(wh), None if bh is None else fptr(bh), Cout,
                             kind, k, stride, pad, act, slope, C.c_void_p(y.data_ptr()), mode, bn_hint,
                             C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, test_api=True)
    return y.float().permute(0, 3, 1, 2).contiguous()
