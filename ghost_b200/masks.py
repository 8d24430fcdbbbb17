"""Blend-mask synthesis from 106-point landmarks, on the device.

Same names and argument meaning as utils/inference/masks.py:5-108 of the reference (face_mask_static,
erode_and_blur's presets); the arithmetic - eyebrow lift, convex hull, cv2.fillConvexPoly, cv2.erode / dilate, border
clip, cv2.GaussianBlur (uint8, fixed point), /255 - runs behind ghost_face_masks (include/ghost_b200.h) and is
bit-exact against the reference's OpenCV calls.  Landmarks come from a landmark model that is out of scope; they are
an input here.  There is no CPU path.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

MASK_SIZE = 224


def choose_params(landmarks, landmarks_tgt):
    """The preset face_mask_static picks when params is None (masks.py:43-64): [erode, sigmaX, sigmaY]."""
    left = np.sum((landmarks[1][0] - landmarks_tgt[1][0], landmarks[2][0] - landmarks_tgt[2][0],
                   landmarks[13][0] - landmarks_tgt[13][0]))
    right = np.sum((landmarks_tgt[17][0] - landmarks[17][0], landmarks_tgt[18][0] - landmarks[18][0],
                    landmarks_tgt[29][0] - landmarks[29][0]))
    offset = max(left, right)
    if offset > 6:
        return [15, 15, 10]
    if offset > 3:
        return [10, 10, 8]
    if offset < -3:
        return [-5, 5, 10]
    return [5, 5, 5]


def face_masks(landmarks, params, device=None, dtype=torch.float64, out=None):
    """Batched device call: landmarks [n,106,2] (any float / int array), params [n,3] ints (erode, sigmaX, sigmaY).
    Returns a device tensor [n,224,224] of `dtype` (float64 for get_final_image, float32 for get_final_video)."""
    if not torch.cuda.is_available():
        raise RuntimeError("ghost_b200 needs a CUDA device (no CPU fallback)")
    lm = np.ascontiguousarray(np.asarray(landmarks, dtype=np.float32).reshape(-1, 106, 2))
    pr = np.ascontiguousarray(np.asarray(params, dtype=np.int32).reshape(-1, 3))
    n = lm.shape[0]
    if pr.shape[0] != n:
        raise ValueError("one (erode, sigmaX, sigmaY) triple per landmark set")
    if dtype not in (torch.float64, torch.float32):
        raise TypeError("mask dtype must be float64 or float32")
    dev = device or torch.device("cuda", torch.cuda.current_device())
    if out is None:
        out = torch.empty((n, MASK_SIZE, MASK_SIZE), dtype=dtype, device=dev)
    elif tuple(out.shape) != (n, MASK_SIZE, MASK_SIZE) or out.dtype != dtype or not out.is_cuda or not out.is_contiguous():
        raise ValueError("out must be a contiguous CUDA tensor [n,224,224] of the requested dtype")
    if n == 0:
        return out
    lib = _lib.load()
    with torch.cuda.device(out.device):
        _lib.check(lib.ghost_face_masks(lm.ctypes.data_as(C.POINTER(C.c_float)), pr.ctypes.data_as(C.POINTER(C.c_int)), n,
                                        C.c_void_p(out.data_ptr()), int(dtype == torch.float64),
                                        C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


def face_mask_static(image, landmarks, landmarks_tgt, params=None):
    """utils/inference/masks.py:39-85: returns (mask, params) when params is None, else mask - numpy float64 in [0,1]
    like the reference (the batched, device-resident form is face_masks)."""
    if tuple(np.shape(image)[:2]) != (MASK_SIZE, MASK_SIZE):
        raise ValueError("face_mask_static works on the 224x224 aligned crop (got %s)" % (np.shape(image)[:2],))
    chosen = choose_params(landmarks, landmarks_tgt) if params is None else [int(params[0]), int(params[1]), int(params[2])]
    mask = face_masks(np.asarray(landmarks)[None], [chosen])[0].cpu().numpy()
    if params is None:
        return mask, chosen
    return mask
