"""CPU oracle for the video paste-back (float mode) - TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: utils/inference/video_processing.py:212-233 (get_final_video) calls kornia==0.5.4
(requirements.txt:11) `invert_affine_transform` and `warp_affine`; kornia is not vendored in the reference
and not installed here, and no reference test pins the result.  This restates kornia 0.5.4's published
algorithm with torch fp32 on the CPU:

    invert_affine_transform(M):  pad to 3x3, torch.inverse, take [:2]
    warp_affine(src, M, (H,W)):  M3 = pad(M); N(h,w) = [[2/(w-1),0,-1],[0,2/(h-1),-1],[0,0,1]]
                                 dst_norm<-src_norm = N(H,W) @ M3 @ inverse(N(h_in,w_in))
                                 grid = F.affine_grid(inverse(dst_norm<-src_norm)[:, :2], align_corners=True)
                                 F.grid_sample(src, grid, bilinear, zeros, align_corners=True)

and the per-face composite of the reference: fp32 blend, truncating uint8 cast, faces of one frame
composing through uint8 (video_processing.py:233,235).
"""
import numpy as np
import torch
import torch.nn.functional as F

from .paste_ref import resize_linear_u8


def _pad3(m):
    out = torch.zeros(m.shape[0], 3, 3, dtype=m.dtype)
    out[:, :2] = m
    out[:, 2, 2] = 1
    return out


def _norm_pix(h, w):
    return torch.tensor([[2.0 / (w - 1), 0, -1], [0, 2.0 / (h - 1), -1], [0, 0, 1]], dtype=torch.float32)[None]


def invert_affine_transform(m):
    return torch.inverse(_pad3(m))[:, :2]


def warp_affine(src, m, dsize):
    B, Cc, h, w = src.shape
    H, W = dsize
    dst_norm_src_norm = _norm_pix(H, W) @ (_pad3(m) @ torch.inverse(_norm_pix(h, w)))
    theta = torch.inverse(dst_norm_src_norm)[:, :2]
    grid = F.affine_grid(theta, [B, Cc, H, W], align_corners=True)
    return F.grid_sample(src, grid, mode="bilinear", padding_mode="zeros", align_corners=True)


def composite_frame_ref(frame, crops256, masks, tfms):
    """One frame of get_final_video: crops256[j] uint8 BGR 256x256x3, masks[j] 224x224, tfms[j] 2x3 float64."""
    H, W = frame.shape[:2]
    result = frame
    for crop, mask, tfm in zip(crops256, masks, tfms):
        swap = resize_linear_u8(crop, (224, 224))
        swap_t = torch.from_numpy(swap).permute(2, 0, 1).unsqueeze(0).type(torch.float32)
        mask_t = torch.from_numpy(np.asarray(mask)).unsqueeze(0).unsqueeze(0).type(torch.float32)
        full = torch.from_numpy(result).permute(2, 0, 1).unsqueeze(0)
        mat = torch.from_numpy(np.asarray(tfm)).unsqueeze(0).type(torch.float32)
        mat_rev = invert_affine_transform(mat)
        sw = warp_affine(swap_t, mat_rev, (H, W))
        mk = warp_affine(mask_t, mat_rev, (H, W))
        result = (mk * sw + (1 - mk) * full).type(torch.uint8).squeeze(0).permute(1, 2, 0).numpy()
    return result
