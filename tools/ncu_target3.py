"""Bandwidth-bound kernels of the headline step for an ncu --set full capture: one paste-back of 64 faces into 64 1080p
frames (frame copy + resize_crops_kernel + paste_faces_kernel), run twice (the second is the warm one)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import ghost_b200
from ghost_b200 import synth

B, H, W = 64, 1080, 1920
rng = np.random.default_rng(100)
crops = torch.from_numpy(synth.crops_u8(B, seed=0)).cuda()
base = synth.frame_u8(H, W, seed=0)
frames = torch.from_numpy(np.stack([np.roll(base, 17 * i, axis=1) for i in range(B)])).cuda()
masks = torch.from_numpy(np.stack([synth.soft_mask(rng) for _ in range(B)])).cuda()
tfms = np.stack([synth.similarity_tfm(H, W, rng, clipped=(i % 20 == 19)) for i in range(B)])
ff = np.arange(B, dtype=np.int32)
out = torch.empty_like(frames)
for _ in range(2):
    ghost_b200.paste_back(frames, crops, masks, tfms, ff, mode="exact", out=out)
    torch.cuda.synchronize()
print("ok")
