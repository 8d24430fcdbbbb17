"""Profiling target of round 2: one un-graphed ArcFace forward (NVTX range "arc") and one un-graphed generator
forward (range "gen") at the bench's batch (B = 64, fp16 storage), after a warm-up pass of both; paste-back of the 64
faces into 1080p frames (range "paste").  Used as
    ncu --set full --nvtx --nvtx-include "gen/" ... python tools/ncu_round2.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
import ghost_b200

B = 64
dev = torch.device("cuda", 0)
G = ghost_b200.AEI_Net("unet", 2, 512, seed=0).eval().half().cuda()
arc = bench.make_arcface(torch.float16)
d = bench.make_inputs(0, "default")
src = torch.from_numpy(d["sources"]).cuda()
crp = torch.from_numpy(d["crops"]).cuda()
frm = torch.from_numpy(d["frames"]).cuda()
msk = torch.from_numpy(d["masks"]).cuda()
out = torch.empty_like(frm)
ff = d["face_frame"]


def once(tag):
    if tag:
        torch.cuda.nvtx.range_push("arc")
    z = arc.embed_hwc(src, out_dtype=torch.float16)
    if tag:
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_pop()
        torch.cuda.nvtx.range_push("gen")
    G.swap_u8(crp, z, write_u8=False)
    if tag:
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_pop()
        torch.cuda.nvtx.range_push("paste")
    yptr, ydt = G.y_nhwc(B, dev)
    ghost_b200.paste_back(frm, (yptr, ydt, B), msk, d["tfms"], ff, mode="exact", out=out)
    torch.cuda.synchronize()
    if tag:
        torch.cuda.nvtx.range_pop()


once(False)
once(True)
print("ok")
