"""Timing of the device mask synthesis (ghost_face_masks) for a batch of faces, per preset (GPU box)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from ghost_b200 import masks as gm, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
lms = np.stack([synth.landmarks106(np.random.default_rng(i), 224, shift=0.0) for i in range(n)])
for preset in ([15, 15, 10], [10, 10, 8], [-5, 5, 10], [5, 5, 5]):
    prs = [preset] * n
    out = torch.empty((n, 224, 224), dtype=torch.float64, device="cuda")
    for _ in range(3):
        gm.face_masks(lms, prs, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(20):
        gm.face_masks(lms, prs, out=out)
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 20 * 1e3
    ms = e0.elapsed_time(e1) / 20
    print("preset %-14s n=%d: %.3f ms per call on the stream (%.1f us / face), %.3f ms wall incl. the host rasteriser; "
          "%.0f masks/s" % (preset, n, ms, ms / n * 1e3, wall, n / wall * 1e3))
