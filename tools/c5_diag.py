"""Diagnostic for configs[4] parity: error of the SR stage alone (same uint8 input crops) in fp16 / bf16 / fp32 mode
against the CPU oracle, one batch of 20."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, ghost_b200
from ghost_b200 import weights as W
from oracle import sr_ref
torch.set_num_threads(os.cpu_count() or 1)
d = bench.make_inputs(0, "default")
G = ghost_b200.AEI_Net("linknet", 3, 512, seed=0).eval().half().cuda()
arc = bench.make_arcface(torch.float16)
sw = ghost_b200.embed_and_swap(G, arc, torch.from_numpy(d["sources"][:20]).cuda(), torch.from_numpy(d["crops"][:20]).cuda())
sw_h = sw.cpu().numpy()
sd = W.init_state_dict(W.sr_spec(), seed=0)
with torch.no_grad():
    ref = sr_ref.face_enhancement_ref({k: v.float() for k, v in sd.items() if v.dtype.is_floating_point}, sw_h, {})
for name, dt in (("fp32", torch.float32), ("fp16", torch.float16), ("bf16", torch.bfloat16)):
    SR = ghost_b200.LIPSPADEGenerator().to(dt).cuda()
    out = SR.enhance_u8(sw).cpu().numpy()
    dd = np.abs(out.astype(int) - ref.astype(int))
    print("SR %s vs CPU fp32 on the same crops: mean %.3f p99 %d p99.9 %d max %d grey levels" % (
        name, dd.mean(), np.percentile(dd, 99), np.percentile(dd, 99.9), dd.max()))
