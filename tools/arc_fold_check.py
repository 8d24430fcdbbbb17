"""iresnet100 with the perturbed weight set (random BatchNorm statistics): the fp16 tensor-core forward against the
fp32 CUDA-core forward of the same library (which keeps bn1 as a separate pass).  Run with GHOST_FOLD_BN1=0 / 1: the
two distances must agree - a wrong border-class bias of the folded bn1 would show up at the image borders."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ghost_b200 as gb
from ghost_b200 import weights as W

B = int(sys.argv[1]) if len(sys.argv) > 1 else 6
sd = W.perturb_(W.init_state_dict(W.iresnet100_spec(), seed=0), seed=1)
g = torch.Generator().manual_seed(8)
x = (torch.rand(B, 3, 112, 112, generator=g) * 2 - 1).cuda()
out = {}
for mode in ("fp32", "fp16", "bf16"):
    net = gb.iresnet100(fp16=False)
    net.load_state_dict(sd)
    net = net.eval().cuda()
    if mode == "fp16":
        net = net.half()
    elif mode == "bf16":
        net = net.bfloat16()
    out[mode] = net(x.to(net.dtype)).float().cpu()
ref = out["fp32"]
for mode in ("fp16", "bf16"):
    z = out[mode]
    rel = ((z - ref).norm(dim=1) / ref.norm(dim=1))
    print("GHOST_FOLD_BN1=%s %s vs fp32: rel err max %.3e mean %.3e, cos min %.6f" % (
        os.environ.get("GHOST_FOLD_BN1", "default"), mode, rel.max().item(), rel.mean().item(),
        torch.nn.functional.cosine_similarity(z, ref, dim=1).min().item()))
