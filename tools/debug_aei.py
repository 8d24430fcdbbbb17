"""Debug helper (GPU box): per-tensor error of the CUDA AEI_Net against the CPU oracle and the bf16 emulation."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ghost_b200
from ghost_b200 import synth, weights as W
from oracle import aei_ref, aei_emul

bb, nb, B = (sys.argv[1] if len(sys.argv) > 1 else "unet"), int(sys.argv[2]) if len(sys.argv) > 2 else 2, 1
mode = sys.argv[3] if len(sys.argv) > 3 else "bf16"
pert = (sys.argv[4] if len(sys.argv) > 4 else "1") == "1"
sd = W.init_state_dict(W.aei_spec(bb, nb, 512), seed=0)
if pert:
    W.perturb_(sd, seed=1)
Xt, z = synth.crops_pm1(B, seed=7), synth.z_ids(B, seed=7)
tr, te = {}, {}
Yr, ar = aei_ref.aei_forward(sd, Xt, z, bb, nb, trace=tr)
Ye, ae = aei_emul.aei_forward_emul(sd, Xt, z, bb, nb, trace=te)
dt = torch.bfloat16 if mode == "bf16" else torch.float32
G = ghost_b200.AEI_Net(bb, nb, 512).eval()
G.load_state_dict(sd)
G = G.to(dt).cuda()
Y, attrs = G(Xt.cuda().to(dt), z.cuda().to(dt))
torch.cuda.synchronize()
def line(name, got, ref, emu):
    d, e, de = (got - ref).abs(), (got - emu).abs(), (emu - ref).abs()
    print("%-6s %-22s |ref|max %.3f  cuda-ref max %.3e mean %.3e | cuda-emul max %.3e mean %.3e | emul-ref max %.3e mean %.3e" %
          (name, tuple(ref.shape), ref.abs().max(), d.max(), d.mean(), e.max(), e.mean(), de.max(), de.mean()))
for k, (a, r, e) in enumerate(zip(attrs, ar, ae), 1):
    line("attr%d" % k, a.float().cpu(), r, e)
for k in range(1, 9):
    for nm in ("m%d" % k, "y%d" % k):
        got = G.intermediate(nm, B).cpu()
        if nm == "y8":
            got = torch.tanh(got) if False else got
        line(nm, got, tr[nm], te[nm])
line("Y", Y.float().cpu(), Yr, Ye)
