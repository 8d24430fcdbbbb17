"""How far is the reference's own fp32 forward from the same arithmetic in fp64? (the fp32 noise floor)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import aei_ref
from oracle.gen_golden import AEI_CASES, aei_case_inputs
for case in AEI_CASES:
    name, bb, nb, pert, B = case
    sd, Xt, z = aei_case_inputs(*case)
    Y32, _ = aei_ref.aei_forward(sd, Xt, z, bb, nb)
    with torch.no_grad():
        sd64 = {k: v.double() for k, v in sd.items() if v.dtype.is_floating_point}
        a64 = aei_ref.encode(sd64, Xt.double(), bb)
        Y64 = aei_ref.generate(sd64, a64, z.double(), nb)
    d = (Y32.double() - Y64).abs()
    print("%-26s fp32-vs-fp64 max %.3e mean %.3e" % (name, d.max(), d.mean()), flush=True)
