"""CPU: the AEI_Net oracle restatement (oracle/aei_ref.py) against the golden vectors recorded from the
reference's own network.AEI_Net.AEI_Net (tests/golden/aei_*.npz, made by oracle/gen_golden.py), and against
the live reference module where /root/reference is mounted."""
import hashlib
import os
import sys

import numpy as np
import pytest
import torch

from ghost_b200 import synth, weights as W
from oracle import aei_ref
from oracle.gen_golden import AEI_CASES, aei_case_inputs


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("case", AEI_CASES, ids=[c[0] for c in AEI_CASES])
def test_oracle_matches_reference_golden(case, golden_dir):
    name, bb, nb, pert, B = case
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    sd, Xt, z = aei_case_inputs(*case)
    assert sha(Xt.numpy()) == str(gold["xt_sha"]) and sha(z.numpy()) == str(gold["z_sha"]), "input generator drifted"
    Y, attrs = aei_ref.aei_forward(sd, Xt, z, bb, nb)
    assert Y.shape == (B, 3, 256, 256)
    assert np.abs(Y.numpy() - gold["Y"]).max() <= 2e-5   # same ops; only thread-count dependent summation order
    for k, a in enumerate(attrs, start=1):
        a = a.numpy()
        assert tuple(a.shape) == tuple(gold["attr%d_shape" % k])
        s = a.reshape(-1)[:: max(1, a.size // 4096)][:4096]
        assert np.abs(s - gold["attr%d_sample" % k]).max() <= 1e-4 * max(1.0, np.abs(s).max())


def test_oracle_matches_live_reference(have_reference):
    if not have_reference:
        pytest.skip("/root/reference is not mounted on this box (goldens cover it)")
    sys.path.insert(0, "/root/reference")
    try:
        from network.AEI_Net import AEI_Net
    finally:
        sys.path.remove("/root/reference")
    torch.manual_seed(3)
    G = AEI_Net("linknet", 1, 512).eval()
    sd = W.perturb_(G.state_dict(), seed=4)
    G.load_state_dict(sd)
    assert list(sd.keys()) == list(W.aei_spec("linknet", 1, 512).keys())
    Xt, z = synth.crops_pm1(1, seed=9), synth.z_ids(1, seed=9)
    with torch.no_grad():
        Y, attr = G(Xt, z.reshape(1, 1, 512))     # z_id may arrive as [B,1,512] (export-onnx.py:30)
    Y2, attr2 = aei_ref.aei_forward(sd, Xt, z, "linknet", 1)
    assert (Y - Y2).abs().max().item() <= 1e-6
    for a, b in zip(attr, attr2):
        assert (a - b).abs().max().item() <= 1e-6


def test_pre_post_conversions():
    u8 = synth.crops_u8(2, seed=1)
    x = aei_ref.transform_target_ref(u8, half=True)
    assert x.dtype == torch.float16 and tuple(x.shape) == (2, 3, 256, 256)
    assert x.min() >= -1 and x.max() <= 1
    # channel flip: output channel 0 (R) comes from BGR index 2
    assert abs(float(x[0, 0, 5, 7]) - (u8[0, 5, 7, 2] / 255. - 0.5) / 0.5) < 2e-3
    y = aei_ref.postprocess_ref(x.float())
    assert y.dtype == np.uint8 and y.shape == (2, 256, 256, 3)
    assert np.abs(y.astype(int) - u8.astype(int)).max() <= 1      # truncation after the fp16 round trip


def test_half_yardstick_matches_live_reference(have_reference, golden_dir):
    """tests/golden/aei_half_ref.npz (how far the reference's own G.half() forward lies from its fp32 forward) is
    recomputed from the unmodified module where /root/reference is mounted; everywhere: it covers every case."""
    rec = np.load(os.path.join(golden_dir, "aei_half_ref.npz"))
    assert sorted(rec.files) == sorted(c[0] for c in AEI_CASES)
    if not have_reference:
        pytest.skip("/root/reference is not mounted on this box (the recorded file covers it)")
    sys.path.insert(0, "/root/reference")
    try:
        from network.AEI_Net import AEI_Net
    finally:
        sys.path.remove("/root/reference")
    name, bb, nb, pert, B = AEI_CASES[0]
    sd, Xt, z = aei_case_inputs(*AEI_CASES[0])
    G = AEI_Net(bb, nb, 512).eval()
    G.load_state_dict(sd)
    with torch.no_grad():
        Y32, _ = G(Xt, z)
        Y16, _ = G.half()(Xt.half(), z.half())
    d = (Y16.float() - Y32).abs()
    assert abs(d.max().item() - rec[name][0]) <= 0.05 * rec[name][0] + 1e-4
    assert abs(d.mean().item() - rec[name][1]) <= 0.05 * rec[name][1] + 1e-6
    # the fp16 evaluation of the uint8 conversion (faceshifter_run.py:20-22 with a half Y) against the restatement
    ref_u8 = ((Y16.permute(0, 2, 3, 1) * 0.5 + 0.5) * 255)[:, :, :, [2, 1, 0]].type(torch.uint8).numpy()
    assert np.array_equal(aei_ref.postprocess_ref_half(Y16), ref_u8)
