"""CPU: the super-resolution oracle restatement (oracle/sr_ref.py) against the golden vectors recorded from the
reference's own LIPSPADEGenerator(TestOptions) in train() mode (tests/golden/sr_*.npz, oracle/gen_golden.py),
and against the live reference module where /root/reference is mounted."""
import hashlib
import os
import sys

import numpy as np
import pytest
import torch

from ghost_b200 import synth, weights as W
from oracle import sr_ref
from oracle.gen_golden import SR_CASES, sr_case_inputs


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("case", SR_CASES, ids=[c[0] for c in SR_CASES])
def test_oracle_matches_reference_golden(case, golden_dir):
    name, B, pert, calls = case
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    sd, crops = sr_case_inputs(name, B, pert)
    assert sha(crops) == str(gold["crops_sha"]), "input generator drifted"
    torch.set_num_threads(max(1, min(8, os.cpu_count() or 1)))
    state = {}
    x = sr_ref.preprocess_ref(crops)
    for _ in range(calls):
        y = sr_ref.sr_forward(sd, x, state)
    assert y.shape == (B, 3, 256, 256)
    ys = y.numpy().reshape(-1)[::97][:8192]
    assert np.abs(ys - gold["Y_sample"]).max() <= 5e-5      # same ops; summation order only
    assert np.abs(y.numpy() - gold["Y"].astype(np.float32)).max() <= 2e-3   # fp16-stored full tensor
    # spectral-norm state after `calls` power iterations
    assert np.abs(state["head_0.conv_0.weight_u"].numpy() - gold["u_head0_conv0"]).max() <= 1e-5
    assert np.abs(state["ups.3.conv_s.weight_u"].numpy() - gold["u_ups3_convs"]).max() <= 1e-5
    assert abs(float(y.abs().mean()) - float(gold["Y_absmean"])) <= 1e-5


def test_oracle_matches_live_reference(have_reference):
    if not have_reference:
        pytest.skip("/root/reference is not mounted on this box (goldens cover it)")
    sys.path.insert(0, "/root/reference")
    try:
        from models.config_sr import TestOptions
        from models.networks.generator import LIPSPADEGenerator
    finally:
        sys.path.remove("/root/reference")
    G = LIPSPADEGenerator(TestOptions())
    sd = W.init_state_dict(W.sr_spec(), seed=5)
    assert list(G.state_dict().keys()) == list(sd.keys())
    for k, v in G.state_dict().items():
        assert tuple(v.shape) == tuple(sd[k].shape), k
    G.load_state_dict(sd)
    G.train()
    crops = synth.crops_u8(1, seed=3)
    x = sr_ref.preprocess_ref(crops)
    state = {}
    with torch.no_grad():
        y1 = G(x)
        y2 = G(x)                      # second call: u has advanced
    r1 = sr_ref.sr_forward(sd, x, state)
    r2 = sr_ref.sr_forward(sd, x, state)
    assert (y1 - r1).abs().max().item() <= 1e-5
    assert (y2 - r2).abs().max().item() <= 1e-5
    assert (y1 - y2).abs().max().item() > 0          # the power iteration does change the output


def test_pre_post_conversions():
    crops = synth.crops_u8(2, seed=1)
    x = sr_ref.preprocess_ref(crops)
    assert x.shape == (2, 3, 256, 256) and float(x.max()) <= 1.0 and float(x.min()) >= 0.0
    assert np.array_equal((x[0, 0].numpy() * 255).round().astype(np.uint8), crops[0, :, :, 2])   # R plane = BGR[2]
    g = torch.tensor([[[[-0.5]], [[0.5]], [[1.5]]]])
    assert sr_ref.postprocess_ref(g).tolist() == [[[[255, 127, 0]]]]
