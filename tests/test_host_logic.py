"""CPU: host-side logic of the operator mirror - state_dict layout, init rules, argument errors, sharding."""
import numpy as np
import pytest
import torch

import ghost_b200
from ghost_b200 import shard_range, synth, weights as W


def test_aei_param_counts_match_the_reference_survey():
    def nparams(spec):
        return sum(int(np.prod(s)) for k, (s, kind) in spec.items()
                   if not (k.endswith("running_mean") or k.endswith("running_var") or k.endswith("num_batches_tracked")))
    assert nparams(W.aei_spec("unet", 2, 512)) == 184338803 or abs(nparams(W.aei_spec("unet", 2, 512)) - 184.34e6) < 1e4
    assert abs(nparams(W.aei_spec("unet", 1, 512)) - 128.47e6) < 1e4
    assert abs(nparams(W.aei_spec("unet", 3, 512)) - 240.22e6) < 1e4
    assert abs(nparams(W.aei_spec("linknet", 3, 512)) - 216.80e6) < 1e4
    assert abs(nparams(W.iresnet100_spec()) - 65.16e6) < 1e4


def test_init_follows_weight_init_rules():
    sd = W.init_state_dict(W.aei_spec("linknet", 1, 512), seed=0)
    assert sd["generator.AADBlk1.add_blocks.0.fc1.weight"].std().item() == pytest.approx(0.001, rel=0.05)
    assert sd["generator.AADBlk1.add_blocks.0.fc1.bias"].abs().max().item() == 0
    w = sd["encoder.conv3.0.weight"]          # xavier_normal: std = sqrt(2 / (fan_in + fan_out))
    assert w.std().item() == pytest.approx((2.0 / ((64 + 128) * 16)) ** 0.5, rel=0.03)
    b = sd["generator.AADBlk8.add_blocks.0.conv1.bias"]   # default conv bias: U(+-1/sqrt(fan_in))
    assert b.abs().max().item() <= 1 / 32 ** 0.5 + 1e-6
    assert torch.equal(sd["encoder.conv1.1.running_var"], torch.ones(32))
    sd2 = W.init_state_dict(W.aei_spec("linknet", 1, 512), seed=0)
    assert all(torch.equal(sd[k], sd2[k]) for k in sd)


def test_module_surface_and_argument_errors_without_gpu():
    with pytest.raises(NotImplementedError):
        ghost_b200.AEI_Net("resnet", 2, 512)
    G = ghost_b200.AEI_Net("unet", 1, 512)
    assert G.eval() is G and G.half().dtype == torch.float16 and G.float().dtype == torch.float32
    with pytest.raises(NotImplementedError):
        G.train()
    with pytest.raises(ValueError):
        G.cpu()
    with pytest.raises(ValueError):        # CPU tensors are rejected: there is no CPU fallback
        G(torch.zeros(1, 3, 256, 256), torch.zeros(1, 512))
    with pytest.raises(TypeError):
        G(np.zeros((1, 3, 256, 256), np.float32), torch.zeros(1, 512))
    sd = G.state_dict()
    with pytest.raises(RuntimeError):
        G.load_state_dict({k: v for k, v in sd.items() if k != "generator.up1.weight"})
    with pytest.raises(RuntimeError):
        G.load_state_dict(dict(sd, extra=torch.zeros(1)))
    G.load_state_dict(dict(sd, extra=torch.zeros(1)), strict=False)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError):
            G.cuda()
        with pytest.raises(RuntimeError):
            ghost_b200.transform_target_to_torch(synth.crops_u8(1))
        with pytest.raises(RuntimeError):
            ghost_b200.paste_back(np.zeros((1, 8, 8, 3), np.uint8), np.zeros((0, 256, 256, 3), np.uint8),
                                  np.zeros((1, 224, 224)), np.zeros((0, 2, 3)), np.zeros(0, np.int32))
    with pytest.raises(ValueError):
        ghost_b200.get_final_image([[synth.crops_u8(1)[0]]], [[None]], np.zeros((8, 8, 3), np.uint8),
                                   [[np.eye(2, 3)]], handler=None, masks=None)
    empty = ghost_b200.get_final_image([], [], synth.frame_u8(16, 24), [], handler=None)
    assert np.array_equal(empty, synth.frame_u8(16, 24))
    with pytest.raises(ValueError):
        ghost_b200.iresnet100(pretrained=True)


@pytest.mark.parametrize("n,world", [(0, 1), (5, 1), (7, 2), (8, 8), (3, 8), (2048, 8), (1000, 3)])
def test_shard_range_partitions_in_order(n, world):
    spans = [shard_range(n, r, world) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    for (a, b), (c, d) in zip(spans, spans[1:]):
        assert b == c and a <= b
    sizes = [b - a for a, b in spans]
    assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def test_sharder_batches_keep_faces_of_a_frame_together():
    s = ghost_b200.FrameSharder(103, rank=1, world=4)
    got = list(s.batches(10, align=4))
    assert got[0][0] == s.lo and got[-1][1] == s.hi
    for (a, b), (c, d) in zip(got, got[1:]):
        assert b == c and (b - a) % 4 == 0


def test_synth_is_deterministic_and_in_range():
    a, b = synth.frame_u8(64, 96, seed=3), synth.frame_u8(64, 96, seed=3)
    assert a.flags["C_CONTIGUOUS"] and np.array_equal(a, b) and a.dtype == np.uint8
    rng = np.random.default_rng(0)
    for clipped in (False, True):
        t = synth.similarity_tfm(1080, 1920, rng, clipped=clipped)
        s = np.hypot(t[0, 0], t[1, 0])
        assert 0.25 <= s <= 0.6 + 1e-9
    m = synth.soft_mask(rng)
    assert m.shape == (224, 224) and m.min() == 0 and m.max() == 1 and m[0].max() == 0
