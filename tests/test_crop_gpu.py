"""GPU parity of the crop extraction through the C ABI: bit-exact against the CPU oracle
(oracle/crop_ref.py, pinned to cv2) - frames of several sizes, faces clipped by the frame border, several faces per
frame - and against cv2 itself where it is installed; plus the frame -> crop -> paste-back round trip."""
import numpy as np
import pytest
import torch

import ghost_b200
from ghost_b200 import synth
from oracle import crop_ref as Cr

pytestmark = pytest.mark.gpu


def _case(F, H, W, per_frame, seed):
    rng = np.random.default_rng(seed)
    frames = np.stack([synth.frame_u8(H, W, seed=seed + i) for i in range(F)])
    tfms, ff = [], []
    for f in range(F):
        for j in range(per_frame):
            tfms.append(synth.similarity_tfm(H, W, rng, clipped=((f + j) % 3 == 0)))
            ff.append(f)
    return frames, np.stack(tfms), np.array(ff, np.int32)


@pytest.mark.parametrize("shape", [(2, 270, 480, 1), (3, 540, 960, 2), (2, 1080, 1920, 4), (1, 333, 517, 2)])
def test_crops_match_oracle(shape):
    F, H, W, per = shape
    frames, tfms, ff = _case(F, H, W, per, seed=30 + H)
    c224, c256 = ghost_b200.crop_faces(torch.from_numpy(frames).cuda(), tfms, ff)
    r224, r256 = Cr.crop_faces_ref(frames, tfms, ff)
    assert np.array_equal(c224.cpu().numpy(), r224), "%d px differ" % int((c224.cpu().numpy() != r224).sum())
    assert np.array_equal(c256.cpu().numpy(), r256), "%d px differ" % int((c256.cpu().numpy() != r256).sum())
    only256 = ghost_b200.crop_faces(frames, tfms, ff, want224=False)[1]
    assert np.array_equal(only256.cpu().numpy(), r256)


def test_crops_match_cv2():
    cv2 = pytest.importorskip("cv2")
    frames, tfms, ff = _case(2, 720, 1280, 3, seed=77)
    c224, c256 = ghost_b200.crop_faces(frames, tfms, ff)
    for i in range(len(ff)):
        ref = cv2.warpAffine(frames[ff[i]], tfms[i], (224, 224), borderValue=0.0)
        assert np.array_equal(c224[i].cpu().numpy(), ref)
        assert np.array_equal(c256[i].cpu().numpy(), cv2.resize(ref, (256, 256)))


def test_crop_then_paste_round_trip():
    """Pasting the unmodified crop back with an all-ones mask must leave the frame within the double-bilinear
    smoothing error, and a zero mask must leave it bit-identical (size-independent property, 1080p)."""
    frames, tfms, ff = _case(2, 1080, 1920, 1, seed=5)
    d_frames = torch.from_numpy(frames).cuda()
    _, c256 = ghost_b200.crop_faces(d_frames, tfms, ff)
    zero = np.zeros((len(ff), 224, 224), np.float64)
    out0 = ghost_b200.paste_back(d_frames, c256, zero, tfms, ff, mode="exact").cpu().numpy()
    assert np.array_equal(out0, frames)
    soft = np.stack([synth.soft_mask(np.random.default_rng(i)) for i in range(len(ff))])
    out1 = ghost_b200.paste_back(d_frames, c256, soft, tfms, ff, mode="exact").cpu().numpy().astype(int)
    d = np.abs(out1 - frames.astype(int))
    assert d.mean() < 3.0, d.mean()        # same content resampled twice: small, smooth differences only


def test_errors():
    frames, tfms, ff = _case(1, 270, 480, 1, seed=1)
    with pytest.raises(ghost_b200._lib.GhostError):
        ghost_b200.crop_faces(frames, tfms, np.array([3], np.int32))
    with pytest.raises(ValueError):
        ghost_b200.crop_faces(frames[0], tfms, ff)
