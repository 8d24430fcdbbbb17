"""CPU: the crop-extraction oracle (oracle/crop_ref.py) bit for bit against cv2.warpAffine(borderValue=0) and
cv2.resize in both directions (224 -> 256 is the one with the border-row subtlety)."""
import numpy as np
import pytest

from ghost_b200 import synth
from oracle import crop_ref as Cr

cv2 = pytest.importorskip("cv2")


def test_warp_affine_const_border_matches_cv2():
    rng = np.random.default_rng(0)
    for t in range(10):
        H, W = (270, 480) if t % 2 else (540, 960)
        frame = synth.frame_u8(H, W, seed=t)
        M = synth.similarity_tfm(H, W, rng, clipped=(t % 3 == 0))
        ref = cv2.warpAffine(frame, M, (224, 224), borderValue=0.0)
        assert np.array_equal(ref, Cr.warp_affine_u8_const0(frame, M, 224, 224)), t


def test_resize_matches_cv2_both_directions():
    rng = np.random.default_rng(1)
    for _ in range(3):
        a = rng.integers(0, 256, (224, 224, 3), dtype=np.uint8)
        b = rng.integers(0, 256, (256, 256, 3), dtype=np.uint8)
        assert np.array_equal(cv2.resize(a, (256, 256)), Cr.resize_linear_u8(a, 256))
        assert np.array_equal(cv2.resize(b, (224, 224)), Cr.resize_linear_u8(b, 224))
