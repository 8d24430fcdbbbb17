"""CPU: the paste-back oracle (oracle/paste_ref.py) against (i) the golden outputs recorded from the reference's
own get_final_image (tests/golden/paste_*.npz), (ii) cv2 itself (resize / warpAffine, both border modes) and
(iii) the host-side mask port against the masks the reference built."""
import hashlib
import os

import numpy as np
import pytest

from ghost_b200 import masks as gmasks, synth
from oracle import paste_ref
from oracle.gen_golden import PASTE_CASES, FakeHandler, paste_case_inputs

cv2 = pytest.importorskip("cv2")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("case", PASTE_CASES, ids=[c[0] for c in PASTE_CASES])
def test_oracle_matches_reference_golden(case, golden_dir):
    name, H, W, nf, seed, clipped, shift = case
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    frame, crops, targets, tfms = paste_case_inputs(*case)
    assert sha(frame) == str(gold["frame_sha"]) and sha(crops) == str(gold["crops_sha"])
    assert sha(tfms) == str(gold["tfm_sha"]), "input generator drifted"
    masks = [m / 255 for m in gold["masks_u8"]]
    out = paste_ref.get_final_image_ref([[c] for c in crops], masks, frame, [[t] for t in tfms])
    assert sha(out) == str(gold["out_sha"])
    y0, y1, x0, x1 = gold["roi"]
    assert np.array_equal(out[y0:y1, x0:x1], gold["out_roi"])


@pytest.mark.parametrize("case", PASTE_CASES[:3], ids=[c[0] for c in PASTE_CASES[:3]])
def test_mask_preset_choice_matches_reference(case, golden_dir):
    """Host logic of ghost_b200.masks (the preset face_mask_static picks, masks.py:43-64); the masks themselves are
    built on the device (tests/test_mask_gpu.py) and their oracle is pinned in tests/test_oracle_mask.py."""
    name, H, W, nf, seed, clipped, shift = case
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    frame, crops, targets, tfms = paste_case_inputs(*case)
    h = FakeHandler(seed, shift)
    for i in range(nf):
        sw = cv2.resize(crops[i], (224, 224))
        lm = h.get_without_detection_without_transform(sw)
        lt = h.get_without_detection_without_transform(targets[i])
        assert list(gmasks.choose_params(lm, lt)) == list(gold["params"][i])


def test_resize_bit_exact_vs_cv2():
    for s in range(3):
        img = synth.crops_u8(1, seed=20 + s)[0]
        assert np.array_equal(paste_ref.resize_linear_u8(img, (224, 224)), cv2.resize(img, (224, 224)))


def test_invert_affine_vs_cv2():
    rng = np.random.default_rng(0)
    for _ in range(10):
        t = synth.similarity_tfm(1080, 1920, rng)
        assert np.array_equal(paste_ref.invert_affine(t), cv2.invertAffineTransform(t))


@pytest.mark.parametrize("clipped", [False, True])
def test_warp_affine_bit_exact_vs_cv2(clipped):
    rng = np.random.default_rng(5 + clipped)
    H, W = 540, 960
    for _ in range(3):
        t = synth.similarity_tfm(H, W, rng, clipped=clipped)
        mrev = cv2.invertAffineTransform(t)
        src = synth.crops_u8(1, seed=int(rng.integers(1000)))[0][:224, :224]
        ref = cv2.warpAffine(src, mrev, (W, H), borderMode=cv2.BORDER_REPLICATE)
        assert np.array_equal(paste_ref.warp_affine_u8_replicate(src, mrev, W, H), ref)
        mask = synth.soft_mask(rng)
        refm = cv2.warpAffine(mask, mrev, (W, H))
        assert np.array_equal(paste_ref.warp_affine_f64_const0(mask, mrev, W, H), refm)


def test_empty_face_list_is_identity():
    frame = synth.frame_u8(64, 96, seed=1)
    assert np.array_equal(paste_ref.get_final_image_ref([], [], frame, []), frame)


def test_zero_mask_is_identity_and_full_mask_replaces():
    rng = np.random.default_rng(2)
    H, W = 300, 400
    frame = synth.frame_u8(H, W, seed=2)
    crop = synth.crops_u8(1, seed=3)
    t = synth.similarity_tfm(H, W, rng)
    out0 = paste_ref.get_final_image_ref([[crop[0]]], [np.zeros((224, 224))], frame, [[t]])
    assert np.array_equal(out0, frame)
    out1 = paste_ref.get_final_image_ref([[crop[0]]], [np.ones((224, 224))], frame, [[t]])
    sw = cv2.warpAffine(cv2.resize(crop[0], (224, 224)), cv2.invertAffineTransform(t), (W, H),
                        borderMode=cv2.BORDER_REPLICATE)
    mt = cv2.warpAffine(np.ones((224, 224)), cv2.invertAffineTransform(t), (W, H))
    inside = mt == 1.0
    assert inside.sum() > 1000 and np.array_equal(out1[inside], sw[inside])
