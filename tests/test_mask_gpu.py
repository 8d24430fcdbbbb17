"""GPU parity of the blend-mask synthesis (utils/inference/masks.py:39-108) through the C ABI:
bit-exact against the masks the reference's face_mask_static produced for the golden paste cases, against the CPU
oracle (oracle/mask_ref.py, itself pinned to cv2 and the reference) on random landmarks for every preset, and end to
end: get_final_image with a landmark handler (masks built on the device) reproduces the reference's golden frame."""
import os

import numpy as np
import pytest
import torch

import ghost_b200
from ghost_b200 import masks as gm, synth
from oracle import mask_ref as M
from oracle.gen_golden import PASTE_CASES, FakeHandler, paste_case_inputs
from test_oracle_mask import PRESETS, case_landmarks

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", PASTE_CASES, ids=[c[0] for c in PASTE_CASES])
def test_masks_match_reference_golden(case, golden_dir):
    gold = np.load(os.path.join(golden_dir, case[0] + ".npz"))
    pairs = case_landmarks(case)
    params = [gm.choose_params(lm, lt) for lm, lt in pairs]
    assert [list(p) for p in params] == [[int(v) for v in row] for row in gold["params"]]
    m64 = gm.face_masks(np.stack([lm for lm, _ in pairs]), params).cpu().numpy()
    m32 = gm.face_masks(np.stack([lm for lm, _ in pairs]), params, dtype=torch.float32).cpu().numpy()
    ref = gold["masks_u8"] / 255
    assert m64.dtype == np.float64 and np.array_equal(m64, ref)
    assert m32.dtype == np.float32 and np.array_equal(m32, ref.astype(np.float32))


def test_masks_match_oracle_all_presets():
    lms, prs = [], []
    for t in range(48):
        r = np.random.default_rng(500 + t)
        lm = synth.landmarks106(r, 224, shift=float(r.uniform(-5, 5)))
        if t % 6 == 0:
            lm = lm + r.uniform(-60, 60, size=(1, 2)).astype(np.float32)     # hull clipped by the crop border
        lms.append(lm)
        prs.append(list(PRESETS[t % 4]))
    got = gm.face_masks(np.stack(lms), prs).cpu().numpy()
    for t in range(48):
        ref = M.face_mask_static((224, 224), lms[t], lms[t], prs[t])
        assert np.array_equal(got[t], ref), "face %d preset %s: %d px differ" % (t, prs[t], int((got[t] != ref).sum()))


def test_face_mask_static_api():
    r = np.random.default_rng(3)
    lm, lt = synth.landmarks106(r, 224, shift=4.0), synth.landmarks106(r, 224, shift=0.0)
    img = np.zeros((224, 224, 3), np.uint8)
    m, p = ghost_b200.masks.face_mask_static(img, lm, lt, None)
    mr, pr = M.face_mask_static((224, 224), lm, lt, None)
    assert list(p) == list(pr) and np.array_equal(m, mr)
    m2 = ghost_b200.masks.face_mask_static(img, lm, lt, [10, 10, 8])
    assert np.array_equal(m2, M.face_mask_static((224, 224), lm, lt, [10, 10, 8]))
    with pytest.raises(ValueError):
        ghost_b200.masks.face_mask_static(np.zeros((256, 256, 3), np.uint8), lm, lt, None)
    with pytest.raises(ghost_b200._lib.GhostError):
        gm.face_masks(lm[None], [[0, 5, 5]])


@pytest.mark.parametrize("case", PASTE_CASES, ids=[c[0] for c in PASTE_CASES])
def test_get_final_image_with_handler_matches_reference_golden(case, golden_dir):
    """The reference's whole post-generator chain for one frame - cv2.resize, landmarks (fake handler), face_mask_static,
    warpAffine composite - with masks and composite on the device."""
    name, H, W, nf, seed, clipped, shift = case
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    frame, crops, targets, tfms = paste_case_inputs(*case)
    out = ghost_b200.get_final_image([[c] for c in crops], [[t] for t in targets], frame, [[t] for t in tfms],
                                     handler=FakeHandler(seed, shift))
    y0, y1, x0, x1 = gold["roi"]
    ref = frame.copy()
    ref[y0:y1, x0:x1] = gold["out_roi"]
    assert np.array_equal(out, ref), "%d pixels differ" % int((out != ref).any(axis=2).sum())
