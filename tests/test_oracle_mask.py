"""CPU: the blend-mask oracle (oracle/mask_ref.py) bit for bit against cv2 itself, against the reference's own
face_mask_static (where /root/reference is mounted) and against the masks the reference produced for the golden
paste-back cases (tests/golden/paste_*.npz); and the host half of the product path (hull + cv::FillConvexPoly
rasterisation + fixed-point Gaussian kernels inside libghost_b200.so, no device work) against the oracle."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from ghost_b200 import _lib, synth
from oracle import mask_ref as M
from oracle.gen_golden import PASTE_CASES, FakeHandler, paste_case_inputs

cv2 = pytest.importorskip("cv2")
PRESETS = [(15, 15, 10), (10, 10, 8), (-5, 5, 10), (5, 5, 5)]


def test_gaussian_blur_matches_cv2():
    rng = np.random.default_rng(0)
    for e, sx, sy in PRESETS:
        for trial in range(2):
            src = rng.integers(0, 256, (224, 224), dtype=np.uint8) if trial == 0 else np.zeros((224, 224), np.uint8)
            if trial:
                cv2.circle(src, (100, 120), 70, 255, -1)
            ref = cv2.GaussianBlur(src, (0, 0), sigmaX=sx, sigmaY=sy)
            assert np.array_equal(ref, M.gaussian_blur_u8(src, sx, sy)), (sx, sy, trial)
        k = M.gaussian_kernel_fixed(sx)
        assert int(k.sum()) == 256 and len(k) == (int(np.rint(sx * 6 + 1)) | 1)


def test_box_morph_matches_cv2():
    img = np.zeros((224, 224), np.uint8)
    cv2.circle(img, (110, 115), 70, 255, -1)
    img[0:20, 50:80] = 255
    img[200:224, 100:224] = 255
    for k in (15, 10, 5):
        assert np.array_equal(cv2.erode(img, np.ones((k, k), "uint8"), iterations=1), M.box_morph(img, k, True))
        assert np.array_equal(cv2.dilate(img, np.ones((k, k), "uint8"), iterations=1), M.box_morph(img, k, False))


def test_fill_convex_poly_matches_cv2():
    rng = np.random.default_rng(1)
    for t in range(300):
        n = int(rng.integers(3, 40))
        lo, hi = ((-60, 284), (5, 219), (0, 224))[t % 3]        # partly outside, inside, touching the border
        pts = rng.integers(lo, hi, size=(n, 2))
        ref = np.zeros((224, 224), np.uint8)
        cv2.fillConvexPoly(ref, cv2.convexHull(pts.astype(np.int32)), 255)
        assert np.array_equal(ref, M.fill_convex_poly(224, 224, M.hull(pts))), t


def case_landmarks(case):
    """The landmark pairs (swap, target) the reference saw for one golden paste case (FakeHandler call sequence)."""
    name, H, W, nf, seed, clipped, shift = case
    frame, crops, targets, tfms = paste_case_inputs(*case)
    h = FakeHandler(seed, shift)
    out = []
    for i in range(nf):
        sw = cv2.resize(crops[i], (224, 224))
        out.append((h.get_without_detection_without_transform(sw), h.get_without_detection_without_transform(targets[i])))
    return out


@pytest.mark.parametrize("case", PASTE_CASES, ids=[c[0] for c in PASTE_CASES])
def test_oracle_matches_reference_golden_masks(case, golden_dir):
    gold = np.load(os.path.join(golden_dir, case[0] + ".npz"))
    for i, (lm, lt) in enumerate(case_landmarks(case)):
        m, p = M.face_mask_static((224, 224), lm, lt, None)
        assert list(p) == [int(v) for v in gold["params"][i]]
        assert np.array_equal(np.rint(m * 255).astype(np.uint8), gold["masks_u8"][i])
        assert np.array_equal(m, gold["masks_u8"][i] / 255)


def test_oracle_matches_live_reference(have_reference):
    if not have_reference:
        pytest.skip("/root/reference is not mounted on this box (goldens cover it)")
    sys.path.insert(0, "/root/reference")
    try:
        from utils.inference import masks as RM
    finally:
        sys.path.remove("/root/reference")
    img = np.zeros((224, 224, 3), np.uint8)
    seen = set()
    for t in range(60):
        r = np.random.default_rng(t)
        lm = synth.landmarks106(r, 224, shift=float(r.uniform(-5, 5)))
        lt = synth.landmarks106(r, 224, shift=0.0)
        ref, p = RM.face_mask_static(img, lm, lt, None)
        got, p2 = M.face_mask_static((224, 224), lm, lt, None)
        assert list(p) == list(p2) and np.array_equal(ref, got), t
        seen.add(tuple(p))
        for params in PRESETS[:2]:
            assert np.array_equal(RM.face_mask_static(img, lm, lt, list(params)),
                                  M.face_mask_static((224, 224), lm, lt, list(params)))
    assert len(seen) >= 3, "the synthetic landmarks should reach several presets: %s" % (seen,)


def host_job(lm, e, sx, sy):
    lib = _lib.load()
    bm = (C.c_uint32 * (224 * 7))()
    kx, ky, nn = (C.c_uint16 * 128)(), (C.c_uint16 * 128)(), (C.c_int * 2)()
    lmf = np.ascontiguousarray(np.asarray(lm, np.float32).reshape(106, 2))
    rc = lib.ghost_test_mask_job(lmf.ctypes.data_as(C.POINTER(C.c_float)), e, sx, sy, bm, kx, ky, nn)
    assert rc == 0, lib.ghost_test_last_error()
    words = np.frombuffer(bm, dtype=np.uint32).reshape(224, 7)
    bits = ((words[:, :, None] >> np.arange(32, dtype=np.uint32)) & 1).reshape(224, 224).astype(np.uint8) * 255
    return bits, np.array(kx[:nn[0]], dtype=np.int64), np.array(ky[:nn[1]], dtype=np.int64)


def test_host_rasteriser_matches_oracle():
    """ghost_test_mask_job (the host half of ghost_face_masks) against the oracle: hull bitmap and Gaussian kernels."""
    for t in range(80):
        r = np.random.default_rng(100 + t)
        lm = synth.landmarks106(r, 224, shift=float(r.uniform(-5, 5)))
        if t % 5 == 0:                                   # faces hanging over the crop border
            lm = lm + r.uniform(-70, 70, size=(1, 2)).astype(np.float32)
        e, sx, sy = PRESETS[t % 4]
        bits, kx, ky = host_job(lm, e, sx, sy)
        mod = 2.7 if e == 15 else (0.5 if e == -5 else 2.0)
        ref = M.hull_mask((224, 224), M.expand_eyebrows(lm, mod))
        assert np.array_equal(bits, ref), t
        assert np.array_equal(kx, M.gaussian_kernel_fixed(sx)) and np.array_equal(ky, M.gaussian_kernel_fixed(sy))
