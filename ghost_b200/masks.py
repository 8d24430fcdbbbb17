"""Host-side blend-mask synthesis from 106-point landmarks (still on the host).

Mirrors utils/inference/masks.py:5-108 of the reference (expand_eyebrows, get_mask, face_mask_static,
erode_and_blur) with cv2 on the CPU, because landmarks come from a landmark model that is out of scope.
The result (float64 224x224 in [0,1]) feeds the paste-back kernel.
"""
import numpy as np

_BROW_L, _EYE_L = [43, 48, 49, 51, 50], [35, 41, 40, 42, 39]
_BROW_R, _EYE_R = [102, 103, 104, 105, 101], [89, 95, 94, 96, 93]

# offset thresholds -> (erode, sigmaX, sigmaY)   (masks.py:47-64)
_PRESETS = ((6, (15, 15, 10)), (3, (10, 10, 8)))


def _lift_eyebrows(lm, mod):
    pts = np.array(lm.copy(), dtype=np.int32)
    for brow, eye in ((_BROW_L, _EYE_L), (_BROW_R, _EYE_R)):
        top, bot = pts[brow], pts[eye]
        pts[brow] = top + mod * 0.5 * (top - bot)
    return pts


def _hull_mask(shape_hw, pts):
    import cv2
    mask = np.zeros(shape_hw, dtype=np.uint8)
    cv2.fillConvexPoly(mask, cv2.convexHull(np.array(pts, np.int32)), 255)
    return mask


def erode_and_blur(mask_input, erode, sigmaX, sigmaY, fade_to_border=True):
    import cv2
    mask = np.copy(mask_input)
    k = np.ones((abs(erode), abs(erode)), "uint8")
    mask = cv2.erode(mask, k, iterations=1) if erode > 0 else cv2.dilate(mask, k, iterations=1)
    if fade_to_border:
        c = sigmaY * 2
        mask[:c, :] = 0
        mask[-c:, :] = 0
        mask[:, :c] = 0
        mask[:, -c:] = 0
    return cv2.GaussianBlur(mask, (0, 0), sigmaX=sigmaX, sigmaY=sigmaY)


def face_mask_static(image, landmarks, landmarks_tgt, params=None):
    """Returns (mask, params) when params is None, else mask - as the reference does."""
    if params is None:
        left = np.sum((landmarks[1][0] - landmarks_tgt[1][0], landmarks[2][0] - landmarks_tgt[2][0],
                       landmarks[13][0] - landmarks_tgt[13][0]))
        right = np.sum((landmarks_tgt[17][0] - landmarks[17][0], landmarks_tgt[18][0] - landmarks[18][0],
                        landmarks_tgt[29][0] - landmarks[29][0]))
        offset = max(left, right)
        chosen = (5, 5, 5)
        if offset < -3:
            chosen = (-5, 5, 10)
        for th, preset in reversed(_PRESETS):
            if offset > th:
                chosen = preset
        erode, sx, sy = chosen
    else:
        erode, sx, sy = params[0], params[1], params[2]
    mod = 2.7 if erode == 15 else (0.5 if erode == -5 else 2.0)
    pts = _lift_eyebrows(landmarks, mod)
    mask = erode_and_blur(_hull_mask(image.shape[:2], pts), erode, sx, sy, True)
    if params is None:
        return mask / 255, [erode, sx, sy]
    return mask / 255
