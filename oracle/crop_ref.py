"""CPU oracle for the crop extraction in front of the generator - TEST INFRASTRUCTURE ONLY.

Integer restatement of the two OpenCV calls the reference makes per (frame, face):
    utils/inference/video_processing.py:159-168  align_img = cv2.warpAffine(frame, M, (224, 224), borderValue=0.0)
    utils/inference/image_processing.py:13-20    (same call for single images)
    utils/inference/video_processing.py:174-188  resize_frames: cv2.resize(crop, (256, 256))
warpAffine: the fixed-point coordinate model of oracle/paste_ref.py (AB_BITS 10, INTER_BITS 5, 15-bit weight table)
with BORDER_CONSTANT 0 - taps outside the frame contribute the border value.  cv2.resize INTER_LINEAR uint8: 11-bit
coefficients; horizontally the index AND the fraction are clamped at the borders, vertically only the two row indices
are clamped and the fraction is kept (that is what makes 224 -> 256 differ from the naive model in the first and last
output row; 256 -> 224 never reaches the clamp).  Pinned bit for bit against cv2 in tests/test_oracle_crop.py.
"""
import numpy as np

from oracle import paste_ref as P


def warp_affine_u8_const0(src, M, W, H):
    """cv2.warpAffine(src_u8, M, (W, H)) - INTER_LINEAR, BORDER_CONSTANT, borderValue 0."""
    sh, sw = src.shape[:2]
    sx, sy, ax, ay = P.warp_coords(M, W, H)
    _, tabi = P.bilinear_tables()
    wgt = tabi[ay, ax]
    s = src.astype(np.int64)
    if s.ndim == 2:
        s = s[..., None]

    def tap(yy, xx):
        ok = (yy >= 0) & (yy < sh) & (xx >= 0) & (xx < sw)
        v = s[np.clip(yy, 0, sh - 1), np.clip(xx, 0, sw - 1)]
        return np.where(ok[..., None], v, 0)
    acc = (tap(sy, sx) * wgt[..., 0:1] + tap(sy, sx + 1) * wgt[..., 1:2] +
           tap(sy + 1, sx) * wgt[..., 2:3] + tap(sy + 1, sx + 1) * wgt[..., 3:4])
    out = (acc + (1 << (P.INTER_REMAP_COEF_BITS - 1))) >> P.INTER_REMAP_COEF_BITS
    out = np.clip(out, 0, 255).astype(np.uint8)
    return out if src.ndim == 3 else out[..., 0]


def resize_tables(s, d):
    """(x index, x a0, x a1, y row0, y row1, y b0, y b1) of cv2.resize INTER_LINEAR uint8 for s -> d (square)."""
    scale = s / d
    xi = np.zeros(d, np.int64)
    xa0, xa1, y0, y1, yb0, yb1 = (np.zeros(d, np.int64) for _ in range(6))
    for i in range(d):
        f = np.float32((i + 0.5) * scale - 0.5)
        si = int(np.floor(f))
        fr = np.float32(f - np.float32(si))
        b0 = int(np.rint(np.float32((np.float32(1) - fr) * np.float32(2048))))
        b1 = int(np.rint(np.float32(fr * np.float32(2048))))
        y0[i], y1[i] = min(max(si, 0), s - 1), min(max(si + 1, 0), s - 1)
        yb0[i], yb1[i] = b0, b1
        if si < 0:
            si, fr = 0, np.float32(0)
        if si >= s - 1:
            si, fr = s - 1, np.float32(0)
        xi[i] = si
        xa0[i] = int(np.rint(np.float32((np.float32(1) - fr) * np.float32(2048))))
        xa1[i] = int(np.rint(np.float32(fr * np.float32(2048))))
    return xi, xa0, xa1, y0, y1, yb0, yb1


def resize_linear_u8(img, d):
    """cv2.resize(img_u8, (d, d)) of a square image, INTER_LINEAR."""
    s = img.shape[0]
    xi, xa0, xa1, y0, y1, yb0, yb1 = resize_tables(s, d)
    v = img.astype(np.int64)
    xi1 = np.minimum(xi + 1, s - 1)
    rows = v[:, xi, :] * xa0[None, :, None] + v[:, xi1, :] * xa1[None, :, None]
    r0, r1 = rows[y0], rows[y1]
    return ((((yb0[:, None, None] * (r0 >> 4)) >> 16) + ((yb1[:, None, None] * (r1 >> 4)) >> 16) + 2) >> 2).astype(np.uint8)


def crop_faces_ref(frames, tfms, face_frame):
    """(crops224, crops256) for every face: warpAffine(frame, M, (224,224), borderValue=0) then resize to 256."""
    c224 = np.stack([warp_affine_u8_const0(frames[f], np.asarray(M, np.float64), 224, 224)
                     for M, f in zip(tfms, face_frame)])
    c256 = np.stack([resize_linear_u8(c, 256) for c in c224])
    return c224, c256
