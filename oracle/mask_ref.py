"""CPU oracle for the blend-mask synthesis - TEST INFRASTRUCTURE ONLY (never imported by ghost_b200/).

Integer / fixed-point restatement, in plain Python + numpy, of what the reference computes with OpenCV in
    utils/inference/masks.py:5-20   expand_eyebrows
    utils/inference/masks.py:23-36  get_mask            (cv2.convexHull + cv2.fillConvexPoly)
    utils/inference/masks.py:39-85  face_mask_static    (preset choice, /255)
    utils/inference/masks.py:88-108 erode_and_blur      (cv2.erode / cv2.dilate, border clip, cv2.GaussianBlur)
The arithmetic lives in OpenCV (opencv-python, un-pinned in requirements.txt; checked here against cv2 4.13.0):
    imgproc/src/convhull.cpp     convexHull            -> hull()        (any exact convex hull gives the same fill)
    imgproc/src/drawing.cpp      FillConvexPoly, Line, LineIterator, clipLine -> fill_convex_poly()
    imgproc/src/morph.dispatch.cpp  erode / dilate with a k x k box, anchor k/2, border ignored -> box_morph()
    imgproc/src/smooth.dispatch.cpp GaussianBlur CV_8U: bit-exact fixed-point kernels (8 fractional bits, error
                                 diffusion, sum == 256), 8.8 horizontal pass, 16.16 vertical pass, round half up,
                                 BORDER_REFLECT_101 -> gaussian_kernel_fixed(), gaussian_blur_u8()

Pinning: tests/test_oracle_mask.py checks every stage bit for bit against cv2 itself and the whole function against
the reference's own face_mask_static (imported from /root/reference where mounted) and the masks recorded in
tests/golden/paste_*.npz.
"""
import math

import numpy as np

_BROW_L, _EYE_L = [43, 48, 49, 51, 50], [35, 41, 40, 42, 39]
_BROW_R, _EYE_R = [102, 103, 104, 105, 101], [89, 95, 94, 96, 93]
XY_SHIFT = 16
XY_ONE = 1 << XY_SHIFT


def expand_eyebrows(lmrks, mod):
    """masks.py:5-20 (int32 truncation on both casts)."""
    pts = np.array(np.asarray(lmrks).copy(), dtype=np.int32)
    for brow, eye in ((_BROW_L, _EYE_L), (_BROW_R, _EYE_R)):
        top, bot = pts[brow], pts[eye]
        pts[brow] = top + mod * 0.5 * (top - bot)
    return pts


def choose_params(landmarks, landmarks_tgt):
    """masks.py:43-64."""
    left = np.sum((landmarks[1][0] - landmarks_tgt[1][0], landmarks[2][0] - landmarks_tgt[2][0],
                   landmarks[13][0] - landmarks_tgt[13][0]))
    right = np.sum((landmarks_tgt[17][0] - landmarks[17][0], landmarks_tgt[18][0] - landmarks[18][0],
                    landmarks_tgt[29][0] - landmarks[29][0]))
    offset = max(left, right)
    if offset > 6:
        return 15, 15, 10
    if offset > 3:
        return 10, 10, 8
    if offset < -3:
        return -5, 5, 10
    return 5, 5, 5


def hull(points):
    """Convex hull of integer points without collinear vertices (Andrew's monotone chain, exact integer cross products)."""
    pts = sorted(set((int(x), int(y)) for x, y in points))
    if len(pts) <= 2:
        return pts

    def cross(o, a, b):
        return (a[0] - o[0]) * (b[1] - o[1]) - (a[1] - o[1]) * (b[0] - o[0])
    lower, upper = [], []
    for p in pts:
        while len(lower) >= 2 and cross(lower[-2], lower[-1], p) <= 0:
            lower.pop()
        lower.append(p)
    for p in reversed(pts):
        while len(upper) >= 2 and cross(upper[-2], upper[-1], p) <= 0:
            upper.pop()
        upper.append(p)
    return lower[:-1] + upper[:-1]


def _clip_line(w, h, x1, y1, x2, y2):
    """cv::clipLine (drawing.cpp): returns (inside, x1, y1, x2, y2)."""
    right, bottom = w - 1, h - 1
    c1 = (x1 < 0) + (x1 > right) * 2 + (y1 < 0) * 4 + (y1 > bottom) * 8
    c2 = (x2 < 0) + (x2 > right) * 2 + (y2 < 0) * 4 + (y2 > bottom) * 8
    if (c1 & c2) == 0 and (c1 | c2) != 0:
        if c1 & 12:
            a = 0 if c1 < 8 else bottom
            x1 += int(float(a - y1) * (x2 - x1) / (y2 - y1))
            y1 = a
            c1 = (x1 < 0) + (x1 > right) * 2
        if c2 & 12:
            a = 0 if c2 < 8 else bottom
            x2 += int(float(a - y2) * (x2 - x1) / (y2 - y1))
            y2 = a
            c2 = (x2 < 0) + (x2 > right) * 2
        if (c1 & c2) == 0 and (c1 | c2) != 0:
            if c1:
                a = 0 if c1 == 1 else right
                y1 += int(float(a - x1) * (y2 - y1) / (x2 - x1))
                x1 = a
                c1 = 0
            if c2:
                a = 0 if c2 == 1 else right
                y2 += int(float(a - x2) * (y2 - y1) / (x2 - x1))
                x2 = a
                c2 = 0
    return (c1 | c2) == 0, x1, y1, x2, y2


def line_pixels(w, h, p1, p2):
    """cv::Line with connectivity 8 = LineIterator(img, pt1, pt2, 8, leftToRight=true) after clipLine."""
    ok, x1, y1, x2, y2 = _clip_line(w, h, p1[0], p1[1], p2[0], p2[1])
    if not ok:
        return []
    dx, dy = x2 - x1, y2 - y1
    if dx < 0:                      # leftToRight: start from the left end point
        x1, y1, dx, dy = x2, y2, -dx, -dy
    sy = 1 if dy >= 0 else -1
    dy = abs(dy)
    out = []
    if dy > dx:                     # y is the major axis
        err, plus, minus = dy - 2 * dx, 2 * dy, -2 * dx
        x, y = x1, y1
        for _ in range(dy + 1):
            out.append((x, y))
            m = err < 0
            err += minus + (plus if m else 0)
            y += sy
            if m:
                x += 1
    else:
        err, plus, minus = dx - 2 * dy, 2 * dx, -2 * dy
        x, y = x1, y1
        for _ in range(dx + 1):
            out.append((x, y))
            m = err < 0
            err += minus + (plus if m else 0)
            x += 1
            if m:
                y += sy
    return out


def fill_convex_poly(w, h, v, value=255):
    """cv::FillConvexPoly (drawing.cpp, line_type 8, shift 0) on a zero image: the outline (cv::Line around the
    polygon, each edge clipped to the image first) plus the scan-line fill between the two edge chains."""
    img = np.zeros((h, w), dtype=np.uint8)
    n = len(v)
    if n == 0:
        return img
    for i in range(n):
        for x, y in line_pixels(w, h, v[i - 1], v[i]):
            img[y, x] = value
    xs, ys = [p[0] for p in v], [p[1] for p in v]
    ymin, ymax, xmin, xmax = min(ys), max(ys), min(xs), max(xs)
    imin = ys.index(ymin)
    if n < 3 or xmax < 0 or ymax < 0 or xmin >= w or ymin >= h:
        return img
    ymax = min(ymax, h - 1)
    edges = n
    e_idx, e_di = [imin, imin], [1, n - 1]
    e_x, e_dx, e_ye = [-XY_ONE, -XY_ONE], [0, 0], [ymin, ymin]
    y = ymin
    delta1 = delta2 = XY_ONE >> 1
    while True:
        for i in range(2):
            if y >= e_ye[i]:
                idx0, di = e_idx[i], e_di[i]
                idx = idx0 + di
                if idx >= n:
                    idx -= n
                while True:
                    edges -= 1
                    if edges < 0:
                        break
                    ty = ys[idx]
                    if ty > y:
                        x_s, x_e = xs[idx0] << XY_SHIFT, xs[idx] << XY_SHIFT
                        e_ye[i] = ty
                        num, den = (x_e - x_s) * 2 + (ty - y), 2 * (ty - y)
                        q = abs(num) // den                     # C++ integer division truncates toward zero
                        e_dx[i] = q if num >= 0 else -q
                        e_x[i] = x_s
                        e_idx[i] = idx
                        break
                    idx0 = idx
                    idx += di
                    if idx >= n:
                        idx -= n
        if edges < 0:
            break
        if y >= 0:
            left, right = (1, 0) if e_x[0] > e_x[1] else (0, 1)
            xx1 = (e_x[left] + delta1) >> XY_SHIFT
            xx2 = (e_x[right] + delta2) >> XY_SHIFT
            if xx2 >= 0 and xx1 < w:
                xx1, xx2 = max(xx1, 0), min(xx2, w - 1)
                if xx1 <= xx2:
                    img[y, xx1:xx2 + 1] = value
        e_x[0] += e_dx[0]
        e_x[1] += e_dx[1]
        y += 1
        if y > ymax:
            break
    return img


def box_morph(img, k, erode=True):
    """cv2.erode / cv2.dilate with np.ones((k, k)): window [p - k//2, p - k//2 + k - 1], pixels outside ignored."""
    h, w = img.shape
    a = k // 2
    pad_val = 255 if erode else 0
    p = np.full((h + k, w + k), pad_val, dtype=np.uint8)
    p[a:a + h, a:a + w] = img
    out = np.full((h, w), pad_val, dtype=np.uint8)
    f = np.minimum if erode else np.maximum
    for dy in range(k):
        for dx in range(k):
            out = f(out, p[dy:dy + h, dx:dx + w])
    return out


def gaussian_kernel_fixed(sigma):
    """getGaussianKernelBitExact + getGaussianKernelFixedPoint_ED (8 fractional bits): odd n = round(6 sigma + 1) | 1."""
    n = int(np.rint(sigma * 3 * 2 + 1)) | 1
    n2 = (n - 1) // 2
    scale2x = -0.125 / (sigma * sigma)
    vals, s, x = [], 0.0, 1 - n
    for _ in range(n2):
        t = math.exp(float(x * x) * scale2x)
        vals.append(t)
        s += t
        x += 2
    s = s * 2.0 + 1.0
    mul1 = 1.0 / s
    res, err, tot = [0] * n, 0.0, 0
    for i in range(n2):
        adj = vals[i] * mul1 * 256.0 + err
        v0 = int(np.rint(adj))
        err = adj - float(v0)
        res[i] = res[n - 1 - i] = v0
        tot += v0
    res[n2] = 256 - 2 * tot
    return np.array(res, dtype=np.int64)


def _reflect101(i, n):
    while i < 0 or i >= n:
        i = -i if i < 0 else 2 * (n - 1) - i
    return i


def gaussian_blur_u8(src, sigma_x, sigma_y):
    """cv2.GaussianBlur(src_u8, (0, 0), sigmaX, sigmaY), BORDER_REFLECT_101, fixed point."""
    kx, ky = gaussian_kernel_fixed(sigma_x), gaussian_kernel_fixed(sigma_y)
    h, w = src.shape
    rx, ry = len(kx) // 2, len(ky) // 2
    s = src.astype(np.int64)
    ix = np.array([[_reflect101(x + k - rx, w) for k in range(len(kx))] for x in range(w)])
    hor = np.zeros((h, w), dtype=np.int64)
    for k in range(len(kx)):
        hor += kx[k] * s[:, ix[:, k]]
    iy = np.array([[_reflect101(y + k - ry, h) for k in range(len(ky))] for y in range(h)])
    ver = np.zeros((h, w), dtype=np.int64)
    for k in range(len(ky)):
        ver += ky[k] * hor[iy[:, k], :]
    return ((ver + 32768) >> 16).astype(np.uint8)


def erode_and_blur(mask, erode, sigma_x, sigma_y, fade_to_border=True):
    """masks.py:88-108."""
    m = box_morph(mask, abs(erode), erode > 0)
    if fade_to_border:
        c = sigma_y * 2
        m[:c, :] = 0
        m[-c:, :] = 0
        m[:, :c] = 0
        m[:, -c:] = 0
    return gaussian_blur_u8(m, sigma_x, sigma_y)


def hull_mask(shape_hw, pts):
    return fill_convex_poly(shape_hw[1], shape_hw[0], hull(pts))


def face_mask_static(image_hw, landmarks, landmarks_tgt, params=None):
    """masks.py:39-85: returns (mask float64 in [0,1], [erode, sigmaX, sigmaY]) when params is None, else the mask."""
    erode, sx, sy = choose_params(landmarks, landmarks_tgt) if params is None else (params[0], params[1], params[2])
    mod = 2.7 if erode == 15 else (0.5 if erode == -5 else 2.0)
    pts = expand_eyebrows(landmarks, mod)
    m = erode_and_blur(hull_mask(image_hw, pts), erode, sx, sy, True) / 255
    return (m, [erode, sx, sy]) if params is None else m
