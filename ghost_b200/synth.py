"""Deterministic synthetic inputs for the tests and the bench: aligned 256x256 crops,
z_id vectors, 1080p-style uint8 frames, frame->224-crop similarity transforms, 106-point landmark
sets and soft blend masks.  numpy / torch only - no reference code, no dataset, no network."""
import math

import numpy as np
import torch


def crops_pm1(B, seed=0):
    """Xt [B,3,256,256] fp32 RGB in [-1,1]."""
    g = torch.Generator().manual_seed(1000 + seed)
    return torch.rand(B, 3, 256, 256, generator=g) * 2 - 1


def z_ids(B, c_id=512, seed=0):
    """ArcFace-like identity vectors: ~N(0,1) per dimension after BN1d (|z| ~ 22)."""
    g = torch.Generator().manual_seed(2000 + seed)
    return torch.randn(B, c_id, generator=g)


def crops_u8(B, seed=0):
    """uint8 BGR [B,256,256,3] crops with smooth + sharp content."""
    rng = np.random.default_rng(3000 + seed)
    out = np.empty((B, 256, 256, 3), np.uint8)
    for b in range(B):
        out[b] = frame_u8(256, 256, seed=int(rng.integers(1 << 30)), cell=8)
    return out


def sources_u8(B, seed=0):
    """uint8 RGB [B,224,224,3] source faces for the ArcFace embedding."""
    rng = np.random.default_rng(4000 + seed)
    out = np.empty((B, 224, 224, 3), np.uint8)
    for b in range(B):
        out[b] = frame_u8(224, 224, seed=int(rng.integers(1 << 30)), cell=8)
    return out


def frame_u8(H, W, seed=0, cell=16):
    """Low-frequency random field (bilinearly upsampled `cell`x) plus fine noise, uint8 BGR [H,W,3]."""
    rng = np.random.default_rng(5000 + seed)
    gh, gw = H // cell + 2, W // cell + 2
    coarse = rng.uniform(20, 235, size=(gh, gw, 3))
    ys = (np.arange(H) + 0.5) / cell
    xs = (np.arange(W) + 0.5) / cell
    y0 = np.floor(ys).astype(np.int64)
    x0 = np.floor(xs).astype(np.int64)
    fy = (ys - y0)[:, None, None]
    fx = (xs - x0)[None, :, None]
    a = coarse[y0][:, x0]
    b = coarse[y0][:, x0 + 1]
    c = coarse[y0 + 1][:, x0]
    d = coarse[y0 + 1][:, x0 + 1]
    img = (1 - fy) * ((1 - fx) * a + fx * b) + fy * ((1 - fx) * c + fx * d)
    img += rng.integers(-12, 13, size=(H, W, 3))
    # a few hard edges so the bilinear paths see sharp content too
    for _ in range(4):
        y, x = int(rng.integers(0, H)), int(rng.integers(0, W))
        h, w = int(rng.integers(8, max(9, H // 4))), int(rng.integers(8, max(9, W // 4)))
        img[y:y + h, x:x + w] = rng.uniform(0, 255, size=3)
    return np.ascontiguousarray(np.clip(np.rint(img), 0, 255).astype(np.uint8))


def similarity_tfm(H, W, rng, clipped=False, scale=None, center=None):
    """2x3 float64 frame->224-crop affine: crop = s*R(theta)*frame + t.  The inverse-mapped 224^2 square lies
    inside the frame, except with clipped=True where its centre sits near the frame border."""
    s_hi = 0.6
    s_lo = max(0.25, 224.0 * math.sqrt(2.0) / (min(H, W) - 8) + 1e-3)
    s = float(rng.uniform(min(s_lo, s_hi), s_hi)) if scale is None else float(scale)
    th = math.radians(float(rng.uniform(-20.0, 20.0)))
    half = 112.0 * math.sqrt(2.0) / s + 2.0   # half-diagonal of the square in frame pixels
    if center is not None:
        cx, cy = float(center[0]), float(center[1])
    elif clipped:
        cx = float(rng.choice([half * 0.3, W - half * 0.3]))
        cy = float(rng.uniform(half * 0.3, H - half * 0.3))
    else:
        cx = float(rng.uniform(half, max(half + 1e-6, W - half)))
        cy = float(rng.uniform(half, max(half + 1e-6, H - half)))
    a, b = s * math.cos(th), s * math.sin(th)
    return np.array([[a, -b, 112.0 - (a * cx - b * cy)],
                     [b, a, 112.0 - (b * cx + a * cy)]], dtype=np.float64)


def soft_mask(rng=None, size=224):
    """float64 [224,224] blend mask in [0,1]: an ellipse with a smooth shoulder, zero at the crop border
    (the shape face_mask_static produces, without needing landmarks)."""
    rng = rng or np.random.default_rng(0)
    cy, cx = size / 2 + rng.uniform(-6, 6), size / 2 + rng.uniform(-6, 6)
    ry, rx = size * rng.uniform(0.30, 0.38), size * rng.uniform(0.24, 0.32)
    y, x = np.mgrid[0:size, 0:size].astype(np.float64)
    r = np.sqrt(((y - cy) / ry) ** 2 + ((x - cx) / rx) ** 2)
    m = np.clip((1.15 - r) / 0.3, 0.0, 1.0)
    m = m * m * (3 - 2 * m)
    m[:10] = 0
    m[-10:] = 0
    m[:, :10] = 0
    m[:, -10:] = 0
    return np.rint(m * 255) / 255     # same value set as the reference's uint8 mask / 255


def landmarks106(rng=None, size=224, shift=0.0):
    """Synthetic 106-point landmark set on a face-like ellipse inside a `size` crop (float32 [106,2])."""
    rng = rng or np.random.default_rng(0)
    cx, cy = size / 2 + shift, size / 2 + 4
    pts = np.zeros((106, 2), np.float32)
    # 0..32: jaw line, lower half of an ellipse from the left temple to the right temple
    t = np.linspace(math.pi * 0.98, math.pi * 0.02, 33)
    pts[:33, 0] = cx + 70 * np.cos(t)
    pts[:33, 1] = cy + 85 * np.sin(t) * 0.9 - 10
    # the rest: interior points (eyes, brows, nose, mouth) scattered inside the upper ellipse
    ang = rng.uniform(0, 2 * math.pi, 73)
    rad = np.sqrt(rng.uniform(0, 1, 73))
    pts[33:, 0] = cx + 55 * rad * np.cos(ang)
    pts[33:, 1] = cy - 15 + 55 * rad * np.sin(ang)
    # eyebrow rows sit above the eye rows so expand_eyebrows pushes them outwards
    for brow, eye, sx in (([43, 48, 49, 51, 50], [35, 41, 40, 42, 39], -1), ([102, 103, 104, 105, 101],
                                                                            [89, 95, 94, 96, 93], 1)):
        xs = cx + sx * 30 + np.linspace(-18, 18, 5)
        pts[brow, 0] = xs
        pts[brow, 1] = cy - 45
        pts[eye, 0] = xs
        pts[eye, 1] = cy - 30
    return pts + rng.uniform(-0.5, 0.5, size=pts.shape).astype(np.float32)
