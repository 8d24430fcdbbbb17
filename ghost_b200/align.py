"""Host glue between a landmark detector's output and the crop extraction (host half):

    smooth_landmarks(kps_arr, n=2)                      utils/inference/video_processing.py:86-108
    estimate_norm(lmk, image_size, mode)                insightface.utils.face_align.estimate_norm as called at
                                                        video_processing.py:133,162 and image_processing.py:18
    landmarks_to_crops(full_frames, kps_array, size)    video_processing.py:157-168 (second half of
                                                        crop_frames_and_get_transforms)

The detector itself (`app.get`) and the video codecs are out of scope; the 5-point landmark tracks come from the
caller.  Everything here is small float arithmetic per face (a 2x2 SVD, a mean over <= 5 frames), so it stays on
the host in numpy, exactly where the reference has it; the pixels (cv2.warpAffine of every crop) go through the
device kernel behind `crop_faces`.

`estimate_norm` restates a dependency the reference does not vendor (insightface, un-pinned in requirements.txt,
on top of scikit-image's SimilarityTransform.estimate = Umeyama's closed form): PARITY UNPINNED for that function -
neither package is in this image; the tests check it through its defining properties instead.
"""
import numpy as np

# The five 112x112 pose templates of insightface.utils.face_align (left profile, left, frontal, right, right profile);
# mode != 'arcface' picks the best of the five, src_map[224] = 2 x the 112 templates.
_SRC_112 = np.array([
    [[51.642, 50.115], [57.617, 49.990], [35.740, 69.007], [51.157, 89.050], [57.025, 89.702]],
    [[45.031, 50.118], [65.568, 50.872], [39.677, 68.111], [45.177, 86.190], [64.246, 86.758]],
    [[39.730, 51.138], [72.270, 51.138], [56.000, 68.493], [42.463, 87.010], [69.537, 87.010]],
    [[46.845, 50.872], [67.382, 50.118], [72.737, 68.111], [48.167, 86.758], [67.236, 86.190]],
    [[54.796, 49.990], [60.771, 50.115], [76.673, 69.007], [55.388, 89.702], [61.257, 89.050]]], dtype=np.float32)
_ARCFACE_112 = np.array([[[38.2946, 51.6963], [73.5318, 51.5014], [56.0252, 71.7366], [41.5493, 92.3655],
                          [70.7299, 92.2041]]], dtype=np.float32)
_SRC_MAP = {112: _SRC_112, 224: _SRC_112 * 2}


def _missing(k):
    return len(k) == 0


def smooth_landmarks(kps_arr, n=2):
    """video_processing.py:86-108.  kps_arr[q][i]: the 5x2 landmarks of target q in frame i, or an empty list when
    the face was not found.  A track is cut wherever a detection is missing on either side or the first / third
    landmark (an eye, the nose) jumps by more than 5 px between neighbouring frames (a scene cut); inside a run,
    frame i becomes the mean over frames i-q..i+q with q = min(n, distance to the nearer end of the run).
    Returns lists shaped like the input (missing entries stay as they were)."""
    result = []
    for track in kps_arr:
        # run boundaries: starts[k] is the first frame of run k
        starts = [0]
        for i in range(1, len(track)):
            a, b = track[i], track[i - 1]
            cut = _missing(a) or _missing(b)
            if not cut:
                cut = (np.linalg.norm(np.asarray(a[0]) - np.asarray(b[0])) > 5 or
                       np.linalg.norm(np.asarray(a[2]) - np.asarray(b[2])) > 5)
            if cut:
                starts.append(i)
        ends = starts[1:] + [len(track)]
        smooth = []
        for lo, hi in zip(starts, ends):
            for i in range(lo, hi):
                q = min(i - lo, hi - 1 - i, n)
                smooth.append(np.mean(np.array(track[i - q:i + 1 + q]), axis=0))
        result.append(smooth)
    return result


def _umeyama(src, dst):
    """Least-squares similarity transform dst ~ s R src + t (Umeyama 1991), float64: the closed form behind
    skimage.transform.SimilarityTransform.estimate.  Returns the 3x3 homogeneous matrix."""
    src = np.asarray(src, dtype=np.float64)
    dst = np.asarray(dst, dtype=np.float64)
    num, dim = src.shape
    src_mean, dst_mean = src.mean(axis=0), dst.mean(axis=0)
    sc, dc = src - src_mean, dst - dst_mean
    A = dc.T @ sc / num
    d = np.ones((dim,), dtype=np.float64)
    if np.linalg.det(A) < 0:
        d[dim - 1] = -1
    T = np.eye(dim + 1, dtype=np.float64)
    U, S, V = np.linalg.svd(A)
    rank = np.linalg.matrix_rank(A)
    if rank == 0:
        return np.nan * T
    if rank == dim - 1:
        if np.linalg.det(U) * np.linalg.det(V) > 0:
            T[:dim, :dim] = U @ V
        else:
            s = d[dim - 1]
            d[dim - 1] = -1
            T[:dim, :dim] = U @ np.diag(d) @ V
            d[dim - 1] = s
    else:
        T[:dim, :dim] = U @ np.diag(d) @ V
    scale = 1.0 / sc.var(axis=0).sum() * (S @ d)
    T[:dim, dim] = dst_mean - scale * (T[:dim, :dim] @ src_mean.T)
    T[:dim, :dim] *= scale
    return T


def estimate_norm(lmk, image_size=112, mode="arcface"):
    """insightface.utils.face_align.estimate_norm: the 2x3 frame -> aligned-crop similarity for five landmarks
    (eyes, nose, mouth corners).  mode='arcface' fits the single ArcFace template scaled to image_size; any other
    mode (the reference passes 'None') tries the five pose templates of size 112 / 224 and keeps the one with the
    smallest summed landmark distance.  Returns (M float64 [2,3], template index)."""
    lmk = np.asarray(lmk)
    if lmk.shape != (5, 2):
        raise AssertionError("lmk must be [5,2], got %s" % (lmk.shape,))
    if mode == "arcface":
        src = _ARCFACE_112 if image_size == 112 else float(image_size) / 112 * _ARCFACE_112
    else:
        if image_size not in _SRC_MAP:
            raise KeyError(image_size)
        src = _SRC_MAP[image_size]
    lmk_h = np.insert(lmk, 2, values=np.ones(5), axis=1)
    best, best_i, best_err = [], [], float("inf")
    for i in range(src.shape[0]):
        M = _umeyama(lmk, src[i])[0:2, :]
        res = (M @ lmk_h.T).T
        err = np.sum(np.sqrt(np.sum((res - src[i]) ** 2, axis=1)))
        if err < best_err:
            best, best_i, best_err = M, i, err
    return best, best_i


def landmarks_to_transforms(kps_array, crop_size=224, n=2):
    """smooth_landmarks + estimate_norm per (target, frame) (video_processing.py:157-162,166-168): returns
    tfm_array[q][i] = M [2,3] float64, or [] where the face is missing / the fit fails."""
    smooth = smooth_landmarks(kps_array, n=n)
    tfm_array = []
    for q in range(len(smooth)):
        row = []
        for k in smooth[q]:
            try:
                M, _ = estimate_norm(k, crop_size, mode="None")
                row.append(M if np.all(np.isfinite(M)) else [])
            except Exception:   # the reference's bare except: an empty / malformed landmark set -> no crop
                row.append([])
        tfm_array.append(row)
    return tfm_array


def landmarks_to_crops(full_frames, kps_array, crop_size=224, n=2):
    """Second half of crop_frames_and_get_transforms (video_processing.py:155-171): smoothed landmarks -> affines ->
    `cv2.warpAffine(frame, M, (crop_size, crop_size), borderValue=0.0)` for every (target, frame), the warps on the
    device (ghost_crop_faces, bit-exact against OpenCV).  Returns (crop_frames, tfm_array) shaped like the
    reference's: lists per target over frames, [] where the target has no face in a frame."""
    if crop_size != 224:
        raise ValueError("the device crop kernel is built for crop_size 224 (the reference's default)")
    from .api import crop_faces
    tfm_array = landmarks_to_transforms(kps_array, crop_size, n)
    frames = np.stack([np.asarray(f, dtype=np.uint8) for f in full_frames])
    crop_frames = [[[] for _ in row] for row in tfm_array]
    where = [(q, i) for q, row in enumerate(tfm_array) for i, M in enumerate(row) if len(M)]
    if where:
        tf = np.stack([tfm_array[q][i] for q, i in where])
        order = np.argsort([i for _, i in where], kind="stable")            # crop_faces wants frame order
        c224, _ = crop_faces(frames, tf[order], np.asarray([where[k][1] for k in order], np.int32), want256=False)
        c224 = c224.cpu().numpy()
        for slot, k in enumerate(order):
            q, i = where[k]
            crop_frames[q][i] = c224[slot]
    return crop_frames, tfm_array
