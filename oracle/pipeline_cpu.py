"""CPU pipeline of the hot path, used as the timed CPU baseline - TEST / BENCH INFRASTRUCTURE ONLY.

Runs, on the host cores, what the reference runs for one batch of swapped frames:
    z_id   = netArc(F.interpolate(normalize_and_torch_batch(src), 0.5, bilinear, align_corners=True))  core.py:44-45
    target = transform_target_to_torch(crops, half=False)                                              core.py:13-26
    Y, _   = G(target, z_id);  uint8 BGR conversion                                      faceshifter_run.py:19-22
    frame  = get_final_image(...)  (cv2.resize, cv2.invertAffineTransform, cv2.warpAffine x2, float64 blend)
                                                                                   image_processing.py:63-76
with the oracle restatements (oracle/aei_ref.py, oracle/iresnet_ref.py) standing in for the reference's torch
modules (they issue the same torch ops) and cv2 - the reference's own dependency - for the paste-back.
"""
import numpy as np
import torch

from . import aei_ref, iresnet_ref, paste_ref

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None


def paste_cv2(frame, crops256, masks, tfms):
    """get_final_image (image_processing.py:51-77) with the masks supplied."""
    if cv2 is None:
        return paste_ref.get_final_image_ref([[c] for c in crops256], masks, frame, [[t] for t in tfms])
    final = frame.copy()
    H, W = frame.shape[:2]
    for crop, mask, tfm in zip(crops256, masks, tfms):
        sw = cv2.resize(crop, (224, 224))
        mat_rev = cv2.invertAffineTransform(tfm)
        swap_t = cv2.warpAffine(sw, mat_rev, (W, H), borderMode=cv2.BORDER_REPLICATE)
        mask_t = np.expand_dims(cv2.warpAffine(mask, mat_rev, (W, H)), 2)
        final = mask_t * swap_t + (1 - mask_t) * final
    return np.array(final, dtype="uint8")


def swap_frames_cpu(aei_sd, arc_sd, backbone, num_blocks, sources_u8, crops_u8, frames, masks, tfms,
                    face_frame=None, face_source=None, return_swapped=False):
    """One batch: sources [S,224,224,3] RGB u8, crops [B,256,256,3] BGR u8, frames [F,H,W,3].
    face_frame[i] (non-decreasing) = frame of face i, face_source[i] = source identity of face i; the defaults
    are one face per frame with its own source (BASELINE configs[1-2]); configs[3] has 4 faces per frame, whose
    faces compose in list order exactly as the loop of get_final_image does (image_processing.py:62-75)."""
    n = len(crops_u8)
    face_frame = np.arange(n) if face_frame is None else np.asarray(face_frame)
    face_source = np.arange(n) if face_source is None else np.asarray(face_source)
    z = iresnet_ref.embed_hwc_ref(arc_sd, sources_u8)[torch.from_numpy(face_source).long()]
    target = aei_ref.transform_target_ref(crops_u8, half=False)
    Y, _ = aei_ref.aei_forward(aei_sd, target, z, backbone, num_blocks)
    out_u8 = aei_ref.postprocess_ref(Y)
    res = []
    for f in range(len(frames)):
        sel = [i for i in range(n) if face_frame[i] == f]
        res.append(paste_cv2(frames[f], [out_u8[i] for i in sel], [masks[i] for i in sel], [tfms[i] for i in sel]))
    res = np.stack(res)
    return (res, out_u8) if return_swapped else res


def swap_frames_sr_cpu(aei_sd, arc_sd, sr_sd, backbone, num_blocks, sources_u8, crops_u8, frames, masks, tfms,
                       sr_state=None):
    """BASELINE configs[4] on the host cores: the swap above, then face_enhancement (one DataLoader batch of <= 20
    crops, video_processing.py:263-285) on the swapped crops, then the paste-back of the enhanced crops."""
    from . import sr_ref
    z = iresnet_ref.embed_hwc_ref(arc_sd, sources_u8)
    target = aei_ref.transform_target_ref(crops_u8, half=False)
    Y, _ = aei_ref.aei_forward(aei_sd, target, z, backbone, num_blocks)
    out_u8 = aei_ref.postprocess_ref(Y)
    with torch.no_grad():
        enh = sr_ref.face_enhancement_ref({k: v.float() for k, v in sr_sd.items() if v.dtype.is_floating_point},
                                          out_u8, sr_state)
    return np.stack([paste_cv2(frames[i], [enh[i]], [masks[i]], [tfms[i]]) for i in range(len(frames))])
