"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py

Imports /root/reference (network.AEI_Net.AEI_Net, utils.inference.image_processing.get_final_image,
utils.inference.masks.face_mask_static) on the CPU and records their outputs on seeded inputs and
seeded weights, so that the GPU box - where /root/reference does not exist - can still check the
oracle restatements (oracle/aei_ref.py, oracle/paste_ref.py) and the CUDA path against the reference.
Inputs and weights are regenerated from the seeds by the tests (ghost_b200/synth.py, ghost_b200/weights.py);
the files hold only the reference's outputs plus input checksums.
"""
import hashlib
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.dont_write_bytecode = True
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

from ghost_b200 import synth, weights as W  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


AEI_CASES = [  # name, backbone, num_blocks, perturbed, batch
    ("aei_unet2_default", "unet", 2, False, 1),
    ("aei_unet2_perturbed", "unet", 2, True, 2),
    ("aei_linknet3_perturbed", "linknet", 3, True, 1),
    ("aei_unet1_perturbed", "unet", 1, True, 1),
    ("aei_linknet3_default", "linknet", 3, False, 1),
    ("aei_unet1_default", "unet", 1, False, 2),
]


def aei_case_inputs(name, backbone, nb, perturbed, B):
    sd = W.init_state_dict(W.aei_spec(backbone, nb, 512), seed=0)
    if perturbed:
        W.perturb_(sd, seed=1)
    return sd, synth.crops_pm1(B, seed=7), synth.z_ids(B, seed=7)


def gen_aei(only_missing=True):
    from network.AEI_Net import AEI_Net
    for name, bb, nb, pert, B in AEI_CASES:
        if only_missing and os.path.exists(os.path.join(GOLD, name + ".npz")):
            continue   # recorded in an earlier round: left byte-identical
        sd, Xt, z = aei_case_inputs(name, bb, nb, pert, B)
        G = AEI_Net(bb, nb, 512).eval()
        G.load_state_dict(sd)
        with torch.no_grad():
            Y, attr = G(Xt, z)
        out = {"Y": Y.numpy().astype(np.float32), "xt_sha": sha(Xt.numpy()), "z_sha": sha(z.numpy())}
        for k, a in enumerate(attr):
            a = a.numpy()
            out["attr%d_sample" % (k + 1)] = a.reshape(-1)[:: max(1, a.size // 4096)][:4096].astype(np.float32)
            out["attr%d_absmean" % (k + 1)] = np.float64(np.abs(a).mean())
            out["attr%d_shape" % (k + 1)] = np.array(a.shape)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
        print(name, "Y mean|.|", float(np.abs(out["Y"]).mean()))


def gen_aei_half():
    """The reference as it is DEPLOYED (inference.py:26-30: G.eval() ... .half()): the unmodified module in fp16 on
    the same inputs and weights, and how far its Y lies from its own fp32 forward.  That distance is the yardstick
    for the 16-bit tensor-core modes on the ill-conditioned ("perturbed") weight sets, where no 16-bit operand
    format can hold 1e-2 (tools/sens16.py): tests/golden/aei_half_ref.npz = per case [max-abs, mean-abs]."""
    from network.AEI_Net import AEI_Net
    out = {}
    for name, bb, nb, pert, B in AEI_CASES:
        sd, Xt, z = aei_case_inputs(name, bb, nb, pert, B)
        G = AEI_Net(bb, nb, 512).eval()
        G.load_state_dict(sd)
        with torch.no_grad():
            Y32, _ = G(Xt, z)
            Y16, _ = G.half()(Xt.half(), z.half())
        d = (Y16.float() - Y32).abs()
        out[name] = np.array([d.max().item(), d.mean().item()], dtype=np.float64)
        print(name, "reference .half() vs its fp32 forward: max %.3e mean %.3e" % (out[name][0], out[name][1]))
    np.savez_compressed(os.path.join(GOLD, "aei_half_ref.npz"), **out)


class FakeHandler(object):
    """Stands in for coordinate_reg.image_infer.Handler (landmark model, out of scope): returns seeded
    synthetic landmarks; the swap and the target crop get slightly different sets like the real model."""

    def __init__(self, seed, shift):
        self.seed, self.shift, self.calls = seed, shift, 0

    def get_without_detection_without_transform(self, img):
        self.calls += 1
        rng = np.random.default_rng(self.seed + self.calls % 2)
        return synth.landmarks106(rng, 224, shift=self.shift if self.calls % 2 == 1 else 0.0)


PASTE_CASES = [  # name, H, W, n_faces, seed, clipped, landmark shift (selects the erode/sigma preset)
    ("paste_small_1face", 270, 480, 1, 11, False, 0.0),
    ("paste_small_clipped", 270, 480, 1, 12, True, 3.0),
    ("paste_mid_3faces", 540, 960, 3, 13, False, -2.0),
    ("paste_1080p_1face", 1080, 1920, 1, 14, False, 1.5),
    ("paste_1080p_4faces_overlap", 1080, 1920, 4, 15, False, 0.0),
]


def paste_case_inputs(name, H, W, nf, seed, clipped, shift):
    rng = np.random.default_rng(seed)
    frame = synth.frame_u8(H, W, seed=seed)
    crops = synth.crops_u8(nf, seed=seed)
    targets = synth.crops_u8(nf, seed=seed + 100)[:, :224, :224].copy()
    tfms = []
    for i in range(nf):
        t = synth.similarity_tfm(H, W, rng, clipped=clipped)
        if "overlap" in name and i > 0:      # pile later faces onto the first one
            t = tfms[0].copy()
            t[:, 2] += rng.uniform(-40, 40, size=2)
        tfms.append(t)
    return frame, crops, targets, np.stack(tfms)


def gen_paste():
    for m in ("matplotlib", "matplotlib.pyplot", "insightface", "insightface.utils", "insightface.utils.face_align"):
        sys.modules.setdefault(m, types.ModuleType(m))
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["insightface"].utils = sys.modules["insightface.utils"]
    sys.modules["insightface.utils"].face_align = sys.modules["insightface.utils.face_align"]
    from utils.inference.image_processing import get_final_image
    from utils.inference import masks as ref_masks
    import cv2
    for name, H, W, nf, seed, clipped, shift in PASTE_CASES:
        frame, crops, targets, tfms = paste_case_inputs(name, H, W, nf, seed, clipped, shift)
        handler = FakeHandler(seed, shift)
        final_frames = [[crops[i]] for i in range(nf)]
        crop_frames = [[targets[i]] for i in range(nf)]
        tfm_arrays = [[tfms[i]] for i in range(nf)]
        out = get_final_image(final_frames, crop_frames, frame, tfm_arrays, handler)
        # the masks the reference built on the way (same handler call sequence)
        h2 = FakeHandler(seed, shift)
        masks_u8, params = [], []
        for i in range(nf):
            sw = cv2.resize(crops[i], (224, 224))
            lm = h2.get_without_detection_without_transform(sw)
            lt = h2.get_without_detection_without_transform(targets[i])
            m, p = ref_masks.face_mask_static(targets[i], lm, lt, None)
            masks_u8.append(np.rint(m * 255).astype(np.uint8))
            assert np.array_equal(masks_u8[-1] / 255, m)
            params.append(p)
        diff = np.any(out != frame, axis=2)
        ys, xs = np.nonzero(diff)
        y0, y1, x0, x1 = (int(ys.min()), int(ys.max()) + 1, int(xs.min()), int(xs.max()) + 1) if len(ys) else (0, 0, 0, 0)
        rec = {"masks_u8": np.stack(masks_u8), "params": np.array(params), "roi": np.array([y0, y1, x0, x1]),
               "out_roi": out[y0:y1, x0:x1].copy(), "out_sha": sha(out), "frame_sha": sha(frame),
               "crops_sha": sha(crops), "tfm_sha": sha(tfms)}
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **rec)
        print(name, "roi", rec["roi"], "params", params, "changed px", int(diff.sum()))


SR_CASES = [  # name, batch, perturbed LIP gates, number of consecutive forwards (spectral-norm state advances)
    ("sr_b2_default", 2, False, 1),
    ("sr_b3_perturbed_2calls", 3, True, 2),
]


def sr_case_inputs(name, B, perturbed):
    sd = W.init_state_dict(W.sr_spec(), seed=0)
    if perturbed:
        W.perturb_sr_(sd, seed=1)
    return sd, synth.crops_u8(B, seed=21)


def gen_sr():
    """LIPSPADEGenerator(TestOptions) of the reference, .train() as inference.py:48, fed like face_enhancement."""
    from models.config_sr import TestOptions
    from models.networks.generator import LIPSPADEGenerator
    for name, B, pert, calls in SR_CASES:
        sd, crops = sr_case_inputs(name, B, pert)
        G = LIPSPADEGenerator(TestOptions())
        G.load_state_dict(sd)
        G.train()
        x = torch.from_numpy(np.ascontiguousarray(crops[..., ::-1])).permute(0, 3, 1, 2).float().div(255)
        with torch.no_grad():
            for _ in range(calls):
                y = G(x)
        u8 = torch.clamp(y * 255, 0, 255).type(torch.uint8).permute(0, 2, 3, 1).numpy()[..., ::-1]
        out = {"Y": y.numpy().astype(np.float16), "Y_absmean": np.float64(y.abs().mean()),
               "Y_sample": y.numpy().reshape(-1)[::97][:8192].astype(np.float32), "u8_sha": sha(u8),
               "crops_sha": sha(crops), "u_head0_conv0": G.head_0.conv_0.weight_u.numpy().copy(),
               "u_ups3_convs": G.ups[3].conv_s.weight_u.numpy().copy()}
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
        print(name, "Y mean|.|", float(out["Y_absmean"]))


def align_case_tracks(seed=0, n_frames=40):
    """Synthetic 5-point landmark tracks (float32 [5,2] per frame, as the detector returns them) for two targets:
    smooth drift + jitter, a scene cut (a jump of ~40 px) at frame 17, a slow pan that crosses the 5 px threshold on
    the nose only at frame 30, missing detections (empty lists) at frames 9 and 25-26 of target 1, and a
    one-frame run between two cuts."""
    rng = np.random.default_rng(seed)
    base = np.array([[80, 100], [150, 98], [116, 140], [88, 175], [144, 173]], np.float32)
    tracks = []
    for q in range(2):
        tr = []
        off = np.array([300.0 + 500 * q, 200.0 + 100 * q])
        for i in range(n_frames):
            off = off + rng.uniform(-1.5, 1.5, 2)
            if i == 17:
                off = off + np.array([40.0, -25.0])
            if i == 21 and q == 0:
                off = off + np.array([9.0, 0.0])
            if i == 22 and q == 0:
                off = off - np.array([9.0, 0.0])
            k = (base * (1.0 + 0.2 * q) + off + rng.uniform(-0.7, 0.7, (5, 2))).astype(np.float32)
            if i == 30:
                k[2] += np.float32(6.0)
            tr.append(k)
        tracks.append(tr)
    for i in (9, 25, 26):
        tracks[1][i] = []
    return tracks


def gen_align():
    """smooth_landmarks of the unmodified reference (utils/inference/video_processing.py:86-108) on the tracks above."""
    for m in ("matplotlib", "matplotlib.pyplot", "insightface", "insightface.utils", "insightface.utils.face_align",
              "kornia", "kornia.geometry", "kornia.geometry.transform"):
        sys.modules.setdefault(m, types.ModuleType(m))
    sys.modules["insightface"].utils = sys.modules["insightface.utils"]
    sys.modules["insightface.utils"].face_align = sys.modules["insightface.utils.face_align"]
    from utils.inference.video_processing import smooth_landmarks
    tracks = align_case_tracks()
    out = smooth_landmarks(tracks, n=2)
    rec = {}
    for q, row in enumerate(out):
        present = np.array([len(k) != 0 for k in row])
        rec["present%d" % q] = present
        rec["smooth%d" % q] = np.stack([np.asarray(k) for k in row if len(k) != 0])
        rec["dtype%d" % q] = np.array(str(np.asarray([k for k in row if len(k) != 0][0]).dtype))
    np.savez_compressed(os.path.join(GOLD, "align_smooth.npz"), **rec)
    print("align_smooth", {k: v.shape for k, v in rec.items()})


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    which = sys.argv[1:] or ["aei", "paste", "sr", "align"]
    if "align" in which:
        gen_align()
    if "sr" in which:
        gen_sr()
    if "aei" in which:
        gen_aei()
        gen_aei_half()
    if "paste" in which:
        gen_paste()
