"""CPU oracle for the pix2pix super-resolution generator - TEST INFRASTRUCTURE ONLY (never imported by ghost_b200/).

A functional, state_dict-driven restatement (plain torch fp32 on the CPU) of what the reference computes in
    models/networks/generator.py:93-137    SPADEGenerator.forward (is_test: phase = 4)
    models/networks/generator.py:311-351   lip2d / SoftGate / SimplifiedLIP
    models/networks/generator.py:353-400   LIPEncoder, LIPSPADEGenerator.encode
    models/networks/architecture.py:52-70  SPADEResnetBlock.forward / shortcut / actvn
    models/networks/normalization.py:94-107 SPADE.forward
    models/networks/sync_batchnorm/batchnorm.py:65-68  SynchronizedBatchNorm2d on one device = F.batch_norm(training)
    torch.nn.utils.spectral_norm (torch, training mode: one power iteration per forward, u and v updated in place)
    utils/inference/video_processing.py:246-285  Frames / face_enhancement pre- and post-processing
with the `.train()` semantics of inference.py:48 (batch statistics in every SPADE, spectral-norm power iteration).

Pinning: the reference holds no golden vectors for this path; the oracle is pinned against the
reference's own LIPSPADEGenerator(TestOptions) run in the build container (oracle/gen_golden.py ->
tests/golden/sr_*.npz); tests/test_oracle_sr.py checks this restatement against those files everywhere and
against the live reference where it is mounted.
"""
import numpy as np
import torch
import torch.nn.functional as F

BLOCKS = [("head_0", 768, 768), ("G_middle_0", 768, 768), ("G_middle_1", 768, 768),
          ("ups.0", 768, 384), ("ups.1", 384, 192), ("ups.2", 192, 96), ("ups.3", 96, 48)]


def spectral_weight(sd, p, state):
    """torch.nn.utils.spectral_norm.SpectralNorm.compute_weight(do_power_iteration=True), eps 1e-12.
    `state` maps key -> current u (updated in place like the module buffer); returns W / sigma."""
    w = sd[p + ".weight_orig"]
    wm = w.reshape(w.shape[0], -1)
    u = state.setdefault(p + ".weight_u", sd[p + ".weight_u"].clone())
    v = F.normalize(torch.mv(wm.t(), u), dim=0, eps=1e-12)
    u_new = F.normalize(torch.mv(wm, v), dim=0, eps=1e-12)
    state[p + ".weight_u"] = u_new
    state[p + ".weight_v"] = v
    sigma = torch.dot(u_new, torch.mv(wm, v))
    return w / sigma


def spade(sd, p, x, seg):
    """SPADE.forward (normalization.py:94-107) with the param-free norm in training mode."""
    normalized = F.batch_norm(x, None, None, None, None, True, 0.1, 1e-5)
    segd = F.interpolate(seg, size=x.shape[2:], mode="nearest")
    actv = F.relu(F.conv2d(segd, sd[p + ".mlp_shared.0.weight"], sd[p + ".mlp_shared.0.bias"], padding=1))
    gamma = F.conv2d(actv, sd[p + ".mlp_gamma.weight"], None, padding=1)
    beta = F.conv2d(actv, sd[p + ".mlp_beta.weight"], None, padding=1)
    return normalized * gamma + beta


def spade_resblk(sd, p, fin, fout, x, seg, state):
    """SPADEResnetBlock.forward (architecture.py:52-70)."""
    if fin != fout:
        x_s = F.conv2d(spade(sd, p + ".norm_s", x, seg), spectral_weight(sd, p + ".conv_s", state), None)
    else:
        x_s = x
    dx = F.conv2d(F.leaky_relu(spade(sd, p + ".norm_0", x, seg), 0.2), spectral_weight(sd, p + ".conv_0", state),
                  sd[p + ".conv_0.bias"], padding=1)
    dx = F.conv2d(F.leaky_relu(spade(sd, p + ".norm_1", dx, seg), 0.2), spectral_weight(sd, p + ".conv_1", state),
                  sd[p + ".conv_1.bias"], padding=1)
    return x_s + dx


def lip_encode(sd, x, taps=None):
    """LIPEncoder.forward (generator.py:353-388)."""
    m = "lip_encoder.model."
    x = F.relu(F.instance_norm(F.conv2d(x, sd[m + "0.weight"], None, padding=1), eps=1e-5))
    for i in range(5):
        li, ci = 3 + 4 * i, 4 + 4 * i
        logit = F.conv2d(x, sd[m + "%d.logit.0.weight" % li], None, padding=1)
        logit = F.instance_norm(logit, weight=sd[m + "%d.logit.1.weight" % li], bias=sd[m + "%d.logit.1.bias" % li],
                                eps=1e-5)
        w = (torch.sigmoid(logit) * 12.0).exp()                      # SoftGate.COEFF = 12 (generator.py:317-324)
        x = F.avg_pool2d(x * w, 3, 2, 1) / F.avg_pool2d(w, 3, 2, 1)  # lip2d (generator.py:311-313)
        x = F.conv2d(x, sd[m + "%d.weight" % ci], sd[m + "%d.bias" % ci], padding=1)
        x = F.instance_norm(x, eps=1e-5)
        if i < 4:
            x = F.relu(x)
        if taps is not None:
            taps["enc%d" % (i + 1)] = x
    return x


def sr_forward(sd, inp, state=None, taps=None):
    """LIPSPADEGenerator.forward(input) with seg = input (generator.py:93-137), train-mode semantics.
    inp [B,3,256,256] RGB in [0,1].  `state` carries the spectral-norm u vectors across calls (mutated)."""
    if state is None:
        state = {}
    seg = inp
    x = lip_encode(sd, inp, taps)
    for k, (name, fin, fout) in enumerate(BLOCKS):
        if k in (1, 3, 4, 5, 6):
            x = F.interpolate(x, scale_factor=2.0, mode="nearest")
        x = spade_resblk(sd, name, fin, fout, x, seg, state)
        if taps is not None:
            taps[name] = x
    x = F.conv2d(F.leaky_relu(x, 0.2), sd["to_rgbs.3.weight"], sd["to_rgbs.3.bias"], padding=1)
    return torch.tanh(x)


def preprocess_ref(crops_bgr_u8):
    """Frames.__getitem__ (video_processing.py:253-256): BGR -> RGB, ToTensor (uint8 / 255, CHW)."""
    x = torch.from_numpy(np.ascontiguousarray(crops_bgr_u8[..., ::-1])).permute(0, 3, 1, 2).float().div(255)
    return x


def postprocess_ref(gen):
    """face_enhancement (video_processing.py:277-280): clamp(g*255, 0, 255) -> uint8 (truncation) -> HWC -> BGR."""
    g = torch.clamp(gen * 255, 0, 255).type(torch.uint8).permute(0, 2, 3, 1).numpy()
    return np.ascontiguousarray(g[..., ::-1])


def face_enhancement_ref(sd, crops_bgr_u8, state=None, batch=20):
    """face_enhancement for one list of crops: batches of 20 in order, each with its own batch statistics."""
    if state is None:
        state = {}
    out = []
    for i in range(0, len(crops_bgr_u8), batch):
        out.append(postprocess_ref(sr_forward(sd, preprocess_ref(crops_bgr_u8[i:i + batch]), state)))
    return np.concatenate(out, 0)
