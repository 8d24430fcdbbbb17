"""Host-side mirror of the reference's operator interface for the face-swap hot path.

Same names, argument meaning and error behaviour as the reference (postworthy/ghost):

    AEI_Net(backbone, num_blocks=2, c_id=256).forward(Xt, z_id) -> (Y, attr)   network/AEI_Net.py:143-159
    iresnet100(fp16=False)(x[B,3,112,112]) -> z_id[B,512]                        arcface_model/iresnet.py (un-vendored)
    transform_target_to_torch(resized_frs, half=True)                           utils/inference/core.py:13-26
    faceshifter_batch(source_emb, target, G) -> uint8 BGR [B,256,256,3]         utils/inference/faceshifter_run.py:5-23
    normalize_and_torch(_batch)                                                 utils/inference/image_processing.py:23-48
    get_final_image(final_frames, crop_frames, full_frame, tfm_arrays, handler) utils/inference/image_processing.py:51-77
    get_final_video(final_frames, crop_frames, full_frames, tfm_array, ...)     utils/inference/video_processing.py:191-243

Everything numerical runs in the CUDA library behind include/ghost_b200.h; torch is used for
device memory, streams and dtype bookkeeping only.  There is no CPU path: tensors must live on
a CUDA device and the library must be built, otherwise the call raises.
"""
import ctypes as C
from collections import OrderedDict

import numpy as np
import torch

from . import _lib
from . import weights as _w

_DT = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16, torch.float16: _lib.F16, torch.uint8: _lib.U8}


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _require_cuda(t, name):
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor, got %s" % (name, type(t).__name__))
    if not t.is_cuda:
        raise ValueError("%s must be a CUDA tensor (ghost_b200 has no CPU path)" % name)


class _Net(object):
    """Weight store + lazily created device plans (one C handle per (device, arithmetic mode))."""

    _kind = None

    def __init__(self, spec, seed):
        self._spec = spec
        self._sd = _w.init_state_dict(spec, seed)
        self._dtype = torch.float32
        self._device = None
        self._handles = {}
        self._graph = False
        self.training = False

    # ---- nn.Module-like surface used by the reference's call sites (inference.py:26-36)
    def eval(self):
        self.training = False
        return self

    def train(self, mode=True):
        if mode:
            raise NotImplementedError("ghost_b200 implements the inference path only (eval-mode BatchNorm)")
        return self.eval()

    def requires_grad_(self, flag=False):
        if flag:
            raise NotImplementedError("ghost_b200 implements the inference path only")
        return self

    def half(self):
        self._dtype = torch.float16
        return self

    def bfloat16(self):
        self._dtype = torch.bfloat16
        return self

    def float(self):
        self._dtype = torch.float32
        return self

    def cuda(self, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("ghost_b200 needs a CUDA device (no CPU fallback)")
        self._device = torch.device("cuda", torch.cuda.current_device() if device is None else
                                    (device if isinstance(device, int) else torch.device(device).index or 0))
        return self

    def to(self, *args, **kwargs):
        for a in list(args) + list(kwargs.values()):
            if isinstance(a, torch.dtype):
                if a not in (torch.float32, torch.float16, torch.bfloat16):
                    raise TypeError("unsupported dtype %s" % a)
                self._dtype = a
            elif isinstance(a, (str, torch.device)):
                d = torch.device(a)
                if d.type != "cuda":
                    raise ValueError("ghost_b200 modules live on CUDA devices only")
                self.cuda(d.index)
        return self

    def cpu(self):
        raise ValueError("ghost_b200 modules live on CUDA devices only (no CPU fallback)")

    @property
    def dtype(self):
        return self._dtype

    def state_dict(self):
        return OrderedDict((k, v.clone()) for k, v in self._sd.items())

    def load_state_dict(self, sd, strict=True):
        missing = [k for k in self._spec if k not in sd]
        unexpected = [k for k in sd if k not in self._spec]
        if strict and (missing or unexpected):
            raise RuntimeError("Error(s) in loading state_dict: missing keys %s, unexpected keys %s" %
                               (missing[:8], unexpected[:8]))
        new = OrderedDict()
        for k, (shape, _) in self._spec.items():
            if k not in sd:
                new[k] = self._sd[k]
                continue
            v = sd[k].detach().to("cpu")
            if tuple(v.shape) != tuple(shape):
                raise RuntimeError("size mismatch for %s: expected %s, got %s" % (k, tuple(shape), tuple(v.shape)))
            new[k] = v.to(torch.int64) if k.endswith("num_batches_tracked") else v.to(torch.float32).contiguous()
        self._sd = new
        self._drop_handles()
        return self

    def enable_graph(self, enable=True):
        """Replay the whole forward as one CUDA graph per batch size."""
        self._graph = bool(enable)
        lib = _lib.load()
        for h in self._handles.values():
            _lib.check(lib.ghost_net_enable_graph(h, int(self._graph)))
        return self

    def _drop_handles(self):
        if self._handles:
            lib = _lib.load()
            torch.cuda.synchronize()
            for h in self._handles.values():
                lib.ghost_net_destroy(h)
        self._handles = {}

    def __del__(self):
        try:
            self._drop_handles()
        except Exception:
            pass

    def _mode(self):
        # the module dtype IS the arithmetic mode, as for the reference module: .float() -> fp32 CUDA cores,
        # .half() -> fp16 storage on tcgen05 (inference.py:30 `G.half()`), .bfloat16() -> bf16 storage on tcgen05
        return {torch.float32: _lib.F32, torch.float16: _lib.F16, torch.bfloat16: _lib.BF16}[self._dtype]

    def _create(self, lib, out, mode):
        raise NotImplementedError

    def _handle(self, device):
        lib = _lib.load()
        key = (device.index, self._mode())
        h = self._handles.get(key)
        if h is not None:
            return h
        with torch.cuda.device(device):
            out = C.c_void_p()
            self._create(lib, out, self._mode())
            try:
                for k, v in self._sd.items():
                    if k.endswith("num_batches_tracked"):
                        continue
                    shape = (C.c_int64 * max(1, v.dim()))(*v.shape)
                    _lib.check(lib.ghost_net_set_weight(out, k.encode(), C.c_void_p(v.data_ptr()), shape, v.dim()))
                _lib.check(lib.ghost_net_finalize(out))
                _lib.check(lib.ghost_net_enable_graph(out, int(self._graph)))
            except Exception:
                lib.ghost_net_destroy(out)
                raise
        self._handles[key] = out
        return out

    def stats(self, B, device=None):
        """(kernel launches, dense FLOPs) of one forward at batch B."""
        lib = _lib.load()
        dev = device or self._device or torch.device("cuda", torch.cuda.current_device())
        n, f = C.c_int(), C.c_double()
        with torch.cuda.device(dev):
            _lib.check(lib.ghost_net_stats(self._handle(dev), B, C.byref(n), C.byref(f)))
        return n.value, f.value

    def profile(self, B, reps=5, device=None):
        """Per-op CUDA-event timing of one forward at batch B: list of dicts (name, cls, ms, flops, bytes)."""
        lib = _lib.load()
        dev = device or self._device or torch.device("cuda", torch.cuda.current_device())
        cap, stride = 1024, 64
        ms, fl, by = (C.c_float * cap)(), (C.c_double * cap)(), (C.c_double * cap)()
        cl, names, n = (C.c_int * cap)(), C.create_string_buffer(cap * stride), C.c_int()
        with torch.cuda.device(dev):
            _lib.check(lib.ghost_net_profile(self._handle(dev), B, reps, cap, ms, fl, by, cl, names, stride,
                                             C.byref(n), _stream()))
        raw = names.raw
        return [dict(name=raw[i * stride:(i + 1) * stride].split(b"\0", 1)[0].decode(), cls=cl[i], ms=ms[i],
                     flops=fl[i], bytes=by[i]) for i in range(min(n.value, cap))]

    def intermediate(self, name, B, device=None, dtype=torch.float32):
        """Named intermediate activation of the last forward at batch B as an NCHW tensor (block-level tests)."""
        lib = _lib.load()
        dev = device or self._device or torch.device("cuda", torch.cuda.current_device())
        shape = (C.c_int * 4)()
        with torch.cuda.device(dev):
            h = self._handle(dev)
            _lib.check(lib.ghost_net_debug_get(h, B, name.encode(), None, _DT[dtype], shape, _stream()))
            out = torch.empty(tuple(shape), dtype=dtype, device=dev)
            _lib.check(lib.ghost_net_debug_get(h, B, name.encode(), _ptr(out), _DT[dtype], shape, _stream()))
        return out

    def __call__(self, *a, **k):
        return self.forward(*a, **k)


class AEI_Net(_Net):
    """Drop-in for network.AEI_Net.AEI_Net (network/AEI_Net.py:143-159), inference only.

    The arithmetic mode follows the module dtype like the reference: `.float()` -> fp32 CUDA-core
    correctness mode (1e-4 contract); `.half()` -> tcgen05 fp16 x fp16 -> fp32 with fp16 activations and
    weights, the dtype the reference deploys (inference.py:30) and the mode that meets the 1e-2 / 1e-3
    contract; `.bfloat16()` -> the same kernels on bf16 storage (same speed, 8 mantissa bits)."""

    def __init__(self, backbone, num_blocks=2, c_id=256, seed=0):
        super().__init__(_w.aei_spec(backbone, num_blocks, c_id), seed)
        self.backbone, self.num_blocks, self.c_id = backbone, num_blocks, c_id
        self.return_attr = True  # the reference always returns the 8 attribute maps

    def _create(self, lib, out, mode):
        _lib.check(lib.ghost_aei_create(C.byref(out), self.backbone.encode(), self.num_blocks, self.c_id, mode))

    def _attr_shapes(self, B):
        blocks = _w.BLOCKS[self.backbone]
        return [(B, blocks[k][2], 2 << k, 2 << k) for k in range(8)]

    def _check_x(self, Xt, name):
        _require_cuda(Xt, name)
        if Xt.dim() != 4 or tuple(Xt.shape[1:]) != (3, 256, 256):
            raise ValueError("%s must be [B,3,256,256], got %s" % (name, tuple(Xt.shape)))
        if Xt.dtype not in (torch.float32, torch.float16, torch.bfloat16):
            raise TypeError("%s has unsupported dtype %s" % (name, Xt.dtype))
        return Xt.contiguous()

    def forward(self, Xt, z_id):
        Xt = self._check_x(Xt, "Xt")
        _require_cuda(z_id, "z_id")
        B = Xt.shape[0]
        if z_id.dtype not in (torch.float32, torch.float16, torch.bfloat16):
            raise TypeError("z_id has unsupported dtype %s" % z_id.dtype)
        z = z_id.reshape(z_id.shape[0], -1).contiguous()  # [B,512] or [B,1,512] (export-onnx.py:30)
        if tuple(z.shape) != (B, self.c_id):
            raise ValueError("z_id must be [%d,%d], got %s" % (B, self.c_id, tuple(z_id.shape)))
        if z.device != Xt.device:
            raise ValueError("Xt and z_id must be on the same device")
        lib = _lib.load()
        with torch.cuda.device(Xt.device):
            h = self._handle(Xt.device)
            Y = torch.empty((B, 3, 256, 256), dtype=self._dtype, device=Xt.device)
            attrs, ptrs = None, None
            if self.return_attr:
                attrs = tuple(torch.empty(s, dtype=self._dtype, device=Xt.device) for s in self._attr_shapes(B))
                ptrs = (C.c_void_p * 8)(*[a.data_ptr() for a in attrs])
            _lib.check(lib.ghost_aei_forward(h, _ptr(Xt), _DT[Xt.dtype], _ptr(z), _DT[z.dtype], _ptr(Y), ptrs,
                                             _DT[self._dtype], B, _stream()))
        return Y, attrs

    def get_attr(self, X):
        X = self._check_x(X, "X")
        B = X.shape[0]
        lib = _lib.load()
        with torch.cuda.device(X.device):
            h = self._handle(X.device)
            attrs = tuple(torch.empty(s, dtype=self._dtype, device=X.device) for s in self._attr_shapes(B))
            ptrs = (C.c_void_p * 8)(*[a.data_ptr() for a in attrs])
            _lib.check(lib.ghost_aei_encode(h, _ptr(X), _DT[X.dtype], ptrs, _DT[self._dtype], B, _stream()))
        return attrs

    def swap_u8(self, crops_bgr, z_id, out=None, write_u8=True):
        """Fused transform_target_to_torch + faceshifter_batch (core.py:13-26, faceshifter_run.py:5-23) without
        the bounce through the host: uint8 BGR crops [B,256,256,3] on the device -> uint8 BGR [B,256,256,3] on
        the device.  z_id is [1,c_id] (tiled over the batch as faceshifter_batch does) or [B,c_id].
        write_u8=False skips the uint8 conversion (the paste-back reads the generator output through y_nhwc)."""
        _require_cuda(crops_bgr, "crops_bgr")
        _require_cuda(z_id, "z_id")
        if crops_bgr.dtype != torch.uint8 or crops_bgr.dim() != 4 or tuple(crops_bgr.shape[1:]) != (256, 256, 3):
            raise ValueError("crops_bgr must be uint8 [B,256,256,3]")
        B = crops_bgr.shape[0]
        z = z_id.reshape(z_id.shape[0], -1).contiguous()
        if z.shape[0] not in (1, B) or z.shape[1] != self.c_id:
            raise ValueError("z_id must be [1,%d] or [%d,%d]" % (self.c_id, B, self.c_id))
        crops_bgr = crops_bgr.contiguous()
        lib = _lib.load()
        with torch.cuda.device(crops_bgr.device):
            h = self._handle(crops_bgr.device)
            if out is None and write_u8:
                out = torch.empty_like(crops_bgr)
            optr = _ptr(out) if write_u8 else None
            _lib.check(lib.ghost_aei_swap_u8(h, _ptr(crops_bgr), _ptr(z), _DT[z.dtype], z.shape[0], optr, B,
                                             _stream()))
        return out if write_u8 else None

    def encode_u8(self, crops_bgr):
        """First half of swap_u8: normalise + attribute encoder (network/AEI_Net.py:72-95) on the current stream.
        The encoder does not depend on z_id, so it can run on a side stream beside the ArcFace network
        (embed_and_swap does that); generate_u8 must be ordered after it."""
        _require_cuda(crops_bgr, "crops_bgr")
        if crops_bgr.dtype != torch.uint8 or crops_bgr.dim() != 4 or tuple(crops_bgr.shape[1:]) != (256, 256, 3):
            raise ValueError("crops_bgr must be uint8 [B,256,256,3]")
        crops_bgr = crops_bgr.contiguous()
        lib = _lib.load()
        with torch.cuda.device(crops_bgr.device):
            _lib.check(lib.ghost_aei_encode_u8(self._handle(crops_bgr.device), _ptr(crops_bgr), crops_bgr.shape[0],
                                               _stream()))

    def generate_u8(self, z_id, B, out=None, write_u8=True):
        """Second half of swap_u8: the AAD generator on the attribute maps encode_u8 left in the plan of batch B."""
        _require_cuda(z_id, "z_id")
        z = z_id.reshape(z_id.shape[0], -1).contiguous()
        if z.shape[0] not in (1, B) or z.shape[1] != self.c_id:
            raise ValueError("z_id must be [1,%d] or [%d,%d]" % (self.c_id, B, self.c_id))
        lib = _lib.load()
        with torch.cuda.device(z.device):
            if out is None and write_u8:
                out = torch.empty((B, 256, 256, 3), dtype=torch.uint8, device=z.device)
            _lib.check(lib.ghost_aei_generate_u8(self._handle(z.device), _ptr(z), _DT[z.dtype], z.shape[0],
                                                 _ptr(out) if write_u8 else None, B, _stream()))
        return out if write_u8 else None

    def y_nhwc(self, B, device=None):
        """(device pointer, ghost dtype) of the NHWC generator output of the last forward at batch B."""
        lib = _lib.load()
        dev = device or self._device or torch.device("cuda", torch.cuda.current_device())
        p, dt = C.c_void_p(), C.c_int()
        with torch.cuda.device(dev):
            _lib.check(lib.ghost_net_y_nhwc(self._handle(dev), B, C.byref(p), C.byref(dt)))
        return p.value, dt.value


class IResNet100(_Net):
    """Drop-in for arcface_model.iresnet.iresnet100(fp16=False) as used at inference.py:33-36."""

    def __init__(self, seed=0):
        super().__init__(_w.iresnet100_spec(), seed)

    def _create(self, lib, out, mode):
        _lib.check(lib.ghost_arcface_create(C.byref(out), mode))

    def forward(self, x):
        _require_cuda(x, "x")
        if x.dim() != 4 or tuple(x.shape[1:]) != (3, 112, 112):
            raise ValueError("x must be [B,3,112,112], got %s" % (tuple(x.shape),))
        if x.dtype not in (torch.float32, torch.float16, torch.bfloat16):
            raise TypeError("x has unsupported dtype %s" % x.dtype)
        x = x.contiguous()
        B = x.shape[0]
        lib = _lib.load()
        with torch.cuda.device(x.device):
            h = self._handle(x.device)
            # the reference's iresnet(fp16=False) returns fp32 whatever the input
            z = torch.empty((B, 512), dtype=torch.float32 if self._dtype == torch.float32 else self._dtype,
                            device=x.device)
            _lib.check(lib.ghost_arcface_forward(h, _ptr(x), _DT[x.dtype], _ptr(z), _DT[z.dtype], B, _stream()))
        return z

    def embed_hwc(self, images, out_dtype=torch.float32):
        """normalize_and_torch(_batch) + F.interpolate(scale 0.5, bilinear, align_corners=True) + netArc in one
        call (image_processing.py:23-48, core.py:44-45,52-54).  images: [B,224,224,3] RGB HWC on the device,
        uint8 (divided by 255) or float32 already in [0,1]."""
        _require_cuda(images, "images")
        if images.dim() != 4 or tuple(images.shape[1:]) != (224, 224, 3):
            raise ValueError("images must be [B,224,224,3]")
        if images.dtype not in (torch.uint8, torch.float32):
            raise TypeError("images must be uint8 or float32")
        images = images.contiguous()
        B = images.shape[0]
        lib = _lib.load()
        with torch.cuda.device(images.device):
            h = self._handle(images.device)
            z = torch.empty((B, 512), dtype=out_dtype, device=images.device)
            _lib.check(lib.ghost_arcface_embed_hwc(h, _ptr(images), int(images.dtype == torch.uint8), _ptr(z),
                                                   _DT[out_dtype], B, _stream()))
        return z


def iresnet100(pretrained=False, progress=True, fp16=False, **kwargs):
    if pretrained:
        raise ValueError("no pretrained weights are bundled; load_state_dict() a backbone.pth state dict")
    net = IResNet100()
    return net.half() if fp16 else net


# ------------------------------------------------------------------------------------------------ pre / post
def transform_target_to_torch(resized_frs, half=True):
    """utils/inference/core.py:13-26: uint8 BGR [B,256,256,3] -> normalised RGB, returned (like the reference)
    as the NCHW *view* of an NHWC tensor.  half=True reproduces the fp16 rounding order of the reference."""
    src = torch.from_numpy(np.ascontiguousarray(resized_frs)) if isinstance(resized_frs, np.ndarray) else resized_frs
    if src.dtype != torch.uint8 or src.dim() != 4 or tuple(src.shape[1:]) != (256, 256, 3):
        raise ValueError("resized_frs must be uint8 [B,256,256,3]")
    if not torch.cuda.is_available():
        raise RuntimeError("ghost_b200 needs a CUDA device (no CPU fallback)")
    src = src.cuda().contiguous()
    B = src.shape[0]
    lib = _lib.load()
    dst = torch.empty((B, 256, 256, 3), dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        _lib.check(lib.ghost_preproc_u8(_ptr(src), _ptr(dst), _lib.F32, B, int(bool(half)), _stream()))
    if half:
        dst = dst.half()  # exact: the kernel already rounded every step to fp16
    return dst.permute(0, 3, 1, 2)


def faceshifter_batch(source_emb, target, G):
    """utils/inference/faceshifter_run.py:5-23."""
    bs = target.shape[0]
    assert target.ndim == 4, "target should have 4 dimentions -- B x C x H x W"
    if bs > 1:
        source_emb = torch.cat([source_emb] * bs)
    keep = G.return_attr
    G.return_attr = False
    try:
        Y, _ = G(target, source_emb)
    finally:
        G.return_attr = keep
    lib = _lib.load()
    out = torch.empty((bs, 256, 256, 3), dtype=torch.uint8, device=Y.device)
    with torch.cuda.device(Y.device):
        p, dt = G.y_nhwc(bs, Y.device)
        _lib.check(lib.ghost_postproc_u8(C.c_void_p(p), dt, _ptr(out), bs, _stream()))
    return out.cpu().numpy()


def normalize_and_torch(image):
    """utils/inference/image_processing.py:23-34 (layout/scale bookkeeping only; the arithmetic that matters,
    normalise + resample + embed, is fused in IResNet100.embed_hwc)."""
    image = torch.tensor(np.array(image, copy=True), dtype=torch.float32).cuda()
    if image.max() > 1.:
        image = image / 255.
    image = image.permute(2, 0, 1).unsqueeze(0)
    return (image - 0.5) / 0.5


def normalize_and_torch_batch(frames):
    """utils/inference/image_processing.py:37-48."""
    batch = torch.from_numpy(np.array(frames, copy=True)).cuda()
    if batch.max() > 1.:
        batch = batch / 255.
    batch = batch.permute(0, 3, 1, 2)
    return (batch - 0.5) / 0.5


# ------------------------------------------------------------------------------------------------ paste-back
def _as_dev(a, dtype, device):
    if isinstance(a, torch.Tensor):
        return a.to(device=device, dtype=dtype).contiguous()
    return torch.from_numpy(np.ascontiguousarray(a)).to(device=device, dtype=dtype)


def paste_back(frames, crops, masks, tfms, face_frame, face_crop=None, face_mask=None, mode="exact", out=None,
               crop_kind=None):
    """Fused inverse-warp + mask composite of swapped crops into full frames (K11).

    frames : uint8 BGR [F,H,W,3] (device tensor or numpy)
    crops  : uint8 BGR [n,256,256,3], or the generator's NHWC output given as (device_ptr, ghost dtype, n)
    masks  : [m,224,224] float64 / float32 in [0,1]
    tfms   : [n_faces,2,3] float64 frame->224-crop affines; face_frame[i] (non-decreasing) is the frame of
             face i, faces of one frame compose in list order; face_crop / face_mask default to arange.
    mode   : "exact" = get_final_image arithmetic (bit-exact), "float" = get_final_video arithmetic.
    Returns the composited frames as a device tensor [F,H,W,3] uint8."""
    if mode not in ("exact", "float"):
        raise ValueError("mode must be 'exact' or 'float'")
    if not torch.cuda.is_available():
        raise RuntimeError("ghost_b200 needs a CUDA device (no CPU fallback)")
    dev = frames.device if isinstance(frames, torch.Tensor) and frames.is_cuda else torch.device(
        "cuda", torch.cuda.current_device())
    fr = _as_dev(frames, torch.uint8, dev)
    if fr.dim() != 4 or fr.shape[3] != 3:
        raise ValueError("frames must be [F,H,W,3] uint8")
    F_, H, W = fr.shape[0], fr.shape[1], fr.shape[2]
    tf = np.ascontiguousarray(np.asarray(tfms, dtype=np.float64).reshape(-1, 2, 3))
    nf = tf.shape[0]
    ff = np.ascontiguousarray(np.asarray(face_frame, dtype=np.int32).reshape(-1))
    fc = np.arange(nf, dtype=np.int32) if face_crop is None else np.ascontiguousarray(np.asarray(face_crop, np.int32))
    fm = np.arange(nf, dtype=np.int32) if face_mask is None else np.ascontiguousarray(np.asarray(face_mask, np.int32))
    if not (len(ff) == len(fc) == len(fm) == nf):
        raise ValueError("tfms, face_frame, face_crop and face_mask must have one entry per face")
    if isinstance(crops, tuple):
        cptr, cdt, ncrops = crops
        ckind = {_lib.BF16: _lib.CROP_Y_BF16, _lib.F16: _lib.CROP_Y_F16}.get(cdt, _lib.CROP_Y_F32)
        keep = None
    else:
        keep = _as_dev(crops, torch.uint8, dev)
        if keep.dim() != 4 or tuple(keep.shape[1:]) != (256, 256, 3):
            raise ValueError("crops must be uint8 [n,256,256,3]")
        cptr, ncrops, ckind = keep.data_ptr(), keep.shape[0], _lib.CROP_U8_BGR
    if crop_kind is not None:
        ckind = crop_kind
    mk = masks if isinstance(masks, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(masks))
    if mk.dtype not in (torch.float64, torch.float32):
        mk = mk.to(torch.float64)
    mk = mk.to(dev).contiguous()
    if mk.dim() != 3 or tuple(mk.shape[1:]) != (224, 224):
        raise ValueError("masks must be [m,224,224]")
    if out is None:
        out = torch.empty_like(fr)
    elif not (isinstance(out, torch.Tensor) and out.is_cuda and out.dtype == torch.uint8 and out.is_contiguous()
              and out.shape == fr.shape and out.device == fr.device):
        raise ValueError("out must be a contiguous uint8 CUDA tensor shaped like frames")
    d = _lib.PasteDesc()
    d.crops, d.crop_kind, d.n_crops = cptr, ckind, ncrops
    d.masks, d.mask_f64, d.n_masks = mk.data_ptr(), int(mk.dtype == torch.float64), mk.shape[0]
    d.frames_in, d.frames_out = fr.data_ptr(), out.data_ptr()
    d.n_frames, d.H, d.W = F_, H, W
    d.n_faces = nf
    d.tfm = tf.ctypes.data_as(C.POINTER(C.c_double))
    d.face_frame = ff.ctypes.data_as(C.POINTER(C.c_int))
    d.face_crop = fc.ctypes.data_as(C.POINTER(C.c_int))
    d.face_mask = fm.ctypes.data_as(C.POINTER(C.c_int))
    d.mode = _lib.PASTE_EXACT if mode == "exact" else _lib.PASTE_FLOAT
    lib = _lib.load()
    with torch.cuda.device(dev):
        _lib.check(lib.ghost_paste_back(C.byref(d), _stream()))
    return out


def crop_faces(frames, tfms, face_frame, want224=True, want256=True):
    """The crop extraction in front of the generator (utils/inference/video_processing.py:159-168 + resize_frames
    :174-188), on the device and bit-exact against the reference's OpenCV calls:

        crops224[i] = cv2.warpAffine(frames[face_frame[i]], tfms[i], (224, 224), borderValue=0.0)
        crops256[i] = cv2.resize(crops224[i], (256, 256))

    frames: uint8 BGR [F,H,W,3] (device tensor or numpy); tfms [n,2,3] float64 frame->crop affines.
    Returns (crops224, crops256) device tensors (None for the one not asked for)."""
    if not torch.cuda.is_available():
        raise RuntimeError("ghost_b200 needs a CUDA device (no CPU fallback)")
    dev = frames.device if isinstance(frames, torch.Tensor) and frames.is_cuda else torch.device(
        "cuda", torch.cuda.current_device())
    fr = _as_dev(frames, torch.uint8, dev)
    if fr.dim() != 4 or fr.shape[3] != 3:
        raise ValueError("frames must be [F,H,W,3] uint8")
    tf = np.ascontiguousarray(np.asarray(tfms, dtype=np.float64).reshape(-1, 2, 3))
    ff = np.ascontiguousarray(np.asarray(face_frame, dtype=np.int32).reshape(-1))
    n = tf.shape[0]
    if len(ff) != n:
        raise ValueError("one frame index per affine")
    if not (want224 or want256):
        raise ValueError("nothing to compute")
    c224 = torch.empty((n, 224, 224, 3), dtype=torch.uint8, device=dev) if want224 else None
    c256 = torch.empty((n, 256, 256, 3), dtype=torch.uint8, device=dev) if want256 else None
    if n == 0:
        return c224, c256
    lib = _lib.load()
    with torch.cuda.device(dev):
        _lib.check(lib.ghost_crop_faces(_ptr(fr), fr.shape[0], fr.shape[1], fr.shape[2],
                                        tf.ctypes.data_as(C.POINTER(C.c_double)), ff.ctypes.data_as(C.POINTER(C.c_int)), n,
                                        _ptr(c224) if want224 else None, _ptr(c256) if want256 else None, _stream()))
    return c224, c256


def match_identities(face_embeds, target_embeds, similarity_th):
    """utils/inference/video_processing.py:141-148: which detected face of a frame belongs to which target identity.
    face_embeds [n_faces,512], target_embeds [n_targets,512]: un-normalised ArcFace outputs on the device.
    Returns (best_idxs [n_targets] int64, matched [n_targets] bool, similarity [n_targets] float32) as numpy -
    matched[j] is the reference's `similarity[best_idx][j] > similarity_th`."""
    _require_cuda(face_embeds, "face_embeds")
    _require_cuda(target_embeds, "target_embeds")
    f = face_embeds.to(torch.float32).contiguous()
    t = target_embeds.to(torch.float32).contiguous()
    if f.dim() != 2 or t.dim() != 2 or f.shape[1] != t.shape[1]:
        raise ValueError("face_embeds [n,d] and target_embeds [t,d] expected")
    nt = t.shape[0]
    idx = torch.empty(nt, dtype=torch.int32, device=f.device)
    sim = torch.empty(nt, dtype=torch.float32, device=f.device)
    ok = torch.empty(nt, dtype=torch.int32, device=f.device)
    lib = _lib.load()
    with torch.cuda.device(f.device):
        _lib.check(lib.ghost_match_identities(_ptr(f), f.shape[0], _ptr(t), nt, f.shape[1], float(similarity_th),
                                              _ptr(idx), _ptr(sim), _ptr(ok), _stream()))
    return idx.cpu().numpy().astype(np.int64), ok.cpu().numpy().astype(bool), sim.cpu().numpy()


def embed_and_swap(G, arc, sources_rgb224, crops_bgr, side=None, event=None, out=None, write_u8=True,
                   face_source=None):
    """z_id = netArc(sources) and the swap of `crops_bgr` with it (core.py:44-45 + faceshifter_batch), with the two
    independent halves overlapped: the attribute encoder of G runs on the stream `side` while the ArcFace network runs
    on the current stream; the AAD generator follows on the current stream.  Same results as
    `G.swap_u8(crops, arc.embed_hwc(sources))`; side=None runs exactly that.
    face_source (int64 device tensor [B], optional): crop i is swapped with the identity of source face_source[i] -
    the reference embeds every source once (core.py:52-54) and reuses it for all the crops of its target."""
    zdt = torch.bfloat16 if G.dtype == torch.bfloat16 else (torch.float16 if G.dtype == torch.float16 else torch.float32)
    if side is None:
        z = arc.embed_hwc(sources_rgb224, out_dtype=zdt)
        if face_source is not None:
            z = z.index_select(0, face_source)
        return G.swap_u8(crops_bgr, z, out=out, write_u8=write_u8)
    main = torch.cuda.current_stream(crops_bgr.device)
    side.wait_stream(main)                       # inputs ready, previous users of the plan's buffers done
    with torch.cuda.stream(side):
        G.encode_u8(crops_bgr)
        event = event or torch.cuda.Event()
        event.record(side)
    z = arc.embed_hwc(sources_rgb224, out_dtype=zdt)
    if face_source is not None:
        z = z.index_select(0, face_source)
    main.wait_event(event)
    return G.generate_u8(z, crops_bgr.shape[0], out=out, write_u8=write_u8)


def paste_boxes(tfms, frame_hw):
    """Bounding boxes the paste-back rewrites, one per face: int32 [n,4] = (x0, y0, x1, y1), x1 / y1 exclusive, all
    zero for a face outside its frame.  Everything outside these boxes leaves get_final_image / get_final_video
    unchanged (image_processing.py:66-75: the warped mask is 0 there)."""
    tf = np.ascontiguousarray(np.asarray(tfms, dtype=np.float64).reshape(-1, 2, 3))
    boxes = np.zeros((tf.shape[0], 4), dtype=np.int32)
    lib = _lib.load()
    _lib.check(lib.ghost_paste_boxes(tf.ctypes.data_as(C.POINTER(C.c_double)), tf.shape[0], int(frame_hw[0]),
                                     int(frame_hw[1]), boxes.ctypes.data_as(C.POINTER(C.c_int))))
    return boxes


def _copy_boxes(dst, src, face_frame, boxes, to_device):
    """Box pixels of [F,H,W,3] uint8 frames between a pinned host tensor and a device tensor of the same shape,
    asynchronously on the current stream (cudaMemcpy2DAsync per box).  Returns the bytes moved."""
    F, H, W = int(dst.shape[0]), int(dst.shape[1]), int(dst.shape[2])
    ff = np.ascontiguousarray(np.asarray(face_frame, dtype=np.int32).reshape(-1))
    bx = np.ascontiguousarray(np.asarray(boxes, dtype=np.int32).reshape(-1, 4))
    lib = _lib.load()
    devt = dst if to_device else src
    with torch.cuda.device(devt.device):
        _lib.check(lib.ghost_copy_boxes(dst.data_ptr(), src.data_ptr(), F, H, W, ff.ctypes.data_as(C.POINTER(C.c_int)),
                                        bx.ctypes.data_as(C.POINTER(C.c_int)), len(ff), 1 if to_device else 0,
                                        _stream()))
    return int(((bx[:, 2] - bx[:, 0]) * (bx[:, 3] - bx[:, 1])).sum()) * 3


class SwapPipeline(object):
    """The per-batch video loop of the reference (utils/inference/core.py:model_inference ->
    faceshifter_batch -> video_processing.py:get_final_video) as one double-buffered device pipeline:

        pinned host batch --H2D--> ArcFace z_id -> AEI_Net swap -> paste-back --D2H--> pinned host frames

    Uploads, compute and downloads run on three CUDA streams over `depth` device slots, so the PCIe copies of
    neighbouring batches overlap the generator.  Results are bit-identical to calling embed_hwc / swap_u8 /
    paste_back one after the other.  Call submit() per batch and flush() before reading the last outputs.

    The networks start as soon as the crops and sources (22 MB per 64 faces) are on the device; the frames follow on
    the upload stream and only the paste-back waits for them.  boxes_only=True composites in place in the caller's
    pinned frame buffer and moves just the face boxes over PCIe (pass out=frames); sr=<LIPSPADEGenerator> inserts
    the --use_sr pass (face_enhancement, batches of sr_batch crops) between the swap and the paste-back."""

    def __init__(self, G, arc, batch, frame_hw, depth=2, mode="exact", mask_dtype=torch.float64, device=None,
                 boxes_only=False, sr=None, sr_batch=20, n_frames=None, n_sources=None):
        if not torch.cuda.is_available():
            raise RuntimeError("ghost_b200 needs a CUDA device (no CPU fallback)")
        self.G, self.arc, self.B, self.mode, self.depth = G, arc, int(batch), mode, int(depth)
        # several faces per frame / several crops per source (BASELINE configs[3]): n_frames frames and n_sources
        # source images per batch of `batch` crops; submit() then takes face_frame / face_source index lists
        self.F = int(batch if n_frames is None else n_frames)
        self.S = int(batch if n_sources is None else n_sources)
        # boxes_only: the frames stay in the caller's pinned host buffer and are composited IN PLACE - only the face
        # boxes cross PCIe (up, composited on the device, down into the same host frames); `out` must be `frames`
        self.boxes_only = bool(boxes_only)
        self.frame_hw = (int(frame_hw[0]), int(frame_hw[1]))
        self.h2d_bytes = self.d2h_bytes = 0       # bytes of the last submit()
        # sr: the --use_sr pass (face_enhancement, batches of sr_batch crops like the reference's DataLoader) between the
        # swap and the paste-back
        self.sr, self.sr_batch = sr, int(sr_batch)
        if sr is not None and self.B % self.sr_batch:
            # the SR generator runs on batch statistics (inference.py:48), so the batch boundaries are part of the
            # result: the reference batches 20 continuously over a target's frame list (video_processing.py:263-285)
            raise ValueError("SwapPipeline(sr=...): batch (%d) must be a multiple of sr_batch (%d) so that the SR "
                             "batches fall where face_enhancement puts them" % (self.B, self.sr_batch))
        self.dev = device or torch.device("cuda", torch.cuda.current_device())
        H, W = frame_hw
        if sr is not None:
            self.swapped = torch.empty((self.B, 256, 256, 3), dtype=torch.uint8, device=self.dev)
            self.enhanced = torch.empty_like(self.swapped)
        u8 = dict(dtype=torch.uint8, device=self.dev)
        self.slots = []
        for _ in range(self.depth):
            self.slots.append({
                "src": torch.empty((self.S, 224, 224, 3), **u8), "crp": torch.empty((self.B, 256, 256, 3), **u8),
                "frm": torch.empty((self.F, H, W, 3), **u8), "out": torch.empty((self.F, H, W, 3), **u8),
                "msk": torch.empty((self.B, 224, 224), dtype=mask_dtype, device=self.dev),
                "ev_in": torch.cuda.Event(), "ev_frm": torch.cuda.Event(), "ev_done": torch.cuda.Event(), "ev_out": torch.cuda.Event(), "used": False})
        self.s_in = torch.cuda.Stream(device=self.dev)
        self.s_out = torch.cuda.Stream(device=self.dev)
        self.s_enc = torch.cuda.Stream(device=self.dev)   # attribute encoder beside the ArcFace network
        self.ev_enc = torch.cuda.Event()
        self.i = 0

    def submit(self, sources, crops, frames, masks, tfms, face_frame, out, face_source=None):
        """sources uint8 RGB [S,224,224,3], crops uint8 BGR [B,256,256,3], frames uint8 BGR [F,H,W,3], masks
        [B,224,224]: pinned host tensors of one batch; tfms / face_frame as in paste_back (one face per entry);
        face_source: int64 device tensor [B] (source of each crop) when S != B;
        out: pinned host tensor [F,H,W,3] that receives the composited frames (valid after flush())."""
        slot = self.slots[self.i % self.depth]
        self.i += 1
        main = torch.cuda.current_stream(self.dev)
        if slot["used"]:
            self.s_in.wait_event(slot["ev_done"])   # the step that last read this slot's inputs has finished
            if self.boxes_only:
                self.s_in.wait_event(slot["ev_out"])  # ... and its boxes have been read back out of slot["frm"]
        with torch.cuda.stream(self.s_in):
            # the networks need only the small tensors (22 MB): they start while the frames (398 MB at 1080p) are
            # still crossing PCIe; only the paste-back waits for those
            slot["src"].copy_(sources, non_blocking=True)
            slot["crp"].copy_(crops, non_blocking=True)
            slot["ev_in"].record(self.s_in)
            if self.boxes_only:
                if out is not frames:
                    raise ValueError("boxes_only composites in place: pass out=frames")
                boxes = paste_boxes(tfms, self.frame_hw)
                nb = _copy_boxes(slot["frm"], frames, face_frame, boxes, to_device=True)
            else:
                slot["frm"].copy_(frames, non_blocking=True)
                nb = frames.numel()
            slot["msk"].copy_(masks, non_blocking=True)
            slot["ev_frm"].record(self.s_in)
            self.h2d_bytes = (nb + sources.numel() + crops.numel() + masks.numel() * masks.element_size())
        main.wait_event(slot["ev_in"])
        if self.sr is None:
            embed_and_swap(self.G, self.arc, slot["src"], slot["crp"], side=self.s_enc, event=self.ev_enc, write_u8=False,
                           face_source=face_source)
            yptr, ydt = self.G.y_nhwc(self.B, self.dev)
            swap_arg = (yptr, ydt, self.B)
        else:
            embed_and_swap(self.G, self.arc, slot["src"], slot["crp"], side=self.s_enc, event=self.ev_enc,
                           out=self.swapped, write_u8=True, face_source=face_source)
            for b0 in range(0, self.B, self.sr_batch):
                self.sr.enhance_u8(self.swapped[b0:b0 + self.sr_batch], out=self.enhanced[b0:b0 + self.sr_batch])
            swap_arg = self.enhanced
        main.wait_event(slot["ev_frm"])
        if slot["used"]:
            main.wait_event(slot["ev_out"])         # this slot's previous result has left the device
        dst = slot["frm"] if self.boxes_only else slot["out"]     # in place on the device too: no frame copy
        paste_back(slot["frm"], swap_arg, slot["msk"], tfms, face_frame, mode=self.mode, out=dst)
        slot["ev_done"].record(main)
        self.s_out.wait_event(slot["ev_done"])
        with torch.cuda.stream(self.s_out):
            if self.boxes_only:
                self.d2h_bytes = _copy_boxes(out, slot["frm"], face_frame, boxes, to_device=False)
            else:
                out.copy_(slot["out"], non_blocking=True)
                self.d2h_bytes = out.numel()
            slot["ev_out"].record(self.s_out)
        slot["used"] = True
        return out

    def flush(self):
        """Make the current stream wait for every outstanding download (then synchronize it to read `out`)."""
        main = torch.cuda.current_stream(self.dev)
        for slot in self.slots:
            if slot["used"]:
                main.wait_event(slot["ev_out"])


def _masks_from_handler(swaps224, crop_frames, handler, params, dtype=torch.float64):
    """Landmarks from the caller's handler (the landmark model is out of scope), masks on the device: the preset
    choice of face_mask_static (masks.py:43-69) per face, then one batched ghost_face_masks call.
    Returns a device tensor [n,224,224]; params[j] is filled in where it was None."""
    from .masks import choose_params, face_masks
    lms = []
    for j, (swap, tgt) in enumerate(zip(swaps224, crop_frames)):
        lm = handler.get_without_detection_without_transform(swap)
        lm_t = handler.get_without_detection_without_transform(tgt)
        if params[j] is None:
            params[j] = choose_params(lm, lm_t)
        lms.append(np.asarray(lm, dtype=np.float32))
    return face_masks(np.stack(lms), [list(p) for p in params], dtype=dtype)


def get_final_image(final_frames, crop_frames, full_frame, tfm_arrays, handler=None, masks=None):
    """utils/inference/image_processing.py:51-77.  Landmark detection is out of scope, so the blend masks can be
    passed directly (masks[i]: float64 224x224); with a `handler` they are built on the host exactly as the
    reference does (cv2.resize + handler + face_mask_static).  Returns uint8 [H,W,3]."""
    n = len(final_frames)
    if n == 0:
        return np.array(full_frame, dtype="uint8")
    if masks is None:
        if handler is None:
            raise ValueError("get_final_image needs either a landmark handler or precomputed masks")
        import cv2
        swaps = [cv2.resize(final_frames[i][0], (224, 224)) for i in range(n)]
        masks = _masks_from_handler(swaps, [crop_frames[i][0] for i in range(n)], handler, [None] * n)
    crops = np.stack([np.asarray(final_frames[i][0], dtype=np.uint8) for i in range(n)])
    tf = np.stack([np.asarray(tfm_arrays[i][0], dtype=np.float64) for i in range(n)])
    if not isinstance(masks, torch.Tensor):
        masks = np.stack(masks).astype(np.float64)
    out = paste_back(np.asarray(full_frame, dtype=np.uint8)[None], crops, masks, tf, np.zeros(n, np.int32),
                     mode="exact")
    return out[0].cpu().numpy()


def get_final_video(final_frames, crop_frames, full_frames, tfm_array, OUT_VIDEO_NAME=None, fps=30.0, handler=None,
                    masks=None, chunk=64):
    """utils/inference/video_processing.py:191-243 compositing (fp32 arithmetic, faces compose through uint8).
    masks[j][i] replaces the landmark handler (out of scope).  Returns the list of composited frames; when
    OUT_VIDEO_NAME is given they are also written with cv2.VideoWriter as the reference does."""
    nF, nJ = len(full_frames), len(crop_frames)
    result = []
    params = [None] * nJ
    for c0 in range(0, nF, chunk):
        c1 = min(nF, c0 + chunk)
        crops, mks, tfs, ffr = [], [], [], []
        for i in range(c0, c1):
            for j in range(nJ):
                # the reference resizes the swapped frame first (video_processing.py:212): for a frame without this face
                # that raises inside its try / except and the iteration ends with params[j] untouched; the reset at
                # :214-216 is only reached by a swapped frame whose crop is missing
                if len(final_frames[j][i]) == 0:
                    continue
                if len(crop_frames[j][i]) == 0:
                    params[j] = None
                    continue
                if masks is not None:
                    m = masks[j][i]
                else:
                    if handler is None:
                        raise ValueError("get_final_video needs either a landmark handler or precomputed masks")
                    import cv2
                    swap = cv2.resize(final_frames[j][i], (224, 224))
                    p = [params[j]]
                    m = _masks_from_handler([swap], [crop_frames[j][i]], handler, p, dtype=torch.float32)[0]
                    params[j] = p[0]
                crops.append(np.asarray(final_frames[j][i], dtype=np.uint8))
                mks.append(m if isinstance(m, torch.Tensor) else torch.from_numpy(np.asarray(m, dtype=np.float32)))
                tfs.append(np.asarray(tfm_array[j][i], dtype=np.float64))
                ffr.append(i - c0)
        frames = np.stack([np.asarray(f, dtype=np.uint8) for f in full_frames[c0:c1]])
        if crops:
            dev = torch.device("cuda", torch.cuda.current_device())
            out = paste_back(frames, np.stack(crops), torch.stack([m.to(dev) for m in mks]), np.stack(tfs),
                             np.asarray(ffr, np.int32), mode="float").cpu().numpy()
        else:
            out = frames
        result.extend(list(out))
    if OUT_VIDEO_NAME:
        import cv2
        vw = cv2.VideoWriter(str(OUT_VIDEO_NAME), cv2.VideoWriter_fourcc(*"mp4v"), fps,
                             (full_frames[0].shape[1], full_frames[0].shape[0]))
        for f in result:
            vw.write(f)
        vw.release()
    return result


# ------------------------------------------------------------------------------------------------------------------
# pix2pix super-resolution
class LIPSPADEGenerator(_Net):
    """Drop-in for models.networks.generator.LIPSPADEGenerator(TestOptions) (generator.py:390-400; options
    models/config_sr.py:3-29) in the state the reference runs it in: `model.netG.train()` (inference.py:48), i.e.
    batch statistics in every SPADE and one spectral-norm power iteration per forward.  `.float()` = fp32 CUDA-core
    mode (the reference's own dtype for this network), `.half()` / `.bfloat16()` = tcgen05 mode on fp16 / bf16
    storage."""

    def __init__(self, opt=None, seed=0):
        super().__init__(_w.sr_spec(), seed)
        self.opt = opt
        self.training = True

    def train(self, mode=True):
        if not mode:
            return self.eval()
        self.training = True
        return self

    def eval(self):
        raise NotImplementedError(
            "ghost_b200.LIPSPADEGenerator runs with train() semantics only, as the reference does (inference.py:48); "
            "the running-statistics mode is not built")

    def _create(self, lib, out, mode):
        _lib.check(lib.ghost_sr_create(C.byref(out), mode))

    def forward(self, input, seg=None):
        if seg is not None and seg is not input:
            raise NotImplementedError("separate segmentation input is not supported (the reference passes seg = input)")
        _require_cuda(input, "input")
        if input.dim() != 4 or tuple(input.shape[1:]) != (3, 256, 256):
            raise ValueError("input must be [B,3,256,256], got %s" % (tuple(input.shape),))
        if input.dtype not in (torch.float32, torch.float16, torch.bfloat16):
            raise TypeError("input has unsupported dtype %s" % input.dtype)
        x = input.contiguous()
        B = x.shape[0]
        lib = _lib.load()
        with torch.cuda.device(x.device):
            h = self._handle(x.device)
            y = torch.empty((B, 3, 256, 256), dtype=self._dtype, device=x.device)
            _lib.check(lib.ghost_sr_forward(h, _ptr(x), _DT[x.dtype], _ptr(y), _DT[self._dtype], B, _stream()))
        return y

    def enhance_u8(self, crops_bgr, out=None):
        """One DataLoader batch of face_enhancement (video_processing.py:263-285) on the device:
        uint8 BGR [B,256,256,3] -> uint8 BGR [B,256,256,3]."""
        _require_cuda(crops_bgr, "crops_bgr")
        if crops_bgr.dtype != torch.uint8 or crops_bgr.dim() != 4 or tuple(crops_bgr.shape[1:]) != (256, 256, 3):
            raise ValueError("crops_bgr must be uint8 [B,256,256,3]")
        crops_bgr = crops_bgr.contiguous()
        B = crops_bgr.shape[0]
        lib = _lib.load()
        with torch.cuda.device(crops_bgr.device):
            h = self._handle(crops_bgr.device)
            if out is None:
                out = torch.empty_like(crops_bgr)
            _lib.check(lib.ghost_sr_enhance_u8(h, _ptr(crops_bgr), _ptr(out), B, _stream()))
        return out

    def state_dict(self):
        """As the reference module: weight_u / weight_v hold the state after the forwards run so far."""
        sd = super().state_dict()
        if self._handles:
            lib = _lib.load()
            (dev_index, _), h = next(iter(self._handles.items()))
            with torch.cuda.device(dev_index):
                for k in sd:
                    if k.endswith(".weight_u") or k.endswith(".weight_v"):
                        buf = np.empty(sd[k].numel(), np.float32)
                        n = C.c_int()
                        _lib.check(lib.ghost_sr_read_buffer(h, k.encode(), buf.ctypes.data_as(C.POINTER(C.c_float)),
                                                            buf.size, C.byref(n), _stream()))
                        sd[k] = torch.from_numpy(buf.copy())
        return sd


class Pix2PixModel(object):
    """Inference surface of models.pix2pix_model.Pix2PixModel (pix2pix_model.py:16-62): `.netG` and
    `model(data, mode='inference2')`."""

    def __init__(self, opt=None, seed=0):
        self.opt = opt
        self.netG = LIPSPADEGenerator(opt, seed=seed)

    def load_state_dict(self, sd, strict=True):
        self.netG.load_state_dict(sd, strict)
        return self

    def __call__(self, data, mode):
        return self.forward(data, mode)

    def forward(self, data, mode):
        if mode != "inference2":
            raise ValueError("|mode| is invalid")      # pix2pix_model.py:62 (only the inference branch is built)
        x = data["label"]
        if not x.is_cuda:
            x = x.cuda()                               # preprocess_input (pix2pix_model.py:64-69)
        return self.netG(x)


def face_enhancement(final_frames, model, device=None):
    """utils/inference/video_processing.py:263-285: every non-empty 256x256 BGR crop of each target's frame list goes
    through the super-resolution generator in DataLoader order, batches of 20 (the crops of a batch share their
    SPADE batch statistics, so the batching is part of the result)."""
    netG = model.netG if hasattr(model, "netG") else model
    dev = device or netG._device or torch.device("cuda", torch.cuda.current_device())
    out_all = []
    for frames in final_frames:
        enhanced = list(frames)
        idx = [i for i, x in enumerate(frames) if not isinstance(x, list)]
        for b0 in range(0, len(idx), 20):
            sel = idx[b0:b0 + 20]
            batch = torch.from_numpy(np.stack([np.ascontiguousarray(frames[i]) for i in sel])).to(dev)
            res = netG.enhance_u8(batch).cpu().numpy()
            for j, i in enumerate(sel):
                enhanced[i] = res[j]
        out_all.append(enhanced)
    return out_all
