"""ctypes binding of the C-ABI shared library (include/ghost_b200.h).

The library is built in-tree (ghost_b200/libghost_b200.so) by `make -C ghost_b200/csrc` or
`__graft_entry__.build()`.  There is no fallback: if the library is missing, loading fails
loudly, and every compute entry point fails when there is no CUDA device.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GHOST_B200_LIB: another build of the same library (the profiling build of tools/igemm_prof.py), never a fallback
LIB_PATH = os.environ.get("GHOST_B200_LIB") or os.path.join(_HERE, "libghost_b200.so")

# ghost_status
OK, ERR_INVALID, ERR_WEIGHTS, ERR_CUDA, ERR_KERNEL = 0, -1, -2, -3, -4
# ghost_dtype
F32, BF16, F16, U8 = 0, 1, 2, 3
# ghost_paste_mode / ghost_crop_kind
PASTE_EXACT, PASTE_FLOAT = 0, 1
CROP_U8_BGR, CROP_Y_BF16, CROP_Y_F32, CROP_Y_F16 = 0, 1, 2, 3
CROP_NOALLOC_MAX_FACES = 32   # GHOST_CROP_NOALLOC_MAX_FACES


class GhostError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("ghost_b200 error %d: %s" % (code, msg))
        self.code = code


class PasteDesc(C.Structure):
    """ghost_paste_desc (include/ghost_b200.h)."""
    _fields_ = [
        ("crops", C.c_void_p), ("crop_kind", C.c_int), ("n_crops", C.c_int),
        ("masks", C.c_void_p), ("mask_f64", C.c_int), ("n_masks", C.c_int),
        ("frames_in", C.c_void_p), ("frames_out", C.c_void_p),
        ("n_frames", C.c_int), ("H", C.c_int), ("W", C.c_int),
        ("n_faces", C.c_int),
        ("tfm", C.POINTER(C.c_double)),
        ("face_frame", C.POINTER(C.c_int)), ("face_crop", C.POINTER(C.c_int)), ("face_mask", C.POINTER(C.c_int)),
        ("mode", C.c_int),
    ]


_vp, _i, _f, _d = C.c_void_p, C.c_int, C.c_float, C.c_double
_fp = C.POINTER(C.c_float)

# name -> (restype, argtypes): every symbol include/ghost_b200.h declares
SIGNATURES = {
    "ghost_version": (_i, []),
    "ghost_last_error": (C.c_char_p, []),
    "ghost_check_errors": (_i, []),
    "ghost_aei_create": (_i, [C.POINTER(_vp), C.c_char_p, _i, _i, _i]),
    "ghost_arcface_create": (_i, [C.POINTER(_vp), _i]),
    "ghost_net_set_weight": (_i, [_vp, C.c_char_p, _vp, C.POINTER(C.c_int64), _i]),
    "ghost_net_finalize": (_i, [_vp]),
    "ghost_aei_forward": (_i, [_vp, _vp, _i, _vp, _i, _vp, C.POINTER(_vp), _i, _i, _vp]),
    "ghost_aei_encode": (_i, [_vp, _vp, _i, C.POINTER(_vp), _i, _i, _vp]),
    "ghost_aei_swap_u8": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i, _vp]),
    "ghost_aei_encode_u8": (_i, [_vp, _vp, _i, _vp]),
    "ghost_aei_generate_u8": (_i, [_vp, _vp, _i, _i, _vp, _i, _vp]),
    "ghost_net_y_nhwc": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_i)]),
    "ghost_arcface_forward": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp]),
    "ghost_arcface_embed_hwc": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp]),
    "ghost_net_enable_graph": (_i, [_vp, _i]),
    "ghost_net_stats": (_i, [_vp, _i, C.POINTER(_i), C.POINTER(_d)]),
    "ghost_net_profile": (_i, [_vp, _i, _i, _i, _fp, C.POINTER(_d), C.POINTER(_d), C.POINTER(_i), C.c_char_p, _i,
                               C.POINTER(_i), _vp]),
    "ghost_net_debug_get": (_i, [_vp, _i, C.c_char_p, _vp, _i, C.POINTER(_i), _vp]),
    "ghost_sr_create": (_i, [C.POINTER(_vp), _i]),
    "ghost_sr_forward": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp]),
    "ghost_sr_enhance_u8": (_i, [_vp, _vp, _vp, _i, _vp]),
    "ghost_sr_read_buffer": (_i, [_vp, C.c_char_p, _fp, _i, C.POINTER(_i), _vp]),
    "ghost_net_destroy": (None, [_vp]),
    "ghost_preproc_u8": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "ghost_postproc_u8": (_i, [_vp, _i, _vp, _i, _vp]),
    "ghost_paste_back": (_i, [C.POINTER(PasteDesc), _vp]),
    "ghost_paste_boxes": (_i, [C.POINTER(_d), _i, _i, _i, C.POINTER(_i)]),
    "ghost_copy_boxes": (_i, [_vp, _vp, _i, _i, _i, C.POINTER(_i), C.POINTER(_i), _i, _i, _vp]),
    "ghost_match_identities": (_i, [_vp, _i, _vp, _i, _i, _f, _vp, _vp, _vp, _vp]),
    "ghost_crop_faces": (_i, [_vp, _i, _i, _i, C.POINTER(_d), C.POINTER(_i), _i, _vp, _vp, _vp]),
    "ghost_crop_faces_noalloc": (_i, [_vp, _i, _i, _i, C.POINTER(_d), C.POINTER(_i), _i, _vp, _vp, _vp]),
    "ghost_peer_alloc": (_i, [C.c_size_t, C.POINTER(_vp), C.c_char_p]),
    "ghost_peer_open": (_i, [C.c_char_p, C.POINTER(_vp)]),
    "ghost_peer_close": (_i, [_vp]),
    "ghost_peer_free": (_i, [_vp]),
    "ghost_peer_copy": (_i, [_vp, _vp, C.c_size_t, _vp]),
    "ghost_peer_signal": (_i, [_vp, C.c_uint32, _vp]),
    "ghost_peer_wait": (_i, [_vp, C.c_uint32, _vp]),
    "ghost_face_masks": (_i, [_fp, C.POINTER(_i), _i, _vp, _i, _vp]),
    "ghost_test_conv": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _vp, _i, _i, _vp]),
    "ghost_test_aad": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _f,
                            _vp, _i, _vp]),
    "ghost_test_last_error": (C.c_char_p, []),
    "ghost_test_last_ms": (_f, []),
    "ghost_test_mask_job": (_i, [_fp, _i, _i, _i, C.POINTER(C.c_uint32), C.POINTER(C.c_uint16), C.POINTER(C.c_uint16),
                                C.POINTER(_i)]),
    "ghost_test_upsample2x": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "ghost_test_upsample2x_stats": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _fp, _fp, _vp]),
}

_lib = None


def load():
    """Load libghost_b200.so (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "ghost_b200: %s is missing - build it with `make -C ghost_b200/csrc` (there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code, test_api=False):
    if code != OK:
        lib = load()
        msg = (lib.ghost_test_last_error() if test_api else lib.ghost_last_error()) or b""
        raise GhostError(code, msg.decode("utf-8", "replace"))


def version():
    return load().ghost_version()
