"""GPU: ghost_crop_faces_noalloc (the allocation-free variant of the crop extraction, scratch in a file-scope
`static __device__` array; no counterpart in the reference) against the
existing ghost_crop_faces entry point, through the C ABI, on two calls in a row - bit for bit, the tolerance every
crop test uses.  Collected last (test_zz_*) so that a failure here cannot hide the other tests under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest
import torch

from ghost_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def _call(fn, frames, tfms, face_frame, want224, want256):
    n = tfms.shape[0]
    dev = frames.device
    c224 = torch.full((n, 224, 224, 3), 7, dtype=torch.uint8, device=dev) if want224 else None
    c256 = torch.full((n, 256, 256, 3), 9, dtype=torch.uint8, device=dev) if want256 else None
    tf = np.ascontiguousarray(tfms, dtype=np.float64)
    ff = np.ascontiguousarray(face_frame, dtype=np.int32)
    rc = fn(C.c_void_p(frames.data_ptr()), frames.shape[0], frames.shape[1], frames.shape[2],
            tf.ctypes.data_as(C.POINTER(C.c_double)), ff.ctypes.data_as(C.POINTER(C.c_int)), n,
            C.c_void_p(c224.data_ptr()) if want224 else None, C.c_void_p(c256.data_ptr()) if want256 else None,
            C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, c224, c256


def _case(seed, n_frames, n, H, W):
    rng = np.random.default_rng(seed)
    frames = torch.from_numpy(np.stack([synth.frame_u8(H, W, seed=seed + i) for i in range(n_frames)])).cuda()
    tfms = np.stack([synth.similarity_tfm(H, W, rng, clipped=(i % 3 == 2)) for i in range(n)])
    face_frame = np.sort(rng.integers(0, n_frames, n)).astype(np.int32)
    return frames, tfms, face_frame


def test_noalloc_crop_matches_existing_entry_point_twice_in_a_row():
    lib = _lib.load()
    # call 1: 5 faces over 3 frames, both outputs; call 2 straight after it on the same stream: 32 faces (the limit),
    # only the 256x256 crops wanted, so the 224x224 intermediate lives in the static scratch the first call also used
    cases = [(_case(11, 3, 5, 540, 960), True, True), (_case(23, 2, _lib.CROP_NOALLOC_MAX_FACES, 720, 1280), False, True)]
    got = [_call(lib.ghost_crop_faces_noalloc, *c, w224, w256) for c, w224, w256 in cases]   # back to back, no sync
    torch.cuda.synchronize()
    ref = [_call(lib.ghost_crop_faces, *c, w224, w256) for c, w224, w256 in cases]
    torch.cuda.synchronize()
    for (rc_g, g224, g256), (rc_r, r224, r256) in zip(got, ref):
        assert rc_g == _lib.OK, lib.ghost_last_error()
        assert rc_r == _lib.OK, lib.ghost_last_error()
        if r224 is not None:
            assert torch.equal(g224, r224)
        assert torch.equal(g256, r256)
        assert int(g256.max()) > 9 or int(g256.min()) < 9          # the sentinel fill was overwritten
    # a third call reuses the scratch the second one left behind
    rc, g224, g256 = _call(lib.ghost_crop_faces_noalloc, *cases[0][0], True, True)
    torch.cuda.synchronize()
    assert rc == _lib.OK and torch.equal(g224, ref[0][1]) and torch.equal(g256, ref[0][2])


def test_noalloc_crop_rejects_sizes_beyond_the_static_scratch():
    lib = _lib.load()
    n = _lib.CROP_NOALLOC_MAX_FACES + 1
    frames, tfms, face_frame = _case(5, 1, n, 360, 640)
    rc, _, _ = _call(lib.ghost_crop_faces_noalloc, frames, tfms, face_frame, True, True)
    assert rc == _lib.ERR_INVALID
    msg = lib.ghost_last_error().decode()
    assert "exceeds the static scratch limit of %d" % _lib.CROP_NOALLOC_MAX_FACES in msg, msg
    rc, _, _ = _call(lib.ghost_crop_faces_noalloc, frames, tfms[:2], np.array([0, 3], np.int32), True, True)
    assert rc == _lib.ERR_INVALID and b"face_frame out of range" in lib.ghost_last_error()
