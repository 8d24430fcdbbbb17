"""CPU: the C-ABI library loads and exports every symbol include/ghost_b200.h declares (no compute calls)."""
import ctypes as C
import os
import re

import pytest

from ghost_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ghost_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ghost_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "libghost_b200.so does not export %s" % n
    # and the binding table covers the header exactly
    assert sorted(_lib.SIGNATURES) == names


def test_version_and_error_reporting_without_gpu():
    lib = _lib.load()
    assert lib.ghost_version() == 100
    h = C.c_void_p()
    assert lib.ghost_aei_create(C.byref(h), b"resnet", 2, 512, _lib.BF16) == _lib.ERR_INVALID
    assert b"resnet" in lib.ghost_last_error()
    assert lib.ghost_aei_create(C.byref(h), b"unet", 0, 512, _lib.BF16) == _lib.ERR_INVALID
    assert lib.ghost_aei_create(C.byref(h), b"unet", 2, 500, _lib.BF16) == _lib.ERR_INVALID
    assert lib.ghost_aei_create(C.byref(h), b"unet", 2, 512, 7) == _lib.ERR_INVALID
    assert lib.ghost_aei_create(C.byref(h), b"unet", 2, 512, _lib.BF16) == _lib.OK and h.value
    # a weight can be handed over and a missing one is reported at finalize (host-side check, no kernel runs)
    import numpy as np
    w = np.zeros((32, 3, 4, 4), np.float32)
    shape = (C.c_int64 * 4)(32, 3, 4, 4)
    assert lib.ghost_net_set_weight(h, b"encoder.conv1.0.weight", C.c_void_p(w.ctypes.data), shape, 4) == _lib.OK
    assert lib.ghost_net_finalize(h) == _lib.ERR_WEIGHTS
    assert b"missing state_dict entry" in lib.ghost_last_error()
    lib.ghost_net_destroy(h)
    with pytest.raises(_lib.GhostError):
        _lib.check(_lib.ERR_INVALID)


def test_sass_is_blackwell_native():
    """The built library carries sm_100a code with tcgen05 / TMA instructions (UTC*MMA, UTMALDG, LDTM)."""
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    out = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM", "UTCBAR"):
        assert mnemonic in out, mnemonic
