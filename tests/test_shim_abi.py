"""The C-ABI library loads and exports every symbol include/pymasc_b200.h declares
(no compute calls: this runs without a GPU)."""
import ctypes
import os
import re

import pytest

import helpers
from pymasc_b200 import _shim, build


def test_library_is_built_in_tree():
    build.build()
    assert os.path.exists(_shim.LIB_PATH)
    assert os.path.dirname(_shim.LIB_PATH).endswith("pymasc_b200")


def test_every_declared_symbol_is_exported():
    header = open(os.path.join(helpers.ROOT, "include", "pymasc_b200.h")).read()
    declared = set(re.findall(r"\b(pmb_[a-z_]+)\s*\(", header))
    declared -= {"pmb_ctx", "pmb_batch"}
    assert declared == set(_shim.EXPORTS)
    lib = ctypes.CDLL(_shim.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name


def test_struct_layouts_match_the_header():
    # 64-bit fields first, no implicit padding surprises
    assert ctypes.sizeof(_shim.ChromArgs) == 8 + 4 * 4 + 8 * 9
    assert ctypes.sizeof(_shim.ChromOut) == 8 * 7
    assert ctypes.sizeof(_shim.Timing) == 4 * 3 + 4 * 2 + 4 + 8 + 4 * 4   # one pad word before the u64


def test_create_fails_cleanly_without_a_device():
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    lib = _shim.load_library()
    h = ctypes.c_void_p()
    rc = lib.pmb_create(ctypes.byref(h), 0)
    assert rc == _shim.PMB_ERR_CUDA and not h.value
    assert b"CUDA" in lib.pmb_last_error(None) or b"device" in lib.pmb_last_error(None)
    with pytest.raises(_shim.DeviceError):
        _shim.Device(0)


def test_sass_uses_tma_bulk_copies_and_popc():
    """The shift-window kernel stages tiles with cp.async.bulk (SASS UBLKCP) and counts with
    POPC + LOP3; checked on the built cubin, no GPU needed."""
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", _shim.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    k3 = sass[sass.index("shift_window_kernel"):]
    for mnemonic in ("UBLKCP", "POPC", "LOP3.LUT", "SHF.R", "SYNCS"):
        assert mnemonic in k3, mnemonic
