"""CPU: the C-ABI library builds, loads and exports every symbol include/dexgg.h declares (no compute calls)."""
import ctypes
import os
import re

from conftest import REPO
from dex_gan_grasp_b200 import _native, build


def declared_symbols():
    src = open(os.path.join(REPO, "include", "dexgg.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dexgg_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    path = build.build()
    lib = ctypes.CDLL(path)
    names = declared_symbols()
    assert len(names) >= 7
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dexgg.h but not exported"
        assert n in _native.SIGNATURES, f"{n} has no ctypes signature in _native.py"
    assert set(_native.SIGNATURES) == set(names)


def test_version_and_error_string_without_gpu():
    L = _native.lib()
    assert L.dexgg_version() == 100
    assert isinstance(L.dexgg_last_error(), bytes)


def test_no_cpu_fallback_without_a_b200():
    import torch
    if torch.cuda.is_available():
        return
    L = _native.lib()
    rc = L.dexgg_device_check()
    assert rc in (-2, -4) and len(L.dexgg_last_error()) > 0
    rc = L.dexgg_bps_encode(None, 1, None, None, 1, 1, 1, None, None, 0, None, 0, None)
    assert rc != 0  # compute entry points refuse to run off-GPU
