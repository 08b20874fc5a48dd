"""The C-ABI shared library loads on a CPU-only box and exports every symbol include/fastb200.h declares.
No compute entry point is called here (no GPU)."""
import ctypes
import os
import re

import pytest

from conftest import PKG, ROOT


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "fastb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fastb200_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def libpath():
    import importlib.util
    spec = importlib.util.spec_from_file_location("fastb200_build", os.path.join(PKG, "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.build()


def test_header_declares_the_documented_surface():
    syms = _header_symbols()
    for must in ("fastb200_window_gather", "fastb200_head_fwd", "fastb200_head_bwd", "fastb200_transformer_fwd",
                 "fastb200_transformer_bwd", "fastb200_ce_fwd_bwd", "fastb200_adamw_flat", "fastb200_sosfiltfilt",
                 "fastb200_resample_poly", "fastb200_sosfiltfilt_resample", "fastb200_sosfiltfilt_resample_supported", "fastb200_trials_from_mat", "fastb200_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol(libpath):
    lib = ctypes.CDLL(libpath)
    for name in _header_symbols():
        assert hasattr(lib, name), f"{name} declared in fastb200.h but not exported"


def test_binding_covers_header_and_struct_mirror_matches(libpath):
    from fast import _native as N
    assert sorted(N.SIGNATURES) == _header_symbols()
    lib = N.lib()
    assert lib.fastb200_version() >= 100
    assert lib.fastb200_sizeof_model() == ctypes.sizeof(N.ModelDesc)


def test_argument_validation_without_a_gpu(libpath):
    """Entry points reject bad arguments before any launch; the message travels through fastb200_last_error."""
    from fast import _native as N
    lib = N.lib()
    assert lib.fastb200_head_packed_floats(None) == -1
    rc = lib.fastb200_window_gather(None, None, None, 1, 64, 100, 250, 125, 1, 64, None)     # window longer than trial
    assert rc < 0 and b"windows run past" in lib.fastb200_last_error()
    rc = lib.fastb200_ce_fwd_bwd(None, None, 4, 8, 5, 1.0, None, None, None, None, None)      # int32 labels unsupported
    assert rc < 0 and b"uint8 or int64" in lib.fastb200_last_error()
    with pytest.raises(RuntimeError, match="uint8 or int64"):
        N.check(rc, "ce")
