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


def test_library_exports_nothing_the_header_does_not_declare(libpath):
    """exports (dynamic symbol table, defined `fastb200_*` functions) == header: no undeclared debug entry points ship."""
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", libpath], capture_output=True, text=True, check=True).stdout
    exported = sorted({ln.split()[-1] for ln in out.splitlines() if ln.split() and ln.split()[-1].startswith("fastb200_")})
    assert exported == _header_symbols()
    blob = open(libpath, "rb").read()
    for knob in (b"FASTB200_TC_DBG", b"FASTB200_WG_SWAP", b"FASTB200_DGRAD", b"FASTB200_TR_GROUP"):
        assert knob not in blob, f"debug environment knob {knob.decode()} is compiled into the shipped library"


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
    # the tensor-core backward validates like its siblings (B, windows inside the trial, item bound) before touching memory
    d = N.ModelDesc()
    d.dim_cnn, d.n_zones, d.window_len, d.n_chan, d.slide_step = 32, 1, 250, 4, 125
    d.zone_size[0] = 4
    rc = lib.fastb200_head_bwd_tc(d, None, None, None, None, 2, 300, 5, 125, None, None, None, None)
    assert rc < 0 and b"windows run past" in lib.fastb200_last_error()
    rc = lib.fastb200_head_bwd_tc(d, None, None, None, None, -1, 800, 5, 125, None, None, None, None)
    assert rc < 0 and b"bad B" in lib.fastb200_last_error()
    rc = lib.fastb200_head_bwd_tc(d, None, None, None, None, 2, 800, 5, 125, None, None, None, None)
    assert rc < 0 and b"NULL buffer" in lib.fastb200_last_error()


def test_head_support_predicate_is_host_logic(libpath):
    """Which Conv4Layers shapes the fused kernels cover (the rest runs as stock torch CUDA ops in the drop-in)."""
    from fast.models.fast import _fused_head_ok
    assert _fused_head_ok(250, [6, 9, 6, 7, 7, 10, 15, 4], 32, True)          # production
    assert _fused_head_ok(250, [4, 2, 2], 16, True) and _fused_head_ok(250, [1], 8, True)
    assert not _fused_head_ok(250, [4], 64, False)                              # reference test_model.py:278-284 uses dim 64
    assert not _fused_head_ok(250, [33], 32, False)                             # zone larger than the kernels' electrode tile
    assert _fused_head_ok(500, [4], 32, False) and not _fused_head_ok(500, [4], 32, True)      # backward keeps more rows resident
    assert not _fused_head_ok(4000, [4], 32, False)
    assert not _fused_head_ok(250, [2] * 17, 32, False)                         # more zones than FASTB200_MAX_ZONES


def test_fused_filter_support_predicate_is_host_logic(libpath):
    """fastb200_sosfiltfilt_resample_supported runs no kernel: the 250 -> 200 Hz chain with scipy's default FIR and padlen is
    the fused kernel's shape; another ratio, FIR alignment, a padlen shorter than the resampler's edge reach, or a row no
    longer than padlen are not (the caller then chains the two kernels)."""
    import numpy as np
    from fast import _native as N
    from fast.data.filters import _kaiser_lowpass
    lib = N.lib()
    dp = ctypes.POINTER(ctypes.c_double)

    def fir(up, down, ntaps=None):
        h = _kaiser_lowpass(up, down) if ntaps is None else np.hamming(ntaps) * up
        half = (len(h) - 1) // 2
        pre = down - half % down
        hp = np.ascontiguousarray(np.concatenate([np.zeros(pre), h]))
        return hp, (half + pre) // down

    h, npr = fir(4, 5)
    sup = lambda T, ns, pad, T_out, h_, up, down, npr_: lib.fastb200_sosfiltfilt_resample_supported(
        T, ns, pad, T_out, h_.ctypes.data_as(dp), len(h_), up, down, npr_)
    assert sup(795, 4, 27, 636, h, 4, 5, npr) == 1
    assert sup(800, 4, 27, 640, h, 4, 5, npr) == 1
    assert sup(795, 1, 27, 636, h, 4, 5, npr) == 1
    assert sup(795, 4, 0, 636, h, 4, 5, npr) == 0            # no odd extension to hold the resampler's zero pads
    assert sup(795, 4, 9, 636, h, 4, 5, npr) == 0            # 9 < the 13-sample reach of the 26-tap sub-filters
    assert sup(27, 4, 27, 22, h, 4, 5, npr) == 0             # T must exceed padlen (scipy's ValueError)
    assert sup(795, 5, 27, 636, h, 4, 5, npr) == 0           # more sections than the kernel carries
    h54, npr54 = fir(5, 4)
    assert sup(795, 4, 27, 994, h54, 5, 4, npr54) == 0       # other ratios take the two-kernel chain
    h63, npr63 = fir(4, 5, 63)
    assert sup(795, 4, 27, 636, h63, 4, 5, npr63) == 0       # a FIR whose centre lands on another polyphase alignment
    h61, npr61 = fir(4, 5, 61)
    assert sup(795, 4, 27, 636, h61, 4, 5, npr61) == 1       # same alignment as the default, fewer taps
    rc = lib.fastb200_sosfiltfilt_resample(None, 4, None, 4, 0, 795, None, 4, 27, 636, None, 106, 4, 5, 11, None)
    assert rc < 0 and b"NULL" in lib.fastb200_last_error()


def test_alt_head_descriptor_validation_without_a_gpu(libpath):
    """fastb200_alt_head_fwd rejects malformed chains before any launch (messages through fastb200_last_error)."""
    from fast import _native as N
    lib = N.lib()
    assert lib.fastb200_sizeof_alt_head() == ctypes.sizeof(N.AltHeadDesc)
    d = N.AltHeadDesc()
    assert lib.fastb200_alt_head_fwd(None, None, None, 1, 8, 500, 1, 125, None, None) < 0
    d.window_len, d.n_zones, d.n_stages = 250, 1, 1
    d.zone_size[0], d.zone_off[0] = 4, 0
    st = d.stage[0]
    st.kind, st.out_ch, st.k, st.pad, st.act, st.pool_kind, st.pool, st.scale_off = 1, 8, 3, 0, 0, 0, 1, -1      # depth-wise first stage
    assert lib.fastb200_alt_head_fwd(d, None, None, 1, 8, 500, 1, 125, None, None) < 0
    assert b"must be dense" in lib.fastb200_last_error()
    st.kind, st.pool_kind, st.pool = 0, 1, 3                                                                       # pool size not instantiated
    assert lib.fastb200_alt_head_fwd(d, None, None, 1, 8, 500, 1, 125, None, None) < 0
    assert b"pool size 3" in lib.fastb200_last_error()
    st.pool = 2
    d.final_kind, d.lin_in, d.lin_out = 2, 999, 4                                                                  # flatten + Linear of the wrong width
    assert lib.fastb200_alt_head_fwd(d, None, None, 1, 8, 500, 1, 125, None, None) < 0
    assert b"flatten+linear" in lib.fastb200_last_error()
    d.final_kind, d.lin_out = 0, 8
    assert lib.fastb200_alt_head_fwd(d, None, None, 1, 8, 100, 1, 125, None, None) < 0                             # window longer than the trial
    assert b"windows run past" in lib.fastb200_last_error()
    assert lib.fastb200_alt_head_fwd(d, None, None, 0, 8, 500, 1, 125, None, None) == 0                            # empty batch: nothing to do


def test_static_workspace_entry_point_validation_without_a_gpu(libpath):
    from fast import _native as N
    lib = N.lib()
    assert lib.fastb200_head_bwd_tc_static_max_floats() * 4 == 32 << 20
    assert lib.fastb200_head_bwd_tc_static(None, None, None, None, None, 1, 800, 5, 125, None, None) < 0
    assert b"NULL" in lib.fastb200_last_error()
