"""fastb200_head_bwd_tc_static: the allocation-free form of the tensor-core head backward (its workspace is a file-scope
device array inside the library) against fastb200_head_bwd_tc with a caller workspace -- through the C ABI, two calls in a row
(the second one starts from the first one's workspace contents).  Collected last on purpose (file name)."""
import pytest
import torch

from conftest import make_prod_cfg

pytestmark = pytest.mark.gpu
TOL = 1e-4                           # the tolerance of tests/test_gpu_parity.py (north_star: 1e-4 of the tensor's largest magnitude)


def rel_err(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def test_static_workspace_backward_matches_caller_workspace_backward_twice():
    from fast import _native as N
    from fast.models import FAST
    lib = N.lib()
    torch.manual_seed(11)
    m = FAST(make_prod_cfg()).cuda()
    flat, d = m._runtime()
    assert lib.fastb200_head_bwd_tc_supported(d) == 1
    T, step, n_terms = 800, 125, 2
    n_win = (T - 250) // step + 1
    st = N.stream_ptr()
    cap = lib.fastb200_head_bwd_tc_static_max_floats()
    assert cap * 4 == 32 << 20

    packed_tc = torch.empty(lib.fastb200_head_packed_tc_bytes(d, n_terms), dtype=torch.uint8, device="cuda")
    N.check(lib.fastb200_head_pack_tc(d, N.ptr(flat), N.ptr(packed_tc), n_terms, st), "pack_tc")
    packed = torch.empty(lib.fastb200_head_packed_floats(d), device="cuda")
    N.check(lib.fastb200_head_pack(d, N.ptr(flat), N.ptr(packed), st), "pack")

    for call, B in enumerate((8, 5)):                   # second call: a smaller batch over the first call's leftovers
        g = torch.Generator().manual_seed(100 + call)
        x = (torch.randn(B, 64, T, generator=g) * 3.0).cuda()
        dfeat = torch.randn(B, n_win, 8, 32, generator=g).cuda()
        feat = torch.empty(B, n_win, 8, 32, device="cuda")
        saved = torch.empty(lib.fastb200_head_saved_floats(d, B, n_win), device="cuda")
        N.check(lib.fastb200_head_fwd_tc_train(d, N.ptr(packed_tc), N.ptr(x), B, T, n_win, step, N.ptr(feat), N.ptr(saved),
                                               n_terms, st), "fwd_tc_train")
        need = lib.fastb200_head_bwd_tc_work_floats(d, B, n_win)
        assert 0 < need <= cap
        work = torch.empty(need, device="cuda")
        g_ref = torch.zeros_like(packed)
        N.check(lib.fastb200_head_bwd_tc(d, N.ptr(packed), N.ptr(x), N.ptr(saved), N.ptr(dfeat), B, T, n_win, step,
                                         N.ptr(g_ref), None, N.ptr(work), st), "bwd_tc")
        g_static = torch.zeros_like(packed)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        N.check(lib.fastb200_head_bwd_tc_static(d, N.ptr(packed), N.ptr(x), N.ptr(saved), N.ptr(dfeat), B, T, n_win, step,
                                                N.ptr(g_static), st), "bwd_tc_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert bool(torch.isfinite(g_static).all()) and float(g_ref.abs().max()) > 0
        e = rel_err(g_static, g_ref)
        assert e < TOL, (call, B, e)
        # and in parameter layout, tensor by tensor
        gf_ref, gf_static = torch.zeros_like(flat), torch.zeros_like(flat)
        N.check(lib.fastb200_head_unpack_grad(d, N.ptr(flat), N.ptr(g_ref), N.ptr(gf_ref), st), "unpack")
        N.check(lib.fastb200_head_unpack_grad(d, N.ptr(flat), N.ptr(g_static), N.ptr(gf_static), st), "unpack")
        torch.cuda.synchronize()
        assert rel_err(gf_static, gf_ref) < TOL


def test_static_workspace_backward_rejects_sizes_beyond_its_array():
    from fast import _native as N
    from fast.models import FAST
    lib = N.lib()
    m = FAST(make_prod_cfg()).cuda()
    _, d = m._runtime()
    B, T, step, n_win = 9, 800, 125, 5
    assert lib.fastb200_head_bwd_tc_work_floats(d, B, n_win) > lib.fastb200_head_bwd_tc_static_max_floats()
    rc = lib.fastb200_head_bwd_tc_static(d, None, None, None, None, B, T, n_win, step, None, N.stream_ptr())
    assert rc < 0
    msg = lib.fastb200_last_error()
    assert b"static workspace holds" in msg and b"32 MiB" in msg
    assert lib.fastb200_head_bwd_tc_static(d, None, None, None, None, 0, T, n_win, step, None, N.stream_ptr()) == 0   # empty batch
