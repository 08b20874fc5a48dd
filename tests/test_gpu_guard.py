"""Guard-band test of every kernel family through the C ABI (compute-sanitizer is closed on this GPU pool, see
profiles/r02_sanitizer_unavailable.txt): each output / workspace buffer lives inside a larger allocation whose guard bands
hold a sentinel bit pattern; after the call the bands must be untouched and the payload fully written where the kernel
promises to write it.  Ragged sizes on purpose (batch 1 / 3 / 67, 795- and 800-sample trials, 7-window step override)."""
import numpy as np
import pytest
import torch

from conftest import make_prod_cfg

pytestmark = pytest.mark.gpu
GUARD = 16384                      # floats either side (64 KiB)
SENT = 0x7FC0DEAD                  # a quiet-NaN bit pattern no kernel produces


class Guarded:
    """`n` elements of `dtype` with sentinel guard bands either side; `.t` is the payload view handed to the kernel."""

    def __init__(self, n, dtype=torch.float32, fill_payload=True):
        self.esz = torch.empty(0, dtype=dtype).element_size()
        self.n = int(n)
        nbytes = self.n * self.esz
        pad = (-nbytes) % 4
        words = (nbytes + pad) // 4
        self.raw = torch.full((2 * GUARD + words,), SENT, dtype=torch.int32, device="cuda")
        self.t = self.raw[GUARD:GUARD + words].view(torch.uint8)[:nbytes].view(dtype)
        if not fill_payload:
            self.t.zero_()

    def check(self, name, fully_written=False):
        lo, hi = self.raw[:GUARD], self.raw[self.raw.numel() - GUARD:]
        assert bool((lo == SENT).all()), f"{name}: write BEFORE the buffer"
        assert bool((hi == SENT).all()), f"{name}: write PAST the buffer"
        if fully_written and self.esz == 4:
            assert not bool((self.t.view(torch.int32) == SENT).any()), f"{name}: part of the output was never written"


@pytest.mark.parametrize("B,T,step", [(1, 795, 125), (3, 800, 125), (67, 795, 125), (5, 800, 50)])
@pytest.mark.parametrize("n_terms", [2, 3])
def test_head_tensor_core_kernels_stay_inside_their_buffers(B, T, step, n_terms):
    from fast import _native as N
    from fast.models import FAST
    lib = N.lib()
    torch.manual_seed(42)
    m = FAST(make_prod_cfg()).cuda()
    flat, d = m._runtime()
    n_win = (T - 250) // step + 1
    st = N.stream_ptr()
    x = Guarded(B * 64 * T)
    x.t.copy_(torch.randn(B * 64 * T, device="cuda"))
    # ---- packs
    packed_tc = Guarded(lib.fastb200_head_packed_tc_bytes(d, n_terms), torch.uint8)
    N.check(lib.fastb200_head_pack_tc(d, N.ptr(flat), N.ptr(packed_tc.t), n_terms, st), "pack_tc")
    packed = Guarded(lib.fastb200_head_packed_floats(d))
    N.check(lib.fastb200_head_pack(d, N.ptr(flat), N.ptr(packed.t), st), "pack")
    # ---- inference forward
    feat = Guarded(B * n_win * 8 * 32)
    N.check(lib.fastb200_head_fwd_tc(d, N.ptr(packed_tc.t), N.ptr(x.t), B, T, n_win, step, N.ptr(feat.t), n_terms, st), "fwd_tc")
    torch.cuda.synchronize()
    feat.check("head_fwd_tc feat", fully_written=True)
    ref = feat.t.clone()
    # ---- training forward (also writes y2 / y3 / y4)
    feat2 = Guarded(B * n_win * 8 * 32)
    saved = Guarded(lib.fastb200_head_saved_floats(d, B, n_win))
    N.check(lib.fastb200_head_fwd_tc_train(d, N.ptr(packed_tc.t), N.ptr(x.t), B, T, n_win, step, N.ptr(feat2.t), N.ptr(saved.t),
                                           n_terms, st), "fwd_tc_train")
    torch.cuda.synchronize()
    feat2.check("head_fwd_tc_train feat", fully_written=True)
    saved.check("head_fwd_tc_train saved")
    assert torch.equal(feat2.t, ref)                                   # SAVE form computes the same features, bit for bit
    # ---- backward (input-gradient chain + weight gradients), no dx: the all-tensor-core path of training
    dfeat = Guarded(B * n_win * 8 * 32)
    dfeat.t.copy_(torch.randn(dfeat.n, device="cuda"))
    gpacked = Guarded(packed.n, fill_payload=False)
    work = Guarded(lib.fastb200_head_bwd_tc_work_floats(d, B, n_win))
    N.check(lib.fastb200_head_bwd_tc(d, N.ptr(packed.t), N.ptr(x.t), N.ptr(saved.t), N.ptr(dfeat.t), B, T, n_win, step,
                                     N.ptr(gpacked.t), None, N.ptr(work.t), st), "bwd_tc")
    gflat = Guarded(flat.numel(), fill_payload=False)
    N.check(lib.fastb200_head_unpack_grad(d, N.ptr(flat), N.ptr(gpacked.t), N.ptr(gflat.t), st), "unpack")
    torch.cuda.synchronize()
    for name, g in (("x", x), ("packed_tc", packed_tc), ("packed", packed), ("saved", saved), ("dfeat", dfeat), ("gpacked", gpacked),
                    ("bwd work", work), ("gflat", gflat), ("feat", feat)):
        g.check(name)
    assert bool(torch.isfinite(gflat.t).all()) and float(gflat.t.abs().max()) > 0
    # ---- with dx (SHAP path): CUDA-core forward (fp32 activations saved) + CUDA-core backward; the tensor-core backward works
    # on operand images and refuses an input gradient
    dx = Guarded(B * 64 * T, fill_payload=False)
    assert lib.fastb200_head_bwd_tc(d, N.ptr(packed.t), N.ptr(x.t), N.ptr(saved.t), N.ptr(dfeat.t), B, T, n_win, step,
                                    N.ptr(gpacked.t), N.ptr(dx.t), N.ptr(work.t), st) < 0
    feat3 = Guarded(B * n_win * 8 * 32)
    saved32 = Guarded(lib.fastb200_head_saved_floats(d, B, n_win))
    N.check(lib.fastb200_head_fwd(d, N.ptr(packed.t), N.ptr(x.t), B, T, n_win, step, N.ptr(feat3.t), N.ptr(saved32.t), st), "fwd ffma")
    gpacked2 = Guarded(packed.n, fill_payload=False)
    N.check(lib.fastb200_head_bwd(d, N.ptr(packed.t), N.ptr(x.t), N.ptr(saved32.t), N.ptr(dfeat.t), B, T, n_win, step,
                                  N.ptr(gpacked2.t), N.ptr(dx.t), st), "bwd ffma dx")
    torch.cuda.synchronize()
    for name, g in (("dx", dx), ("gpacked (dx path)", gpacked2), ("saved fp32", saved32), ("feat ffma", feat3)):
        g.check(name)
    feat3.check("feat ffma", fully_written=True)
    assert bool(torch.isfinite(dx.t).all())
    assert float((feat3.t - ref).abs().max()) <= 1e-4 * float(ref.abs().max())
    # both paths produce the same weight gradients up to the split-bf16 rounding of the tensor-core path
    a, b = gpacked.t, gpacked2.t
    assert float((a - b).abs().max()) <= 1e-3 * float(b.abs().max()) + 1e-6


@pytest.mark.parametrize("B", [1, 3, 67, 300])
def test_token_path_loss_and_optimiser_stay_inside_their_buffers(B):
    from fast import _native as N
    from fast.models import FAST
    lib = N.lib()
    torch.manual_seed(1)
    m = FAST(make_prod_cfg()).cuda()
    flat, d = m._runtime()
    st = N.stream_ptr()
    n_win = 5
    feat = Guarded(B * n_win * 8 * 32)
    feat.t.copy_(torch.randn(feat.n, device="cuda") * 0.1)
    logits = Guarded(B * 5)
    saved = Guarded(max(1, lib.fastb200_transformer_saved_floats(d, B, n_win)))
    step_dev = torch.zeros(1, dtype=torch.int32, device="cuda")
    N.check(lib.fastb200_transformer_fwd(d, N.ptr(flat), N.ptr(feat.t), B, n_win, N.MODE_DEFAULT, 0.1, 7, 0, N.ptr(step_dev),
                                         N.ptr(logits.t), N.ptr(saved.t), st), "tr_fwd")
    y = torch.randint(0, 5, (B,), device="cuda").to(torch.uint8)
    loss, ncorr = Guarded(1), Guarded(1, torch.int32)
    dlogits, pred = Guarded(B * 5), Guarded(B, torch.int64)
    N.check(lib.fastb200_ce_fwd_bwd(N.ptr(logits.t), N.ptr(y), 1, B, 5, 1.0, N.ptr(loss.t), N.ptr(dlogits.t), N.ptr(pred.t),
                                    N.ptr(ncorr.t), st), "ce")
    grad = Guarded(flat.numel(), fill_payload=False)
    dfeat = Guarded(feat.n)
    N.check(lib.fastb200_transformer_bwd(d, N.ptr(flat), N.ptr(feat.t), N.ptr(saved.t), N.ptr(dlogits.t), B, n_win, N.MODE_DEFAULT,
                                         0.1, 7, 0, N.ptr(step_dev), N.ptr(grad.t), N.ptr(dfeat.t), st), "tr_bwd")
    p = Guarded(flat.numel()); p.t.copy_(flat)
    mom, var = Guarded(flat.numel(), fill_payload=False), Guarded(flat.numel(), fill_payload=False)
    lr = torch.tensor([5e-4], device="cuda")
    N.check(lib.fastb200_adamw_flat(N.ptr(p.t), N.ptr(grad.t), N.ptr(mom.t), N.ptr(var.t), flat.numel(), N.ptr(lr), 0, 0.9, 0.999,
                                    1e-8, 0.01, N.ptr(step_dev), 1, st), "adamw")
    torch.cuda.synchronize()
    logits.check("logits", True); dlogits.check("dlogits", True); dfeat.check("dfeat", True)
    for name, g in (("feat", feat), ("tr saved", saved), ("loss", loss), ("n_correct", ncorr), ("pred", pred), ("grad", grad),
                    ("params", p), ("exp_avg", mom), ("exp_avg_sq", var)):
        g.check(name)
    assert int(step_dev) == 1 and bool(torch.isfinite(p.t).all()) and 0 <= int(ncorr.t) <= B


@pytest.mark.parametrize("rows,T", [(1, 795), (67, 800), (640, 795), (33, 120)])
def test_filter_kernels_stay_inside_their_buffers(rows, T):
    import ctypes as C
    from fast import _native as N
    from fast.data import filters as F
    lib = N.lib()
    sos = F.bandpass_sos() if T > 200 else np.ascontiguousarray(F.bandpass_sos(2, 4.0, 40.0, 250.0))
    edge = F.default_padlen(sos)
    x = Guarded(rows * T)
    x.t.copy_(torch.randn(rows * T, device="cuda") * 20 + 30)
    y32, y64 = Guarded(rows * T), Guarded(rows * T, torch.float64)
    sp = sos.ctypes.data_as(C.POINTER(C.c_double))
    st = N.stream_ptr()
    N.check(lib.fastb200_sosfiltfilt(N.ptr(x.t), 4, N.ptr(y32.t), 4, rows, T, sp, sos.shape[0], edge, st), "sosfiltfilt f32")
    N.check(lib.fastb200_sosfiltfilt(N.ptr(x.t), 4, N.ptr(y64.t), 8, rows, T, sp, sos.shape[0], edge, st), "sosfiltfilt f64")
    torch.cuda.synchronize()
    y32.check("sosfiltfilt f32 out", True); y64.check("sosfiltfilt f64 out"); x.check("filter input")
    # the Python wrappers (fused band-pass + resample, and the two-kernel chain) on a guarded input view
    xin = x.t.view(rows, T)
    out = F.sosfiltfilt_resample(sos, xin, 4, 5, out_dtype=torch.float32)
    chain = F.resample_poly(F.sosfiltfilt(sos, xin, out_dtype=torch.float64), 4, 5, out_dtype=torch.float32)
    torch.cuda.synchronize()
    x.check("filter input after the fused kernel")
    assert out.shape == (rows, -(-T * 4 // 5)) and torch.equal(out, chain)


@pytest.mark.parametrize("head", ["CVBlock", "EEGNet_Encoder", "HeadConv_Paper_Version"])
def test_alternative_head_kernel_stays_inside_its_output(head):
    """fastb200_alt_head_fwd through the C ABI: every feature written, nothing outside the buffer (ragged batch, step override),
    and the descriptor checks reject a window length the flatten + Linear head cannot take."""
    from fast import _native as N
    from fast.models import FAST
    torch.manual_seed(1)
    m = FAST(make_prod_cfg(head=head, dropout=0.0)).cuda().eval()
    runner = m.head._alt_runner()
    B, T, step = 7, 800, 50
    x = Guarded(B * 64 * T)
    x.t.copy_(torch.randn(B * 64 * T, device="cuda"))
    with torch.no_grad():
        ref = m.forward_head(x.t.view(B, 64, T), step)                      # builds the descriptor + packed parameters
    n_win = (T - 250) // step + 1
    F_out = runner.desc.lin_out
    feat = Guarded(B * n_win * 8 * F_out)
    N.check(N.lib().fastb200_alt_head_fwd(runner.desc, N.ptr(runner.packed), N.ptr(x.t), B, 64, T, n_win, step, N.ptr(feat.t),
                                          N.stream_ptr()), "alt_head_fwd")
    torch.cuda.synchronize()
    feat.check("alt_head_fwd feat", fully_written=True)
    x.check("x")
    assert torch.equal(feat.t.view_as(ref), ref)
    lib = N.lib()
    assert lib.fastb200_alt_head_fwd(runner.desc, N.ptr(runner.packed), N.ptr(x.t), B, 64, 300, 3, 125, N.ptr(feat.t), N.stream_ptr()) < 0
    assert b"windows run past" in lib.fastb200_last_error()
    if head == "CVBlock":                                                   # flatten + Linear(256): only 250-sample windows fit
        import copy
        d2 = copy.copy(runner.desc)
        d2.window_len = 200
        assert lib.fastb200_alt_head_fwd(d2, N.ptr(runner.packed), N.ptr(x.t), B, 64, T, 2, 125, N.ptr(feat.t), N.stream_ptr()) < 0
        assert b"flatten+linear" in lib.fastb200_last_error()
