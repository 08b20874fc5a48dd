"""Parity of the CUDA path (through the C ABI) with the CPU oracle and the golden vectors of the real reference.

Tolerances (SURVEY.md section 8d / BASELINE.json north_star): bit-exact for the index kernels and the trial re-layout;
max|a-b|/max|b| <= 1e-4 for logits, head features and every gradient tensor; <= 1e-5 for filter outputs.
"""
import hashlib

import numpy as np
import pytest
import torch

from conftest import fingerprint, make_prod_cfg, make_small_cfg, rel_err, small_state_dict
from oracle import fast_oracle as O
from oracle import filter_oracle as FO

pytestmark = pytest.mark.gpu
TOL = 1e-4          # logits / features / gradients (north_star)
TOL_FILTER = 1e-5   # filter outputs (north_star)


def _sha(t):
    return hashlib.sha256(t.detach().cpu().contiguous().numpy().tobytes()).hexdigest()


def _small_model(golden, **over):
    from fast.models import FAST
    m = FAST(make_small_cfg(**over))
    m.load_state_dict(small_state_dict(golden))
    return m.cuda()


def _prod_model(**over):
    from fast.models import FAST
    torch.manual_seed(42)
    return FAST(make_prod_cfg(**over)).cuda()


def _prod_x():
    return torch.randn(4, 64, 800, generator=torch.Generator().manual_seed(43))


# ---------------------------------------------------------------------------------------------- index kernels
def test_window_gather_bit_exact(golden):
    from fast import _native as N
    cfg = make_prod_cfg()
    x = _prod_x()
    chan = torch.tensor(np.concatenate(list(O.zone_index(cfg).values())), dtype=torch.int32)
    xd, cd = x.cuda(), chan.cuda()
    out = torch.empty(4 * 5, 64, 250, device='cuda')
    N.check(N.lib().fastb200_window_gather(N.ptr(xd), N.ptr(cd), N.ptr(out), 4, 64, 800, 250, 125, 5, 64, N.stream_ptr()), "wg")
    xw = O.window_unfold(x, 250, 125)
    off = 0
    for area, idx in O.zone_index(cfg).items():
        got = out[:, off:off + len(idx)]
        assert torch.equal(got.cpu(), O.zone_gather(xw, idx))
        assert _sha(got) == str(golden[f"prod/gather_sha/{area}"])
        off += len(idx)
    # identity selection = the plain unfold
    ident = torch.arange(64, dtype=torch.int32, device='cuda')
    N.check(N.lib().fastb200_window_gather(N.ptr(xd), N.ptr(ident), N.ptr(out), 4, 64, 800, 250, 125, 5, 64, N.stream_ptr()), "wg")
    assert _sha(out) == str(golden["prod/unfold_sha"])
    # ragged: 795 samples, stride override, single window
    x2 = x[:, :, :795].contiguous().cuda()
    n = (795 - 250) // 50 + 1
    out2 = torch.empty(4 * n, 64, 250, device='cuda')
    N.check(N.lib().fastb200_window_gather(N.ptr(x2), N.ptr(ident), N.ptr(out2), 4, 64, 795, 250, 50, n, 64, N.stream_ptr()), "wg")
    assert torch.equal(out2.cpu(), O.window_unfold(x[:, :, :795], 250, 50))


# ---------------------------------------------------------------------------------------------- small config
@pytest.mark.parametrize("mode", ["default", "train_head", "train_transformer"])
def test_small_logits_loss_grads(golden, mode):
    m = _small_model(golden)
    m.train()                                                     # dropout = 0.0 in this config
    x, y = torch.from_numpy(golden["small/x"]).cuda(), torch.from_numpy(golden["small/y"]).cuda()
    logits = m(x, forward_mode=mode)
    loss = torch.nn.CrossEntropyLoss()(logits, y)
    loss.backward()
    assert logits.dtype == torch.float32 and logits.shape == (3, 3)
    assert rel_err(logits, golden[f"small/{mode}/logits"]) < TOL
    assert abs(float(loss.detach()) - float(golden[f"small/{mode}/loss"])) < 1e-5
    for k, p in m.named_parameters():
        gk = f"small/{mode}/grad/{k}"
        if gk in golden.files:
            assert p.grad is not None, k
            assert rel_err(p.grad, golden[gk]) < TOL, (k, rel_err(p.grad, golden[gk]))
        else:
            assert p.grad is None, f"{k} must get no gradient in mode {mode} (reference tests/test_model.py:155-164)"


def test_small_features_step_override_and_input_grad(golden):
    m = _small_model(golden).eval()
    x = torch.from_numpy(golden["small/x"]).cuda()
    assert rel_err(m.forward_head(x), golden["small/feat"]) < TOL
    f50 = m.forward_head(x, step_override=50)
    assert tuple(f50.shape) == golden["small/feat_step50"].shape
    assert rel_err(f50, golden["small/feat_step50"]) < TOL
    xr = x.clone().requires_grad_(True)
    m(xr).logsumexp(1).sum().backward()
    assert rel_err(xr.grad, golden["small/dx_lse"]) < TOL
    with pytest.raises(RuntimeError):
        m.forward_transformer(f50)                                # pos_embedding broadcast fails in the reference too


# ---------------------------------------------------------------------------------------------- production config
def test_prod_forward(golden):
    m = _prod_model().eval()
    x = _prod_x().cuda()
    with torch.no_grad():
        feat = m.forward_head(x)
        assert feat.shape == (4, 5, 8, 32)
        assert rel_err(feat, golden["prod/feat"]) < TOL
        logits = m(x)
        assert rel_err(logits, golden["prod/logits"]) < TOL
        l795 = m(x[:, :, :795].contiguous())
        assert torch.equal(l795, logits)                          # samples 750.. are never read
        assert rel_err(m(x, forward_mode='train_head'), golden["prod/train_head_logits"]) < TOL
        # inputs x30: the head dominates the logits (the regime where reduced-precision operands would show)
        assert rel_err(m(x * 30.0), golden["prod/logits_x30"]) < TOL
        # batch of 1, and micro-batched head == full head (reference tests/test_model.py:69, 202-215)
        assert rel_err(m(x[:1]), golden["prod/logits"][:1]) < TOL
        assert torch.equal(m.batched_forward_head(x, 125, 1), m.forward_head(x, 125))


@pytest.fixture
def head_backward():
    from fast.models import fast as FM
    old = FM.get_head_backward()
    yield FM.set_head_backward
    FM.set_head_backward(old)


TRAIN_PATHS = [("ffma", "fp32"), ("tc", "bf16x2"), ("tc", "bf16x3")]   # (head backward, head forward arithmetic)


@pytest.mark.parametrize("bwd,prec", TRAIN_PATHS)
def test_prod_backward(golden, head_backward, head_precision, bwd, prec):
    head_backward(bwd)                                            # 'ffma': CUDA-core forward + backward; 'tc': tcgen05 forward
    head_precision(prec)                                          # (saves y2..y4), input-gradient chain and weight gradients
    m = _prod_model().eval()                                      # golden gradients were taken in eval mode
    x = _prod_x().cuda().requires_grad_(True)
    y = torch.from_numpy(golden["prod/y"]).cuda()
    loss = torch.nn.CrossEntropyLoss()(m(x), y)
    loss.backward()
    assert abs(float(loss) - float(golden["prod/loss"])) < 1e-5
    worst = 0.0
    for i, (k, p) in enumerate(m.named_parameters()):
        assert p.grad is not None, k
        gfp, mine = golden["prod/grad_fp"][i], fingerprint(p.grad)
        scale = max(gfp[3], 1e-30)
        assert abs(mine[3] - gfp[3]) / scale < 1e-3, k
        assert abs(mine[2] - gfp[2]) / (scale * np.sqrt(p.numel())) < 1e-3, k
        if f"prod/grad/{k}" in golden.files:
            e = rel_err(p.grad, golden[f"prod/grad/{k}"])
            worst = max(worst, e)
            assert e < TOL, (k, e)
    assert rel_err(x.grad[0, 3], golden["prod/dx_b0_c3"]) < TOL
    dfp = fingerprint(x.grad)
    assert abs(dfp[3] - golden["prod/dx_fp"][3]) / golden["prod/dx_fp"][3] < 1e-3


@pytest.mark.parametrize("bwd,prec", TRAIN_PATHS)
def test_prod_backward_vs_oracle_all_tensors(head_backward, head_precision, bwd, prec):
    """Every one of the 94 gradient tensors against the oracle's autograd on a fresh input (not only the golden subset)."""
    head_backward(bwd)
    head_precision(prec)
    m = _prod_model().eval()
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(77)
    x = torch.randn(3, 64, 795, generator=g) * 5.0
    y = torch.tensor([3, 0, 4], dtype=torch.uint8)
    _, logits_o, grads_o, dx_o = O.loss_and_grads(sd, make_prod_cfg(), x, y, input_grad=True)
    xr = x.cuda().requires_grad_(True)
    logits = m(xr)
    torch.nn.CrossEntropyLoss()(logits, y.cuda()).backward()
    assert rel_err(logits, logits_o) < TOL
    for k, p in m.named_parameters():
        assert rel_err(p.grad, grads_o[k]) < TOL, (k, rel_err(p.grad, grads_o[k]))
    assert rel_err(xr.grad, dx_o) < TOL


def test_shapes_modes_and_edge_cases(golden):
    m = _prod_model()
    x = torch.randn(4, 64, 800, device='cuda')
    for mode in ('default', 'train_head', 'train_transformer'):
        out = m(x, forward_mode=mode)
        assert out.shape == (4, 5) and out.dtype == torch.float32 and torch.isfinite(out).all()
    with pytest.raises(NotImplementedError):
        m(x, forward_mode='nope')
    assert m(x[:0]).shape == (0, 5)                               # empty batch
    with pytest.raises(RuntimeError):
        m(torch.randn(2, 64, 100, device='cuda'))                 # shorter than one window
    with pytest.raises(RuntimeError):
        m(torch.randn(2, 32, 800, device='cuda'))                 # wrong electrode count
    # non-contiguous and float64 inputs are accepted like the reference accepts them (cast / copy on the way in)
    xt = torch.randn(4, 800, 64, device='cuda').transpose(1, 2)
    m.eval()
    with torch.no_grad():
        assert torch.equal(m(xt), m(xt.contiguous()))
    # zero input / huge input stay finite
    with torch.no_grad():
        assert torch.isfinite(m(torch.zeros(2, 64, 800, device='cuda'))).all()
        assert torch.isfinite(m(torch.full((2, 64, 800), 1e4, device='cuda'))).all()


def test_conv4layers_and_head_standalone(golden):
    """Conv4Layers(channels, dim)(x) and Head(...)(x) called on their own (reference tests/test_model.py:256-285, 362-404)."""
    from fast.models import Conv4Layers, Head
    torch.manual_seed(5)
    enc = Conv4Layers(6, 32).cuda()
    x = torch.randn(7, 6, 300, device='cuda')
    out = enc(x)
    assert out.shape == (7, 32)
    sd = {k: v.detach().cpu() for k, v in enc.state_dict().items()}
    ref = O.conv4layers(x.cpu(), sd["cnn1.weight"], sd["cnn1.bias"], sd["cnn2.weight"], sd["cnn3.weight"], sd["cnn4.weight"])
    assert rel_err(out, ref) < TOL
    out.sum().backward()
    assert all(p.grad is not None for p in enc.parameters())
    cfg = make_small_cfg()
    head = Head('Conv4Layers', cfg.electrodes, cfg.zone_dict, 16).cuda()
    assert head.index_dict['Frontal'].tolist() == [0, 1, 2, 3]
    assert head(torch.randn(5, 8, 250, device='cuda')).shape == (5, 3, 16)


def test_batch_invariance_large(golden):
    """Size-independent property at BASELINE config-2 scale: a trial's logits do not depend on its batch."""
    m = _prod_model().eval()
    x = torch.randn(2048, 64, 795, device='cuda', generator=torch.Generator(device='cuda').manual_seed(1))
    with torch.no_grad():
        full = m(x)
        parts = torch.cat([m(c) for c in x.split(300)])
    assert torch.equal(full, parts)
    assert torch.isfinite(full).all()
    # hundreds of items per persistent CTA: the tensor-core head (default) against the exact FFMA head on every trial
    from fast.models import fast as FM
    old = FM.get_head_precision()
    try:
        feats = {}
        for mode in ("fp32", "bf16x3", "bf16x2"):
            FM.set_head_precision(mode)
            with torch.no_grad():
                feats[mode] = m.forward_head(x)
        assert rel_err(feats["bf16x3"], feats["fp32"]) < 5e-6
        assert rel_err(feats["bf16x2"], feats["fp32"]) < 3e-5
    finally:
        FM.set_head_precision(old)


# ---------------------------------------------------------------------------------------------- loss / optimiser
def test_ce_fwd_bwd_vs_torch():
    from fast import _native as N
    lib = N.lib()
    for B, NC, dt in ((64, 5, torch.uint8), (1, 5, torch.int64), (301, 3, torch.uint8)):
        g = torch.Generator().manual_seed(B)
        logits = (torch.randn(B, NC, generator=g) * 3).requires_grad_(True)
        y = torch.randint(0, NC, (B,), generator=g).to(dt)
        ref = torch.nn.CrossEntropyLoss()(logits, y if dt == torch.int64 else y)
        ref.backward()
        ld, yd = logits.detach().cuda(), y.cuda()
        loss = torch.zeros(1, device='cuda'); dl = torch.empty(B, NC, device='cuda')
        pred = torch.empty(B, device='cuda', dtype=torch.int64); nc = torch.zeros(1, device='cuda', dtype=torch.int32)
        N.check(lib.fastb200_ce_fwd_bwd(N.ptr(ld), N.ptr(yd), yd.element_size(), B, NC, 1.0, N.ptr(loss), N.ptr(dl), N.ptr(pred),
                                        N.ptr(nc), N.stream_ptr()), "ce")
        assert abs(float(loss) - float(ref)) < 1e-6 * max(1.0, abs(float(ref)))
        assert rel_err(dl, logits.grad) < 1e-5
        assert torch.equal(pred.cpu(), logits.detach().argmax(1))
        assert int(nc) == int((logits.detach().argmax(1) == y.long()).sum())


def test_adamw_flat_vs_torch_golden(golden):
    from fast import _native as N
    lib = N.lib()
    p = torch.from_numpy(golden["train/adamw_p0"]).cuda()
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    step = torch.zeros(1, device='cuda', dtype=torch.int32)
    lr = torch.from_numpy(golden["train/adamw_lr"].astype(np.float32)).cuda()
    for s in range(4):
        g = torch.from_numpy(golden["train/adamw_g"][s]).cuda()
        N.check(lib.fastb200_adamw_flat(N.ptr(p), N.ptr(g), N.ptr(m), N.ptr(v), p.numel(), N.ptr(lr), 1, 0.9, 0.999, 1e-8, 0.01,
                                        N.ptr(step), 1, N.stream_ptr()), "adamw")
        assert int(step) == s + 1
        assert rel_err(p, golden[f"train/adamw_p{s + 1}"]) < 1e-6          # includes the lr == 0 step


# ---------------------------------------------------------------------------------------------- data side
def test_trials_from_mat_bit_exact(golden):
    from fast.data import trials_from_mat
    xm = np.random.default_rng(5).standard_normal((795, 64, 6)) * 30.0
    out = trials_from_mat(xm)
    assert out.shape == (6, 64, 800) and out.dtype == torch.float32
    assert _sha(out) == str(golden["data/mat_sha"])
    out32 = trials_from_mat(xm.astype(np.float32), target_len=795)
    assert np.array_equal(out32.cpu().numpy(), np.transpose(xm.astype(np.float32), (2, 1, 0)))
    assert trials_from_mat(xm[:, :, :0]).shape == (0, 64, 800)


def test_sosfiltfilt_vs_scipy_golden(filter_golden):
    from fast.data import sosfiltfilt
    g = filter_golden
    x = torch.from_numpy(g["x"]).cuda()
    y64 = sosfiltfilt(g["sos"], x, out_dtype=torch.float64)
    assert rel_err(y64, g["y_sosfiltfilt"]) < 1e-9                 # fp64 scan against scipy's fp64 serial recursion
    y32 = sosfiltfilt(g["sos"], x)
    assert y32.dtype == torch.float32 and rel_err(y32, g["y_sosfiltfilt"]) < TOL_FILTER
    ylp = sosfiltfilt(g["sos_lp"], x[:, :, :120].contiguous(), out_dtype=torch.float64)
    assert rel_err(ylp, g["y_lp"]) < 1e-9
    with pytest.raises(ValueError):
        sosfiltfilt(g["sos"], x[:, :, :27].contiguous())
    # DC-heavy input (offset 3000): the case an fp32 recursion fails
    rng = np.random.default_rng(9)
    xd = (rng.standard_normal((5, 64, 795)) * 20 + 3000).astype(np.float32)
    ref = FO.sosfiltfilt(g["sos"], xd)
    assert rel_err(sosfiltfilt(g["sos"], torch.from_numpy(xd).cuda()), ref) < TOL_FILTER


def test_resample_poly_vs_scipy_golden(filter_golden):
    from fast.data import bandpass_resample, resample_poly
    g = filter_golden
    y = torch.from_numpy(g["y_sosfiltfilt"]).cuda()
    r = resample_poly(y, 4, 5)
    assert r.shape[-1] == 636 and rel_err(r, g["y_resample_4_5"]) < 1e-11
    assert rel_err(resample_poly(y, 125, 128), g["y_resample_125_128"]) < 1e-11
    assert rel_err(resample_poly(y, 8, 10), g["y_resample_4_5"]) < 1e-11      # gcd reduction
    x = torch.from_numpy(g["x"]).cuda()
    assert rel_err(resample_poly(x, 4, 5, out_dtype=torch.float64), g["y_resample_4_5_f32in"]) < 1e-6
    assert torch.equal(resample_poly(x, 3, 3), x)
    out = bandpass_resample(x)
    assert out.dtype == torch.float32 and rel_err(out, g["y_resample_4_5"]) < TOL_FILTER


def test_fused_filter_resample_equals_the_two_call_chain(filter_golden):
    """The one-kernel band-pass + 4/5 resample (filtered row stays in shared memory) gives the bits of the two-kernel chain
    with its float64 intermediate; the phase-major 4/5 resampler agrees with the generic kernel (forced by a custom
    window of another alignment) and with the oracle on ragged lengths; unsupported shapes fall back to two calls."""
    from fast.data import resample_poly, sosfiltfilt, sosfiltfilt_resample
    from fast.data.filters import _kaiser_lowpass
    g = filter_golden
    sos = g["sos"]
    x = torch.from_numpy(g["x"]).cuda()
    for odt in (torch.float32, torch.float64):
        fused = sosfiltfilt_resample(sos, x, 4, 5, out_dtype=odt)
        chain = resample_poly(sosfiltfilt(sos, x, out_dtype=torch.float64), 4, 5, out_dtype=odt)
        assert fused.dtype == odt and torch.equal(fused, chain)
    assert rel_err(sosfiltfilt_resample(sos, x, 4, 5, out_dtype=torch.float64), g["y_resample_4_5"]) < 1e-9
    xd = x.double()
    assert torch.equal(sosfiltfilt_resample(sos, xd, 4, 5), resample_poly(sosfiltfilt(sos, xd), 4, 5))
    # fewer sections, ragged lengths (T_out not a multiple of 4), one row
    for T in (100, 333, 794, 1001):
        xr = torch.randn(3, 5, T, device='cuda', dtype=torch.float64, generator=torch.Generator(device='cuda').manual_seed(T))
        ref = FO.resample_poly(FO.sosfiltfilt(g["sos_lp"], xr.cpu().numpy()), 4, 5)
        out = sosfiltfilt_resample(g["sos_lp"], xr, 4, 5)
        assert out.shape[-1] == -(-T * 4 // 5) and rel_err(out, ref) < 1e-9, T
        assert rel_err(resample_poly(xr, 4, 5), FO.resample_poly(xr.cpu().numpy(), 4, 5)) < 1e-11, T
    # a user FIR of a different length / alignment takes the generic kernel: same answer as scipy
    from scipy import signal
    y = torch.from_numpy(g["y_sosfiltfilt"]).cuda()
    for ntaps in (61, 63, 141):           # 61: phase-major kernel (same alignment as the default); 63, 141: generic kernel
        w = signal.firwin(ntaps, 0.2, window=('kaiser', 5.0))
        ref = signal.resample_poly(g["y_sosfiltfilt"], 4, 5, axis=-1, window=w)
        assert rel_err(resample_poly(y, 4, 5, window=w), ref) < 1e-11, ntaps
    # the default FIR passed explicitly is the default path
    assert torch.equal(resample_poly(y, 4, 5, window=_kaiser_lowpass(4, 5) / 4), resample_poly(y, 4, 5))
    # unsupported by the fused kernel (ratio 125/128; padlen 0): two-call chain, same contract
    assert rel_err(sosfiltfilt_resample(sos, x, 125, 128, out_dtype=torch.float64), g["y_resample_125_128"]) < 1e-9
    p0 = sosfiltfilt_resample(sos, xd, 4, 5, padlen=0)
    assert torch.equal(p0, resample_poly(sosfiltfilt(sos, xd, padlen=0), 4, 5))


def test_bandpass_resample_host_stream_equals_direct():
    """The host-to-host streaming form (copies of neighbouring blocks overlap the kernel) returns, block for block and in
    order, the bits of the direct call; ragged block sizes and an empty iterable are fine."""
    from fast.data import bandpass_resample, bandpass_resample_host
    g = torch.Generator().manual_seed(11)
    blocks = [(torch.randn(n, 8, 795, generator=g) * 20).pin_memory() for n in (7, 1, 12, 5, 3)]
    outs = list(bandpass_resample_host(blocks, depth=2))
    assert len(outs) == len(blocks)
    for b, o in zip(blocks, outs):
        assert o.dtype == torch.float32 and o.device.type == "cpu" and o.shape == (b.shape[0], 8, 636)
        assert torch.equal(o, bandpass_resample(b.cuda()).cpu())
    assert torch.equal(next(iter(bandpass_resample_host([blocks[2].numpy()], depth=3))), outs[2])
    assert list(bandpass_resample_host([])) == []


def test_filter_properties_at_scale(filter_golden):
    """BASELINE config-3 sized batch: linearity and time-reversal symmetry of the zero-phase filter."""
    from fast.data import sosfiltfilt
    sos = filter_golden["sos"]
    gen = torch.Generator(device='cuda').manual_seed(4)
    # float64 inputs: with float32 inputs the odd extension is rounded in float32 (as scipy does) and linearity
    # only holds to 1e-7
    a = torch.randn(400, 64, 795, device='cuda', generator=gen, dtype=torch.float64) * 20
    b = torch.randn(400, 64, 795, device='cuda', generator=gen, dtype=torch.float64) * 20
    ya, yb = sosfiltfilt(sos, a), sosfiltfilt(sos, b)
    yab = sosfiltfilt(sos, 2 * a - 3 * b)
    assert rel_err(yab, 2 * ya - 3 * yb) < 1e-9
    # rows are independent: any row equals the same row filtered alone
    assert torch.equal(sosfiltfilt(sos, a[123, 7:8].contiguous(), out_dtype=torch.float64), ya[123, 7:8])
    # a constant is rejected by the band-pass (steady-state initial conditions => no start-up transient)
    c = torch.full((2, 795), 1234.5, device='cuda')
    assert float(sosfiltfilt(sos, c, out_dtype=torch.float64).abs().max()) < 1e-6


# ---------------------------------------------------------------------------------------------- fused training step
def _oracle_train(sd, cfg, batches, lrs):
    sd = {k: v.clone() for k, v in sd.items()}
    m = {k: torch.zeros_like(v) for k, v in sd.items()}
    v = {k: torch.zeros_like(t) for k, t in sd.items()}
    losses, first_grads = [], None
    for step, ((x, y), lr) in enumerate(zip(batches, lrs), 1):
        loss, _, grads, _ = O.loss_and_grads(sd, cfg, x, y)
        if first_grads is None:
            first_grads = grads
        losses.append(float(loss))
        for k in sd:
            O.adamw_step(sd[k], grads[k], m[k], v[k], step, lr)
    return losses, sd, first_grads


@pytest.mark.parametrize("use_graph", [False, True])
def test_fused_trainer_matches_oracle_loop(use_graph):
    from fast.train import FusedTrainer
    m = _prod_model(dropout=0.0).train()
    cfg = make_prod_cfg(dropout=0.0)
    sd0 = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(11)
    batches = [(torch.randn(8, 64, 795, generator=g) * 3, torch.randint(0, 5, (8,), generator=g).to(torch.uint8)) for _ in range(4)]
    factors = [0.1, 0.0, 0.5, 1.0]
    ft = FusedTrainer(m, lr=5e-4, lr_factors=factors, use_graph=use_graph)
    losses = []
    for x, y in batches:
        loss, _ = ft.train_step(x.cuda(), y.cuda())
        losses.append(float(loss))
    ref_losses, ref_sd, g1 = _oracle_train(sd0, cfg, batches, [5e-4 * f for f in factors])
    np.testing.assert_allclose(losses, ref_losses, rtol=2e-5, atol=2e-6)
    for k, p in m.state_dict().items():
        # parameters after 4 AdamW updates: compare the UPDATE (p - p0), which is what the step computes.  Adam divides
        # by sqrt(v): where the true gradient is zero (e.g. the key bias of every attention layer) the update is the
        # sign of rounding noise in ANY implementation, so only elements with a resolvable gradient are compared.
        du, dr = p.cpu() - sd0[k], ref_sd[k] - sd0[k]
        ok = g1[k].abs() > 1e-3 * g1[k].abs().max()
        assert ok.any(), k
        assert float((du - dr)[ok].abs().max()) <= 2e-2 * float(dr[ok].abs().max()) + 1e-7, k
    with pytest.raises(RuntimeError):
        ft.train_step(batches[0][0].cuda(), batches[0][1].cuda())          # past the end of the LR schedule
    loss, n_correct, pred = ft.eval_step(batches[0][0].cuda(), batches[0][1].cuda())
    m.eval()
    with torch.no_grad():
        logits = m(batches[0][0].cuda())
    assert torch.equal(pred, logits.argmax(1))
    assert int(n_correct) == int((pred.cpu() == batches[0][1].long()).sum())


def test_fused_trainer_host_batches_equal_device_batches():
    """train_step / eval_step fed pinned HOST batches (uploaded on the trainer's side stream through two staging pairs, so that
    consecutive steps overlap copy and kernels) follow the same loss trajectory as the same batches given as CUDA tensors."""
    from fast.train import FusedTrainer
    g = torch.Generator().manual_seed(21)
    batches = [(torch.randn(8, 64, 795, generator=g) * 3, torch.randint(0, 5, (8,), generator=g).to(torch.uint8)) for _ in range(5)]
    out = {}
    for where in ("cuda", "host"):
        m = _prod_model(dropout=0.0).train()
        ft = FusedTrainer(m, lr=5e-4, use_graph=True)
        losses = []
        for x, y in batches:
            xb, yb = (x.cuda(), y.cuda()) if where == "cuda" else (x.pin_memory(), y.pin_memory())
            loss, _ = ft.train_step(xb, yb)
            losses.append(loss.clone())                             # read later: the calls stay asynchronous
        x, y = batches[0]
        l, nc, pred = ft.eval_step(*((x.cuda(), y.cuda()) if where == "cuda" else (x.pin_memory(), y.pin_memory())))
        out[where] = ([float(v) for v in losses], float(l), int(nc), pred.cpu())
    np.testing.assert_allclose(out["host"][0], out["cuda"][0], rtol=2e-5, atol=2e-6)
    assert abs(out["host"][1] - out["cuda"][1]) < 2e-5 and out["host"][2] == out["cuda"][2]
    assert torch.equal(out["host"][3], out["cuda"][3])


@pytest.mark.parametrize("B", [6, 700])
def test_dropout_train_mode_is_consistent(B):
    """p = 0.1: masks are regenerated identically in the backward (directional derivative check through the C ABI; batch 6 runs
    the six-warps-per-trial schedule of the token kernels, batch 700 the warp-per-trial one) and change between steps; eval
    mode ignores them."""
    from fast import _native as N
    m = _prod_model(dropout=0.1).train()
    flat, desc = m._runtime()
    lib = N.lib()
    n_win = 5
    gen = torch.Generator(device='cuda').manual_seed(2)
    feat = torch.randn(B, n_win, 8, 32, device='cuda', generator=gen)
    logits = torch.empty(B, 5, device='cuda'); saved = torch.empty(lib.fastb200_transformer_saved_floats(desc, B, n_win), device='cuda')

    def fwd(params, seed, off, out, sv=None):
        N.check(lib.fastb200_transformer_fwd(desc, N.ptr(params), N.ptr(feat), B, n_win, 0, 0.1, seed, off, None, N.ptr(out),
                                             N.ptr(sv), N.stream_ptr()), "tfwd")
    fwd(flat, 5, 9, logits, saved)
    l2 = torch.empty_like(logits); fwd(flat, 5, 9, l2)
    assert torch.equal(logits, l2)                                         # same (seed, offset) -> same masks
    fwd(flat, 5, 10, l2)
    assert not torch.equal(logits, l2)                                     # new offset -> new masks
    w = torch.randn(B, 5, device='cuda', generator=gen) * (6.0 / B)        # L = sum(w * logits)
    grads = torch.zeros_like(flat); dfeat = torch.empty_like(feat)
    N.check(lib.fastb200_transformer_bwd(desc, N.ptr(flat), N.ptr(feat), N.ptr(saved), N.ptr(w), B, n_win, 0, 0.1, 5, 9, None,
                                         N.ptr(grads), N.ptr(dfeat), N.stream_ptr()), "tbwd")
    d = torch.randn(flat.numel(), device='cuda', generator=gen, dtype=torch.float64)
    is_head = torch.zeros(flat.numel(), dtype=torch.bool, device='cuda')
    for name, _, off, cnt in m._fp.items:
        if name.startswith("head."):
            is_head[off:off + cnt] = True
    d[is_head] = 0
    eps = 1e-3
    lp, lm = torch.empty_like(logits), torch.empty_like(logits)
    fwd((flat.double() + eps * d).float(), 5, 9, lp)
    fwd((flat.double() - eps * d).float(), 5, 9, lm)
    num = float(((lp.double() - lm.double()) * w.double()).sum() / (2 * eps))
    ana = float((grads.double() * d).sum())
    assert abs(num - ana) <= 2e-2 * max(abs(ana), 1.0), (num, ana)
    x = torch.randn(8, 64, 800, device='cuda', generator=gen)
    a, b = m(x), m(x)
    assert not torch.equal(a, b)                                           # train mode: fresh masks per call
    m.eval()
    with torch.no_grad():
        assert torch.equal(m(x), m(x))


@pytest.mark.parametrize("B", [1, 7, 64, 301, 700, 2500])
def test_token_path_backward_vs_oracle_autograd_every_schedule(B):
    """fastb200_transformer_fwd / _bwd through the C ABI against fp64 autograd of the oracle's forward_transformer: batch sizes
    that run the 6-warps-per-trial schedule (one slot, two slots with an idle one, several rounds) and the warp-per-trial
    schedule (partly filled CTAs, several rounds).  All gradients of the token path, dfeat, and accumulation into `grads`."""
    from fast import _native as N
    m = _prod_model(dropout=0.0).eval()
    cfg = make_prod_cfg(dropout=0.0)
    flat, desc = m._runtime()
    lib = N.lib()
    n_win = 5
    gen = torch.Generator(device='cuda').manual_seed(100 + B)
    feat = torch.randn(B, n_win, 8, 32, device='cuda', generator=gen)
    w = torch.randn(B, 5, device='cuda', generator=gen) / B
    logits = torch.empty(B, 5, device='cuda')
    saved = torch.empty(lib.fastb200_transformer_saved_floats(desc, B, n_win), device='cuda')
    N.check(lib.fastb200_transformer_fwd(desc, N.ptr(flat), N.ptr(feat), B, n_win, 0, 0.0, 0, 0, None, N.ptr(logits), N.ptr(saved),
                                         N.stream_ptr()), "tfwd")
    base = torch.randn(flat.numel(), device='cuda', generator=gen) * 1e-3            # the call ACCUMULATES into grads
    grads = base.clone()
    dfeat = torch.full_like(feat, float('nan'))
    N.check(lib.fastb200_transformer_bwd(desc, N.ptr(flat), N.ptr(feat), N.ptr(saved), N.ptr(w), B, n_win, 0, 0.0, 0, 0, None,
                                         N.ptr(grads), N.ptr(dfeat), N.stream_ptr()), "tbwd")
    sd = {k: v.detach().double().clone().requires_grad_(not k.startswith("head.")) for k, v in m.state_dict().items()
          if v.is_floating_point()}
    f64 = feat.double().requires_grad_(True)
    ref = O.forward_transformer(sd, cfg, f64)
    assert rel_err(logits, ref.detach()) < TOL
    (ref * w.double()).sum().backward()
    assert rel_err(dfeat, f64.grad) < TOL
    got = grads - base
    for name, _, off, cnt in m._fp.items:
        if name.startswith("head."):
            assert torch.equal(got[off:off + cnt], torch.zeros_like(got[off:off + cnt])), name
            continue
        e = rel_err(got[off:off + cnt], sd[name].grad.reshape(-1))
        assert e < TOL, (name, e)
    # head frozen ('train_transformer'): no dfeat wanted
    g2 = torch.zeros_like(flat)
    N.check(lib.fastb200_transformer_bwd(desc, N.ptr(flat), N.ptr(feat), N.ptr(saved), N.ptr(w), B, n_win, 0, 0.0, 0, 0, None,
                                         N.ptr(g2), None, N.stream_ptr()), "tbwd (no dfeat)")
    assert rel_err(g2, got) < 1e-5


# ---------------------------------------------------------------------------------------------- tensor-core head
# fastb200_head_fwd_tc (split-bf16 tcgen05 MMAs) against the fp64 oracle.  Bounds: 3 terms = 24 mantissa bits (fp32
# equivalent; the residual is the 1.5e-7 erfc approximation of the fused GELU), 2 terms = 16 bits.  Both far inside TOL.
TC_TOL = {"fp32": 2e-6, "bf16x3": 5e-6, "bf16x2": 3e-5}


@pytest.fixture
def head_precision():
    from fast.models import fast as FM
    old = FM.get_head_precision()
    yield FM.set_head_precision
    FM.set_head_precision(old)


def _oracle_feat64(model, cfg, x):
    sd = {k: v.detach().cpu().double() for k, v in model.state_dict().items()}
    return O.forward_head(sd, cfg, x.double().cpu())


@pytest.mark.parametrize("mode", ["fp32", "bf16x3", "bf16x2"])
def test_tc_head_features_vs_fp64_oracle(golden, head_precision, mode):
    head_precision(mode)
    cfg = make_prod_cfg(dropout=0.0)
    m = _prod_model(dropout=0.0).eval()
    x = _prod_x()
    with torch.no_grad():
        feat = m.forward_head(x.cuda())
        assert rel_err(feat, golden["prod/feat"]) < TOL                      # the real reference's output
        assert rel_err(feat, _oracle_feat64(m, cfg, x)) < TC_TOL[mode]
        # large, offset inputs (uV-scale EEG with a DC offset): operand splitting has to hold its relative accuracy
        xb = x * 40.0 + 25.0
        assert rel_err(m.forward_head(xb.cuda()), _oracle_feat64(m, cfg, xb)) < TC_TOL[mode]
        assert rel_err(m(x.cuda() * 30.0), golden["prod/logits_x30"]) < TOL


@pytest.mark.parametrize("mode", ["bf16x3", "bf16x2"])
def test_tc_head_batches_windows_and_determinism(head_precision, mode):
    head_precision(mode)
    cfg = make_prod_cfg(dropout=0.0)
    m = _prod_model(dropout=0.0).eval()
    g = torch.Generator().manual_seed(7)
    with torch.no_grad():
        for B, T in ((1, 795), (3, 800), (37, 795), (300, 1050)):           # 1050 samples: 7 windows (forward_head only)
            x = torch.randn(B, 64, T, generator=g) * 3.0
            f = m.forward_head(x.cuda())
            assert f.shape == (B, (T - 250) // 125 + 1, 8, 32)
            assert rel_err(f, _oracle_feat64(m, cfg, x)) < TC_TOL[mode]
            assert torch.equal(f, m.forward_head(x.cuda()))                 # fixed-order reductions: bit reproducible
            assert torch.equal(m.batched_forward_head(x.cuda(), 125, 16), f)  # per-item result independent of the batch
        x = torch.randn(5, 64, 800, generator=g)
        f50 = m.forward_head(x.cuda(), 50)                                   # step_override (reference fast.py:243-246)
        assert f50.shape == (5, 12, 8, 32)
        assert rel_err(f50, O.forward_head({k: v.detach().cpu().double() for k, v in m.state_dict().items()}, cfg,
                                           x.double(), 50)) < TC_TOL[mode]
        assert m.forward_head(x[:0].cuda()).shape == (0, 5, 8, 32)


def test_tc_head_zone_sizes_and_fallback(head_precision):
    """Zones of 1, 8, 9 and 16 electrodes (both electrode-tile layouts of the kernel); shapes the tensor-core kernel does not
    cover (window_len 120, dim_cnn 16, a 20-electrode zone) silently take the FFMA kernel -- still a CUDA kernel."""
    from fast.models import FAST
    from fast import _native as N
    head_precision("bf16x3")
    names = [f"E{i}" for i in range(40)]
    zones = {"one": names[:1], "eight": names[1:9], "nine": names[9:18], "sixteen": names[18:34], "rest": names[34:40]}
    base = dict(electrodes=names, zone_dict=zones, dim_cnn=32, dim_token=32, seq_len=500, window_len=250, slide_step=125,
                head="Conv4Layers", n_classes=4, num_layers=1, num_heads=4, dropout=0.0)
    x = torch.randn(6, 40, 500, generator=torch.Generator().manual_seed(3)) * 2.0
    for over, tc in ((dict(), True), (dict(window_len=120, slide_step=60), False), (dict(dim_cnn=16), False),
                     (dict(zone_dict={"big": names[:20], "rest": names[20:]}), False)):
        cfg = make_small_cfg(**{**base, **over})
        torch.manual_seed(11)
        m = FAST(cfg).cuda().eval()
        _, desc = m._runtime()
        assert bool(N.lib().fastb200_head_tc_supported(desc)) == tc
        with torch.no_grad():
            f = m.forward_head(x.cuda())
        assert rel_err(f, _oracle_feat64(m, cfg, x)) < TC_TOL["bf16x3"]


def test_head_bwd_tc_zone_sizes_and_both_bias_gradient_paths(head_backward, head_precision):
    """Tensor-core backward against the CUDA-core backward for zones of 1, 8, 9, 15 and 16 electrodes: zones of <= 15 take the
    collapsed-bias gradient out of the MMAs (a constant-1 electrode slot in the A tile), a 16-electrode zone has no free slot
    and sums the landed dL/dy2 planes in the converter warps."""
    from fast.models import FAST
    names = [f"E{i}" for i in range(49)]
    zones = {"one": names[:1], "eight": names[1:9], "nine": names[9:18], "fifteen": names[18:33], "sixteen": names[33:49]}
    cfg = make_small_cfg(electrodes=names, zone_dict=zones, dim_cnn=32, dim_token=32, seq_len=500, window_len=250, slide_step=125,
                         head="Conv4Layers", n_classes=4, num_layers=1, num_heads=4, dropout=0.0)
    torch.manual_seed(12)
    m = FAST(cfg).cuda().train()
    g = torch.Generator().manual_seed(6)
    x = (torch.randn(9, 49, 500, generator=g) * 3.0).cuda()
    y = torch.randint(0, 4, (9,), generator=g).cuda()
    grads = {}
    for bwd in ("ffma", "tc"):
        head_backward(bwd)
        head_precision("fp32" if bwd == "ffma" else "bf16x2")
        m.zero_grad(set_to_none=True)
        torch.nn.CrossEntropyLoss()(m(x), y).backward()
        grads[bwd] = {k: p.grad.clone() for k, p in m.named_parameters()}
    for k in grads["ffma"]:
        e = rel_err(grads["tc"][k], grads["ffma"][k])
        assert e < TOL, (k, e)


def test_head_bwd_tc_weight_gradients_vs_ffma_at_scale(head_backward, head_precision):
    """cnn3 / cnn4 weight gradients from the tcgen05 kernel against the CUDA-core kernel on a batch large enough that every
    persistent CTA walks many items (TMEM accumulation across items, staging ring and tile double-buffer wrap-around)."""
    m = _prod_model(dropout=0.0).train()
    g = torch.Generator().manual_seed(5)
    x = (torch.randn(96, 64, 795, generator=g) * 4.0).cuda()
    y = torch.randint(0, 5, (96,), generator=g).cuda()
    grads = {}
    for bwd in ("ffma", "tc"):
        head_backward(bwd)
        head_precision("fp32" if bwd == "ffma" else "bf16x2")
        m.zero_grad(set_to_none=True)
        torch.nn.CrossEntropyLoss()(m(x), y).backward()
        grads[bwd] = {k: p.grad.clone() for k, p in m.named_parameters()}
    for k in grads["ffma"]:
        e = rel_err(grads["tc"][k], grads["ffma"][k])
        assert e < TOL, (k, e)
    worst = max(rel_err(grads["tc"][k], grads["ffma"][k]) for k in grads["tc"])
    assert worst < TOL / 2, worst            # measured on B200: 1.1e-5 (bf16x2 forward, 2-term input-gradient chain and weight gradients)
