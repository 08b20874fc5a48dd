"""The CPU oracle (oracle/) against the golden vectors produced by the REAL reference (tests/golden/make_golden.py)
and against scipy.  CPU only."""
import hashlib

import numpy as np
import pytest
import torch

from conftest import fingerprint, make_prod_cfg, make_small_cfg, rel_err, small_state_dict
from oracle import fast_oracle as O
from oracle import filter_oracle as FO


def _sha(t):
    a = t.detach().contiguous().numpy() if isinstance(t, torch.Tensor) else np.ascontiguousarray(t)
    return hashlib.sha256(a.tobytes()).hexdigest()


def _prod_model():
    """Production-config parameters: our drop-in module consumes the RNG stream exactly like the reference, so
    manual_seed(42) reproduces the reference's initial weights (checked against the golden fingerprints)."""
    from fast.models import FAST
    torch.manual_seed(42)
    return FAST(make_prod_cfg())


def _prod_x():
    return torch.randn(4, 64, 800, generator=torch.Generator().manual_seed(43))


# ---------------------------------------------------------------------------------------------- small config
def test_small_logits_loss_grads_all_modes(golden):
    cfg, sd = make_small_cfg(), small_state_dict(golden)
    x, y = torch.from_numpy(golden["small/x"]), torch.from_numpy(golden["small/y"])
    for mode in ("default", "train_head", "train_transformer"):
        loss, logits, grads, _ = O.loss_and_grads(sd, cfg, x, y, mode)
        assert rel_err(logits, golden[f"small/{mode}/logits"]) < 1e-5
        assert abs(float(loss) - float(golden[f"small/{mode}/loss"])) < 1e-5
        for k in sd:
            gk, nk = f"small/{mode}/grad/{k}", f"small/{mode}/nograd/{k}"
            if gk in golden.files:
                assert grads[k] is not None, (mode, k)
                assert rel_err(grads[k], golden[gk]) < 1e-4, (mode, k)
            else:
                assert nk in golden.files
                assert grads[k] is None or float(grads[k].abs().max()) == 0.0, (mode, k)


def test_small_features_and_step_override(golden):
    cfg, sd = make_small_cfg(), small_state_dict(golden)
    x = torch.from_numpy(golden["small/x"])
    assert rel_err(O.forward_head(sd, cfg, x), golden["small/feat"]) < 1e-5
    f50 = O.forward_head(sd, cfg, x, step_override=50)
    assert f50.shape == golden["small/feat_step50"].shape
    assert rel_err(f50, golden["small/feat_step50"]) < 1e-5


def test_small_input_gradient(golden):
    cfg, sd = make_small_cfg(), small_state_dict(golden)
    x = torch.from_numpy(golden["small/x"]).clone().requires_grad_(True)
    O.forward(sd, cfg, x).logsumexp(1).sum().backward()
    assert rel_err(x.grad, golden["small/dx_lse"]) < 1e-4


# ---------------------------------------------------------------------------------------------- production config
def test_prod_init_matches_reference_rng_stream(golden):
    m = _prod_model()
    sd = m.state_dict()
    assert list(sd.keys()) == list(golden["prod/sd_keys"])
    assert [str(tuple(v.shape)) for v in sd.values()] == list(golden["prod/sd_shapes"])
    assert sum(v.numel() for v in sd.values()) == int(golden["prod/n_params"]) == 191781
    fp = np.stack([fingerprint(v) for v in sd.values()])
    np.testing.assert_allclose(fp, golden["prod/sd_fp"], rtol=1e-12, atol=1e-12)


def test_prod_index_kernels_bit_exact(golden):
    cfg = make_prod_cfg()
    x = _prod_x()
    np.testing.assert_allclose(fingerprint(x), golden["prod/x_fp"], rtol=0, atol=0)
    xw = O.window_unfold(x, 250, 125)
    assert _sha(xw) == str(golden["prod/unfold_sha"])
    assert np.array_equal(xw[7].numpy()[:, :8], golden["prod/unfold_row7"])
    zi = O.zone_index(cfg)
    assert np.array_equal(np.concatenate(list(zi.values())), golden["prod/zone_index"])
    for a, idx in zi.items():
        assert _sha(O.zone_gather(xw, idx)) == str(golden[f"prod/gather_sha/{a}"])


def test_prod_logits_features_grads(golden):
    cfg = make_prod_cfg()
    sd = {k: v.detach() for k, v in _prod_model().state_dict().items()}
    x, y = _prod_x(), torch.from_numpy(golden["prod/y"])
    assert rel_err(O.forward_head(sd, cfg, x), golden["prod/feat"]) < 1e-5
    assert rel_err(O.forward(sd, cfg, x), golden["prod/logits"]) < 1e-5
    assert rel_err(O.forward(sd, cfg, x[:, :, :795].contiguous()), golden["prod/logits_795"]) < 1e-5
    assert np.array_equal(golden["prod/logits"], golden["prod/logits_795"])          # samples 750.. are never read
    assert rel_err(O.forward(sd, cfg, x, "train_head"), golden["prod/train_head_logits"]) < 1e-5
    assert rel_err(O.forward(sd, cfg, x * 30.0), golden["prod/logits_x30"]) < 1e-5
    loss, _, grads, dx = O.loss_and_grads(sd, cfg, x, y, input_grad=True)
    assert abs(float(loss) - float(golden["prod/loss"])) < 1e-5
    keys = list(golden["prod/grad_keys"])
    for i, k in enumerate(keys):
        gfp = golden["prod/grad_fp"][i]
        mine = fingerprint(grads[k])
        scale = max(gfp[3], 1e-30)
        assert abs(mine[3] - gfp[3]) / scale < 1e-3, k
        assert abs(mine[2] - gfp[2]) / (scale * np.sqrt(grads[k].numel())) < 1e-3, k
        if f"prod/grad/{k}" in golden.files:
            assert rel_err(grads[k], golden[f"prod/grad/{k}"]) < 1e-4, k
    assert rel_err(dx[0, 3], golden["prod/dx_b0_c3"]) < 1e-4


# ---------------------------------------------------------------------------------------------- trainer arithmetic
def test_cosine_scheduler(golden):
    np.testing.assert_allclose(O.cosine_scheduler(1, 0.1, 200, 5, warmup_epochs=10), golden["train/cosine_200x5_w10"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(O.cosine_scheduler(1, 0.1, 5, 5, warmup_epochs=0), golden["train/cosine_5x5_w0"], rtol=0, atol=1e-15)
    from fast.train.trainer import cosine_scheduler
    np.testing.assert_allclose(cosine_scheduler(1, 0.1, 200, 5, warmup_epochs=10), golden["train/cosine_200x5_w10"], rtol=0, atol=1e-15)
    s = cosine_scheduler(1, 0.1, 200, 5, warmup_epochs=10)
    assert s[0] == 0.0 and abs(s[-1] - 0.1) < 1e-3
    with pytest.raises(AssertionError):
        cosine_scheduler(1, 0.1, 5, 5, warmup_epochs=10)             # reference quirk


def test_adamw_oracle(golden):
    p = torch.from_numpy(golden["train/adamw_p0"]).clone()
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for step in range(4):
        O.adamw_step(p, torch.from_numpy(golden["train/adamw_g"][step]), m, v, step + 1, float(golden["train/adamw_lr"][step]))
        assert rel_err(p, golden[f"train/adamw_p{step + 1}"]) < 1e-6


def test_data_layout_oracle(golden):
    rng = np.random.default_rng(5)
    xm = rng.standard_normal((795, 64, 6)) * 30.0
    out = O.trials_from_mat_layout(xm, 800)
    assert out.dtype == np.float32 and out.shape == (6, 64, 800)
    assert _sha(out) == str(golden["data/mat_sha"])
    assert np.array_equal(out[2, 5, 780:], golden["data/mat_sample"])
    assert np.array_equal(O.labels_from_onehot(golden["data/onehot"]), golden["data/labels"])


# ---------------------------------------------------------------------------------------------- filter oracle
def test_filter_oracle_vs_scipy_golden(filter_golden):
    g = filter_golden
    sos, x = g["sos"], g["x"]
    assert FO.filtfilt_padlen(sos) == 27
    np.testing.assert_allclose(FO.sos_steady_state(sos), g["zi"], rtol=1e-12, atol=1e-14)
    y = FO.sosfiltfilt(sos, x)
    assert y.dtype == np.float64
    assert rel_err(y, g["y_sosfiltfilt"]) < 1e-11
    assert rel_err(FO.sosfiltfilt(g["sos_lp"], x[:, :, :120]), g["y_lp"]) < 1e-11
    np.testing.assert_allclose(FO.kaiser_lowpass(4, 5), g["h_4_5"], rtol=1e-10, atol=1e-14)
    yf = g["y_sosfiltfilt"]
    assert rel_err(FO.resample_poly(yf[:1, :2], 4, 5), g["y_resample_4_5"][:1, :2]) < 1e-11
    assert rel_err(FO.resample_poly(yf[:1, :1], 125, 128), g["y_resample_125_128"][:1, :1]) < 1e-11


def test_filter_oracle_vs_live_scipy():
    signal = pytest.importorskip("scipy.signal")
    rng = np.random.default_rng(3)
    x = rng.standard_normal((3, 333)).astype(np.float32) * 10 + 100
    for sos in (signal.butter(4, [4, 40], 'bandpass', fs=250, output='sos'),
                signal.butter(3, 0.5, 'highpass', fs=250, output='sos'),
                signal.butter(1, 30, 'lowpass', fs=250, output='sos')):
        assert rel_err(FO.sosfiltfilt(sos, x), signal.sosfiltfilt(sos, x, axis=-1)) < 1e-9
    with pytest.raises(ValueError):
        FO.sosfiltfilt(signal.butter(4, [4, 40], 'bandpass', fs=250, output='sos'), x[:, :27])
    assert rel_err(FO.resample_poly(x[:1].astype(np.float64), 2, 3), signal.resample_poly(x[:1].astype(np.float64), 2, 3, axis=-1)) < 1e-11
