"""Round-2 parity tests of the drop-in surface around the hot path (all through the CUDA kernels / C ABI):

* the three alternative heads inside FAST (reference fast.py:32-196) against goldens generated from the real reference
  (tests/golden/make_golden.py --alt): fresh state_dict incl. BatchNorm buffers, eval and train mode, running statistics
  after one step;
* ``inference_on_loader`` and ``EEG_Encoder_Module.training_step / validation_step / configure_optimizers``
  (reference src/fast/train/trainer.py:30-93) against the CPU oracle loop;
* the test-set edge pad (reference src/fast/data/preprocess.py:110-114), bit exact;
* the model under ``torch.autocast('cuda', torch.bfloat16)`` (the reference's ``precision='bf16-mixed'``,
  scripts/train_fast.py:135,277): fp32 logits and gradients within the same 1e-4.
"""
import hashlib
import os

import numpy as np
import pytest
import torch

from conftest import ROOT, make_prod_cfg, make_small_cfg, rel_err
from oracle import fast_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-4


@pytest.fixture(scope="module")
def alt():
    return np.load(os.path.join(ROOT, "tests", "golden", "alt_golden.npz"))


@pytest.fixture()
def no_tf32():
    """The alternative heads are stock cuDNN / cuBLAS composites: keep them in full fp32 for the 1e-4 comparison."""
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


# ---------------------------------------------------------------------------------------------- alternative heads
@pytest.mark.parametrize("head", ["CVBlock", "EEGNet_Encoder", "HeadConv_Paper_Version"])
def test_alternative_heads_inside_fast_vs_reference(alt, head, no_tf32):
    from fast.models import FAST
    torch.manual_seed(11)
    m = FAST(make_small_cfg(head=head))
    sd = m.state_dict()
    for k, v in sd.items():                                         # same RNG stream AND the probe's BatchNorm side effect
        assert torch.equal(v, torch.from_numpy(alt[f"{head}/sd0/{k}"])), k
    m = m.cuda()
    x, y = torch.from_numpy(alt["x"]).cuda(), torch.from_numpy(alt["y"]).cuda()
    with torch.no_grad():
        assert rel_err(m.eval()(x), alt[f"{head}/eval0_logits"]) < TOL
    m.train()
    for mod in m.modules():
        if isinstance(mod, torch.nn.Dropout):
            mod.p = 0.0
    logits = m(x)
    loss = torch.nn.CrossEntropyLoss()(logits, y)
    loss.backward()
    assert rel_err(logits, alt[f"{head}/train_logits"]) < TOL
    assert abs(float(loss.detach()) - float(alt[f"{head}/loss"])) < 1e-5
    gmax = max(float(np.abs(alt[f"{head}/grad/{k}"]).max()) for k, _ in m.named_parameters())
    for k, p in m.named_parameters():
        ref = alt[f"{head}/grad/{k}"]
        assert p.grad is not None, k
        if np.abs(ref).max() < 1e-4 * gmax:
            # scale / shift directly in front of another BatchNorm (and conv weights in front of one): the true gradient is 0
            # and both sides hold rounding noise (~1e-7): compare absolutely against the largest gradient of the model
            assert float((p.grad.cpu() - torch.from_numpy(ref)).abs().max()) < 1e-5 * gmax, k
        else:
            # cuDNN's BatchNorm / conv backward reassociates; 5e-4 of the tensor's max on these stock composites
            assert rel_err(p.grad, ref) < 5e-4, (k, rel_err(p.grad, ref))
    for k, v in m.state_dict().items():
        if "running_" in k:
            assert rel_err(v, alt[f"{head}/sd1/{k}"]) < TOL, k
        elif "num_batches" in k:
            assert int(v) == int(alt[f"{head}/sd1/{k}"]), k
    with torch.no_grad():
        m.eval()
        assert rel_err(m(x), alt[f"{head}/eval1_logits"]) < TOL
        assert rel_err(m.forward_head(x), alt[f"{head}/feat_eval"]) < TOL


@pytest.mark.parametrize("head", ["CVBlock", "EEGNet_Encoder", "HeadConv_Paper_Version"])
def test_alternative_heads_fused_eval_kernel_vs_stock_modules(head, no_tf32):
    """fastb200_alt_head_fwd (eval mode: one launch for unfold + gather + every stage of every zone) against the stock torch
    modules it replaces, production montage (8 zones of 4..15 electrodes), non-trivial BatchNorm statistics, a ragged batch,
    the configured step and a step override; training mode and autograd keep the stock modules."""
    from fast import _native as N
    from fast.models import FAST
    torch.manual_seed(3)
    m = FAST(make_prod_cfg(head=head, dropout=0.0)).cuda()
    g = torch.Generator().manual_seed(4)
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.running_mean.copy_(torch.randn(mod.num_features, generator=g) * 0.5)
            mod.running_var.copy_(torch.rand(mod.num_features, generator=g) + 0.5)
            mod.weight.data.copy_(torch.randn(mod.num_features, generator=g))
            mod.bias.data.copy_(torch.randn(mod.num_features, generator=g) * 0.3)
    m.eval()
    x = (torch.randn(37, 64, 800, generator=g) * 3).cuda()
    lib = N.lib()

    def stock(step):
        xw = x.unfold(-1, 250, step)
        B, C, n, T = xw.shape
        rows = xw.permute(0, 2, 1, 3).reshape(B * n, C, T)
        f = torch.stack([enc(rows[:, m.head.index_dict[a].cuda()]) for a, enc in m.head.encoders.items()], dim=1)
        return f.reshape(B, n, f.shape[1], f.shape[2])

    with torch.no_grad():
        for step in (125, 50):
            n0 = lib.fastb200_launch_count()
            got = m.forward_head(x, None if step == 125 else step)
            assert lib.fastb200_launch_count() == n0 + 1              # the fused kernel ran, once
            ref = stock(step)
            assert got.shape == ref.shape
            assert rel_err(got, ref) < TOL, (head, step, rel_err(got, ref))
        rows = x[:, :, :250].contiguous()
        assert rel_err(m.head(rows), stock(125)[:, 0]) < TOL          # Head.forward on (R, C, T) rows
        # a parameter / buffer change invalidates the packed copy
        first = next(iter(m.head.encoders.values()))
        bn = next(mod for mod in first.modules() if isinstance(mod, torch.nn.BatchNorm2d))
        bn.running_mean.add_(0.25)
        assert rel_err(m.forward_head(x), stock(125)) < TOL
        logits = m(x)
        assert logits.shape == (37, 5) and bool(torch.isfinite(logits).all())
    n0 = lib.fastb200_launch_count()
    m.forward_head(x[:2])                                            # autograd wanted: stock modules
    m.train()
    with torch.no_grad():
        m.forward_head(x[:2])                                        # training mode: batch statistics, stock modules
    assert lib.fastb200_launch_count() == n0


# ---------------------------------------------------------------------------------------------- trainer-facing module
def _batches(n, b, seed):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(b, 64, 800, generator=g) * 2, torch.randint(0, 5, (b,), generator=g).to(torch.uint8)) for _ in range(n)]


def test_inference_on_loader_matches_oracle():
    from fast.data import BasicDataset
    from fast.models import FAST
    from fast.train import inference_on_loader
    cfg = make_prod_cfg()
    torch.manual_seed(42)
    m = FAST(cfg)
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(5)
    X = (torch.randn(21, 64, 800, generator=g) * 30).numpy()        # x30: the head dominates the logits
    Y = torch.randint(0, 5, (21,), generator=g).to(torch.uint8).numpy()
    loader = torch.utils.data.DataLoader(BasicDataset(X, Y), batch_size=8, shuffle=False)       # 8, 8, 5: a ragged tail
    pred, real = inference_on_loader(m, loader)
    assert pred.dtype == np.int64 and pred.shape == (21,) and real.dtype == np.uint8
    assert np.array_equal(real, Y)
    with torch.no_grad():
        ref = O.forward(sd, cfg, torch.from_numpy(X))
    top2 = ref.topk(2, dim=1).values
    clear = (top2[:, 0] - top2[:, 1]) > 1e-4 * ref.abs().max()       # argmax is only defined up to the parity tolerance
    assert clear.sum() >= 15
    assert np.array_equal(pred[clear.numpy()], ref.argmax(1).numpy()[clear.numpy()])
    assert not m.training and next(m.parameters()).is_cuda           # the reference leaves the model in eval mode on the GPU


def test_confusion_metric_vs_sklearn():
    from sklearn.metrics import accuracy_score, f1_score
    from fast.train.trainer import _ConfusionMetric
    g = torch.Generator().manual_seed(3)
    acc, f1 = _ConfusionMetric(5, 'accuracy').cuda(), _ConfusionMetric(5, 'f1').cuda()
    all_p, all_t = [], []
    for n in (64, 64, 24):
        logits = torch.randn(n, 5, generator=g)
        t = torch.randint(0, 4, (n,), generator=g)                   # class 4 never a target: exercises the 'absent class' rule
        a, f = acc(logits.cuda(), t.cuda()), f1(logits.cuda(), t.cuda())
        p = logits.argmax(1).numpy()
        assert abs(float(a) - accuracy_score(t.numpy(), p)) < 1e-6
        # torchmetrics' macro F1 averages over classes present in target or prediction = sklearn over those labels
        labels = np.union1d(t.numpy(), p)
        assert abs(float(f) - f1_score(t.numpy(), p, average='macro', labels=labels)) < 1e-6
        all_p.append(p); all_t.append(t.numpy())
    P, T = np.concatenate(all_p), np.concatenate(all_t)
    assert abs(float(acc.compute()) - accuracy_score(T, P)) < 1e-6
    assert abs(float(f1.compute()) - f1_score(T, P, average='macro', labels=np.union1d(T, P))) < 1e-6
    acc.reset()
    assert int(acc.confusion.sum()) == 0


def test_eeg_encoder_module_steps_match_oracle_loop():
    """training_step -> backward -> AdamW.step -> global_step += 1 -> LambdaLR.step (Lightning's order for
    interval='step'), three steps, against the oracle's loss_and_grads + adamw_step at the lr the scheduler really applies."""
    from fast.train import EEG_Encoder_Module
    from fast.train.fused import lightning_lr_factors
    cfg = make_prod_cfg(dropout=0.0)
    torch.manual_seed(42)
    mod = EEG_Encoder_Module(cfg, max_epochs=12, niter_per_ep=2).cuda()
    mod.train()
    assert isinstance(mod.loss, torch.nn.CrossEntropyLoss) and len(mod.cosine_lr_list) == 24
    sd = {k: v.detach().cpu().clone() for k, v in mod.model.state_dict().items()}
    mom = {k: torch.zeros_like(v) for k, v in sd.items()}
    var = {k: torch.zeros_like(v) for k, v in sd.items()}
    (opt,), (sch,) = mod.configure_optimizers()
    assert isinstance(opt, torch.optim.AdamW) and opt.defaults["lr"] == 0.0005 and opt.defaults["weight_decay"] == 0.01
    assert sch["interval"] == "step"
    assert len(opt.param_groups[0]["params"]) == 94
    factors = lightning_lr_factors(mod.cosine_lr_list)
    batches = _batches(3, 8, 17)
    for step, (x, y) in enumerate(batches, 1):
        assert abs(opt.param_groups[0]["lr"] - 5e-4 * factors[step - 1]) < 1e-12        # what Lightning would run this step at
        opt.zero_grad()
        loss = mod.training_step((x.cuda(), y.cuda()), step - 1)
        assert loss.dim() == 0 and float(loss) > 0
        loss.backward()
        opt.step()
        mod.global_step += 1
        sch["scheduler"].step()
        ref_loss, ref_logits, grads, _ = O.loss_and_grads(sd, cfg, x, y)
        assert abs(float(loss) - float(ref_loss)) < 2e-5
        assert abs(float(mod.logged["train_loss"]) - float(ref_loss)) < 2e-5
        assert abs(float(mod.logged["train_acc"]) - float((ref_logits.argmax(1) == y.long()).float().mean())) < 1e-6
        for k in sd:
            O.adamw_step(sd[k], grads[k], mom[k], var[k], step, 5e-4 * factors[step - 1])
    # validation_step: loss + accumulating metrics, no dropout, on the updated weights
    mod.eval()
    x, y = batches[0]
    with torch.no_grad():
        vloss = mod.validation_step((x.cuda(), y.cuda()), 0)
        ref = O.forward(sd, cfg, x)
    ref_v = O.cross_entropy_mean(ref, y)
    assert abs(float(vloss) - float(ref_v)) < 5e-5
    assert abs(float(mod.logged["val_loss"]) - float(ref_v)) < 5e-5
    assert 0.0 <= float(mod.logged["val_acc"]) <= 1.0
    assert int(mod.val_f1.confusion.sum()) == 8 and int(mod.val_acc.confusion.sum()) == 8


# ---------------------------------------------------------------------------------------------- data: test-set pad (P2)
def test_pad_edge_ncT_bit_exact(alt):
    from fast.data.preprocess import _pad_edge_ncT
    rng = np.random.default_rng(9)
    xt = rng.standard_normal((5, 64, 795)) * 25.0
    out = _pad_edge_ncT(xt, 800)
    assert out.shape == (5, 64, 800) and out.dtype == torch.float32
    assert hashlib.sha256(out.cpu().contiguous().numpy().tobytes()).hexdigest() == str(alt["pad/sha"])
    assert np.array_equal(out[3, 7, 790:].cpu().numpy(), alt["pad/tail"])
    ref = np.pad(xt, ((0, 0), (0, 0), (0, 5)), 'edge').astype(np.float32)          # preprocess.py:110-114
    assert np.array_equal(out.cpu().numpy(), ref)
    # float32 input, already long enough: unchanged apart from the dtype
    x32 = xt.astype(np.float32)[:, :, :20]
    assert np.array_equal(_pad_edge_ncT(x32, 16).cpu().numpy(), x32)


# ---------------------------------------------------------------------------------------------- autocast (bf16-mixed)
def test_forward_backward_under_bf16_autocast(golden):
    from fast.models import FAST
    torch.manual_seed(42)
    cfg = make_prod_cfg(dropout=0.0)
    m = FAST(cfg).cuda().train()
    x = torch.randn(4, 64, 800, generator=torch.Generator().manual_seed(43))
    y = torch.from_numpy(golden["prod/y"])
    with torch.autocast('cuda', dtype=torch.bfloat16):
        logits = m(x.cuda())
        feat = m.forward_head(x.cuda())
        loss = torch.nn.CrossEntropyLoss()(logits.float(), y.cuda())
    assert logits.dtype == torch.float32 and feat.dtype == torch.float32       # kernels are fp32-in / fp32-out: no bf16 output
    loss.backward()
    assert rel_err(logits, golden["prod/logits"]) < TOL
    assert rel_err(feat, golden["prod/feat"]) < TOL
    keys = [str(k) for k in golden["prod/grad_keys"]]
    named = dict(m.named_parameters())
    for k in keys:
        gk = f"prod/grad/{k}"
        if gk in golden.files:
            assert named[k].grad.dtype == torch.float32
            assert rel_err(named[k].grad, golden[gk]) < TOL, k
    # a bf16 INPUT (what an upstream autocast op would hand over) is cast, not rejected
    with torch.autocast('cuda', dtype=torch.bfloat16), torch.no_grad():
        lb = m(x.cuda().bfloat16())
    assert lb.dtype == torch.float32 and rel_err(lb, golden["prod/logits"]) < 2e-2


# ---------------------------------------------------------------------------------------------- FusedTrainer host state
def test_fused_trainer_follows_mode_toggles_and_parameter_moves():
    """A captured step graph bakes in by-value arguments and buffer addresses: toggling train()/eval(), changing grad_scale,
    a stand-alone ``model.head(...)`` call or a parameter swap must never leave the trainer replaying stale state."""
    from fast.models import FAST
    from fast.train import FusedTrainer
    torch.manual_seed(42)
    m = FAST(make_prod_cfg(dropout=0.1)).cuda().eval()
    ft = FusedTrainer(m, lr=5e-4, use_graph=True)
    g = torch.Generator().manual_seed(2)
    x = (torch.randn(8, 64, 800, generator=g) * 3).cuda()
    y = torch.randint(0, 5, (8,), generator=g).to(torch.uint8).cuda()
    for _ in range(2):
        ft.train_step(x, y)                                          # eval mode: captured with p_drop = 0
    m.train()
    for _ in range(2):
        ft.train_step(x, y)                                          # train mode: must be a different graph (p_drop = 0.1)
    drops = sorted(w["p_drop"] for w in ft._ws.values() if isinstance(w, dict) and "dlogits" in w)
    assert drops == [0.0, pytest.approx(0.1)]
    ft.grad_scale = 0.5
    ft.train_step(x, y)
    assert sorted(w["grad_scale"] for w in ft._ws.values() if isinstance(w, dict) and "dlogits" in w) == [0.5, 1.0, 1.0]
    ft.grad_scale = 1.0
    # stand-alone Head call: adopts its slice of the owner's flat buffer instead of pulling the parameters out of it
    ptr0, ver0 = ft.flat.data_ptr(), m._fp.version
    with torch.no_grad():
        f1 = m.head(x[:, :, :250].contiguous())
        f2 = m.forward_head(x)[:, 0]
    assert torch.equal(f1, f2)
    assert m._runtime()[0].data_ptr() == ptr0 and m._fp.version == ver0
    # swapping the Parameter objects (load_state_dict(assign=True)) re-creates the flat buffer: the trainer follows it
    sd = {k: v.detach().clone() * 0.5 for k, v in m.state_dict().items()}
    m.load_state_dict(sd, assign=True)
    m.eval()
    loss0, _, _ = ft.eval_step(x, y)
    loss0 = float(loss0)
    with torch.no_grad():
        ref = float(torch.nn.CrossEntropyLoss()(m(x), y.long()))
    assert abs(loss0 - ref) < 1e-5
    assert ft.flat.data_ptr() == m._runtime()[0].data_ptr() != ptr0
    m.train()
    before = m.state_dict()["last_layer.weight"].clone()
    ft.train_step(x, y)
    assert not torch.equal(before, m.state_dict()["last_layer.weight"])          # the update landed in the LIVE parameters


# ---------------------------------------------------------------------------------------------- GPU-direct ingestion (f3)
def _fake_dataset(root, sid="01", n_train=12, n_val=4, seed=0):
    """Two MATLAB v5 files in the competition's layout: struct epo_* with x (T, C, N) float64 and one-hot y (5, N)."""
    import scipy.io
    rng = np.random.default_rng(seed)
    parts = {}
    for folder, key, n in (("Training set", "epo_train", n_train), ("Validation set", "epo_validation", n_val)):
        os.makedirs(os.path.join(root, folder), exist_ok=True)
        x = rng.standard_normal((795, 64, n)) * 20.0 + rng.uniform(-50, 50, (1, 64, 1))
        y = np.eye(5)[:, rng.integers(0, 5, n)]
        scipy.io.savemat(os.path.join(root, folder, f"Data_Sample{sid}.mat"), {key: {"x": x, "y": y}})
        parts[key] = (x, y)
    return parts


def test_gpu_direct_ingestion_matches_reference_loader_arithmetic(tmp_path):
    from fast.data import load_subject_train_val, subject_train_val_device, load_training_set
    from fast.train import cv as CV
    parts = _fake_dataset(str(tmp_path))
    # the reference's arithmetic (preprocess.py:170-190): transpose, float32, five edge samples, one-hot argmax, concatenate
    ref_x, ref_y = [], []
    for key in ("epo_train", "epo_validation"):
        x, y = parts[key]
        xr = np.pad(np.transpose(x, (2, 1, 0)).astype(np.float32), ((0, 0), (0, 0), (0, 5)), 'edge')
        ref_x.append(xr); ref_y.append(y.argmax(0).astype(np.uint8))
    ref_x, ref_y = np.concatenate(ref_x), np.concatenate(ref_y)
    X, Y = load_subject_train_val(str(tmp_path), "01")
    assert isinstance(X, np.ndarray) and X.dtype == np.float32 and Y.dtype == np.uint8
    assert X.shape == (16, 64, 800) and np.array_equal(X, ref_x) and np.array_equal(Y, ref_y)
    Xd, Yd = subject_train_val_device(str(tmp_path), "01")
    assert Xd.is_cuda and Yd.is_cuda and Yd.dtype == torch.uint8
    assert np.array_equal(Xd.cpu().numpy(), ref_x) and np.array_equal(Yd.cpu().numpy(), ref_y)
    Xt, Yt = load_training_set(str(tmp_path))
    assert np.array_equal(Xt, ref_x[:12]) and np.array_equal(Yt, ref_y[:12])
    # fused band-pass + resample on the way in (north_star (c) feeding (a)): against scipy on the reference-layout array
    from scipy import signal
    Xf, _ = subject_train_val_device(str(tmp_path), "01", bandpass=dict(low=4.0, high=40.0, order=4, up=4, down=5))
    sos = signal.butter(4, [4, 40], 'bandpass', fs=250, output='sos')
    want = signal.resample_poly(signal.sosfiltfilt(sos, ref_x[:3].astype(np.float64), axis=-1), 4, 5, axis=-1)
    assert Xf.shape == (16, 64, 640) and Xf.dtype == torch.float32
    assert rel_err(Xf[:3], want) < 1e-5
    # the CV runner consumes the device tensors in place (no host copy of the trials)
    cfg = make_prod_cfg(dropout=0.0)
    seen = []

    def load(sid):
        out = subject_train_val_device(str(tmp_path), sid)
        seen.append(out[0].data_ptr())
        return out

    res = CV.run_cv_jobs(cfg, [(0, "01", 0)], load, str(tmp_path / "cv"), max_epochs=2, batch_size=8, n_folds=2, seed=3)
    assert len(res) == 1 and len(seen) == 1 and np.isfinite(res[0]["best_val_acc"])
    assert os.path.exists(tmp_path / "cv" / "sub-01" / "fold-0-best.pth")


# ---------------------------------------------------------------------------------------------- end-to-end predict (row-strided upload)
def test_predict_from_pinned_host_ships_only_the_samples_read():
    from fast.models import FAST
    torch.manual_seed(42)
    m = FAST(make_prod_cfg()).cuda().eval()
    g = torch.Generator().manual_seed(6)
    x = torch.randn(70, 64, 795, generator=g) * 30
    want = m.predict(x.cuda())
    xp = x.pin_memory()
    got = m.predict(xp, chunk=32)                                    # 32 + 32 + 6: ragged tail, two staging buffers reused
    assert torch.equal(got, want)
    assert m._last_h2d_bytes == 70 * 64 * 750 * 4                    # 750 of the 795 samples per row (windows end at 500 + 250)
    assert torch.equal(m.predict(x, chunk=32), want)                 # pageable host memory: plain full-row copies, same result
    assert m._last_h2d_bytes == 70 * 64 * 795 * 4
    x800 = torch.nn.functional.pad(x, (0, 5), mode="replicate").pin_memory()
    assert torch.equal(m.predict(x800, chunk=64), want)              # the reference's 800-sample layout: same logits
    assert m._last_h2d_bytes == 70 * 64 * 750 * 4
    assert m.predict(xp[:0]).shape == (0,)
