"""Generate golden vectors from the REAL reference (run in the build container only).

    python tests/golden/make_golden.py

Imports ``fast.models.fast`` from /root/reference/src (needs torch + einops only; ``fast.data``
needs h5py, which is absent, so a stub module is injected to read the montage constants) and
scipy 1.18.1, and writes

    tests/golden/fast_golden.npz     model path: small config (full tensors) + production config
    tests/golden/filter_golden.npz   scipy sosfiltfilt / resample_poly outputs
    tests/golden/alt_golden.npz      (``--alt`` only) the three alternative heads inside FAST, the test-set edge pad, and a
                                     mock of Lightning's optimiser-then-scheduler loop over the reference's LambdaLR

The GPU box has no /root/reference: tests read only the .npz files.  Inputs are regenerated in the
tests from the seeds recorded here (torch CPU generator / numpy default_rng are deterministic for a
given version, and both boxes run the same image); the golden file also stores input checksums so a
generator drift would be detected rather than silently compared.
"""
import hashlib
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/src"
HERE = os.path.dirname(os.path.abspath(__file__))


def _import_reference():
    sys.modules.setdefault("h5py", types.ModuleType("h5py"))     # constants only; no HDF5 I/O here
    sys.path.insert(0, REF)
    import fast.models.fast as ref_fast                            # noqa
    import fast.data.preprocess as ref_pre                        # noqa
    sys.path.pop(0)
    return ref_fast, ref_pre


def sha(t):
    a = t.detach().cpu().contiguous().numpy() if isinstance(t, torch.Tensor) else np.ascontiguousarray(t)
    return hashlib.sha256(a.tobytes()).hexdigest()


def probe(n, seed=12345):
    """Fixed pseudo-random projection vector used to fingerprint big tensors."""
    return np.random.default_rng(seed).standard_normal(n)


def fingerprint(t):
    a = t.detach().double().numpy().ravel()
    return np.array([a.sum(), np.abs(a).sum(), float(a @ probe(a.size)), np.abs(a).max()])


def alt_main():
    """Goldens added in round 2: alternative heads (reference fast.py:32-196) run inside the reference FAST, the test-set
    edge pad (preprocess.py:110-114) and the LambdaLR step order (trainer.py:48-54).  Separate file so that the round-1
    vectors stay byte-identical."""
    from transformers import PretrainedConfig
    ref_fast, _ = _import_reference()
    out = {}
    electrodes = ['Fp1', 'Fp2', 'F3', 'F4', 'C3', 'C4', 'O1', 'O2']
    zones = {'Frontal': ['Fp1', 'Fp2', 'F3', 'F4'], 'Central': ['C3', 'C4'], 'Occipital': ['O1', 'O2']}
    g = torch.Generator().manual_seed(21)
    x = torch.randn(4, 8, 500, generator=g) * 2.0
    y = torch.tensor([0, 2, 1, 1], dtype=torch.uint8)
    out["x"], out["y"] = x.numpy(), y.numpy()
    for head in ("CVBlock", "EEGNet_Encoder", "HeadConv_Paper_Version"):
        cfg = PretrainedConfig(electrodes=electrodes, zone_dict=zones, dim_cnn=16, dim_token=16, seq_len=500,
                               window_len=250, slide_step=125, head=head, n_classes=3, num_layers=1, num_heads=4,
                               dropout=0.0)
        torch.manual_seed(11)
        m = ref_fast.FAST(cfg)
        for k, v in m.state_dict().items():
            out[f"{head}/sd0/{k}"] = v.numpy().copy()                 # fresh state_dict incl. BatchNorm buffers
        out[f"{head}/eval0_logits"] = m.eval()(x).detach().numpy()    # eval mode on the fresh running statistics
        m.train()
        for mod in m.modules():                                       # the heads carry their own Dropout(p>0): switch it off
            if isinstance(mod, torch.nn.Dropout):
                mod.p = 0.0
        logits = m(x)
        loss = torch.nn.CrossEntropyLoss()(logits, y)
        loss.backward()
        out[f"{head}/train_logits"] = logits.detach().numpy()
        out[f"{head}/loss"] = loss.detach().numpy()
        for k, p in m.named_parameters():
            out[f"{head}/grad/{k}"] = p.grad.numpy().copy()
        for k, v in m.state_dict().items():
            if "running_" in k or "num_batches" in k:
                out[f"{head}/sd1/{k}"] = v.numpy().copy()             # BatchNorm running statistics after one train step
        out[f"{head}/eval1_logits"] = m.eval()(x).detach().numpy()
        out[f"{head}/feat_eval"] = m.forward_head(x).detach().numpy()

    # test-set edge pad, (N, C, T) layout as the v7.3 files hold it (preprocess.py:110-114)
    rng = np.random.default_rng(9)
    xt = rng.standard_normal((5, 64, 795)) * 25.0
    pad_width = 800 - xt.shape[2]
    xp = np.pad(xt, ((0, 0), (0, 0), (0, pad_width)), 'edge').astype(np.float32)
    out["pad/sha"] = np.array(sha(xp))
    out["pad/tail"] = xp[3, 7, 790:].copy()

    # LambdaLR order: the reference's scheduler indexes cosine_lr_list[global_step - 1] and Lightning steps the optimiser,
    # increments global_step, then steps the scheduler (interval='step').  Mock of exactly that loop over torch's LambdaLR.
    sys.modules.setdefault("torchmetrics", types.ModuleType("torchmetrics"))
    sched_src = open("/root/reference/src/fast/train/trainer.py").read().split("class EEG_Encoder_Module")[0]
    ns = {}
    exec(compile(sched_src.replace("import torchmetrics", "").replace("import lightning as pl", "")
                 .replace("from ..models import FAST", ""), "trainer_head", "exec"), ns)
    lst = ns["cosine_scheduler"](1, 0.1, 12, 3, warmup_epochs=10)

    class Mod:
        global_step = 0
    mod = Mod()
    p = torch.nn.Parameter(torch.zeros(3))
    opt = torch.optim.AdamW([p], lr=0.0005)
    sch = torch.optim.lr_scheduler.LambdaLR(opt, lambda epoch: lst[mod.global_step - 1])
    used = []
    for _ in range(len(lst)):
        used.append(opt.param_groups[0]["lr"])
        p.grad = torch.ones(3)
        opt.step()
        mod.global_step += 1
        sch.step()
    out["lr/list"] = np.asarray(lst)
    out["lr/used"] = np.asarray(used)
    np.savez_compressed(os.path.join(HERE, "alt_golden.npz"), **out)
    print("wrote alt_golden.npz", os.path.getsize(os.path.join(HERE, "alt_golden.npz")))


def main():
    from transformers import PretrainedConfig
    ref_fast, ref_pre = _import_reference()
    out = {}

    # ---------------------------------------------------------------- small config, full tensors
    electrodes = ['Fp1', 'Fp2', 'F3', 'F4', 'C3', 'C4', 'O1', 'O2']
    zones = {'Frontal': ['Fp1', 'Fp2', 'F3', 'F4'], 'Central': ['C3', 'C4'], 'Occipital': ['O1', 'O2']}
    cfg = PretrainedConfig(electrodes=electrodes, zone_dict=zones, dim_cnn=16, dim_token=16, seq_len=500,
                           window_len=250, slide_step=125, head="Conv4Layers", n_classes=3, num_layers=1,
                           num_heads=4, dropout=0.0)
    torch.manual_seed(7)
    model = ref_fast.FAST(cfg)
    # move head features away from the tiny random-init regime so the head is visible in the logits
    g = torch.Generator().manual_seed(8)
    x = torch.randn(3, 8, 500, generator=g) * 4.0
    y = torch.tensor([0, 2, 1], dtype=torch.uint8)
    for k, v in model.state_dict().items():
        out["small/sd/" + k] = v.numpy().copy()
    out["small/x"] = x.numpy()
    out["small/y"] = y.numpy()
    model.train()
    for mode in ("default", "train_head", "train_transformer"):
        model.zero_grad()
        logits = model(x, forward_mode=mode)
        loss = torch.nn.CrossEntropyLoss()(logits, y)
        loss.backward()
        out[f"small/{mode}/logits"] = logits.detach().numpy()
        out[f"small/{mode}/loss"] = loss.detach().numpy()
        for k, p in model.named_parameters():
            if p.grad is not None:
                out[f"small/{mode}/grad/{k}"] = p.grad.numpy().copy()
            else:
                out[f"small/{mode}/nograd/{k}"] = np.zeros(0)
    out["small/feat"] = model.forward_head(x).detach().numpy()
    out["small/feat_step50"] = model.forward_head(x, step_override=50).detach().numpy()
    xr = x.clone().requires_grad_(True)
    model.zero_grad()
    model(xr).logsumexp(1).sum().backward()
    out["small/dx_lse"] = xr.grad.numpy().copy()

    # ---------------------------------------------------------------- production config
    E, Z = ref_pre.Electrodes, ref_pre.Zones
    out["prod/electrodes"] = np.array(E)
    out["prod/zone_names"] = np.array(list(Z.keys()))
    out["prod/zone_sizes"] = np.array([len(v) for v in Z.values()])
    out["prod/zone_index"] = np.concatenate([[E.index(c) for c in v] for v in Z.values()])
    pcfg = PretrainedConfig(electrodes=E, zone_dict=Z, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8,
                            dropout=0.1)
    torch.manual_seed(42)
    pm = ref_fast.FAST(pcfg)
    pm.eval()
    sd = pm.state_dict()
    out["prod/sd_keys"] = np.array(list(sd.keys()))
    out["prod/sd_shapes"] = np.array([str(tuple(v.shape)) for v in sd.values()])
    out["prod/sd_fp"] = np.stack([fingerprint(v) for v in sd.values()])
    out["prod/n_params"] = np.array(sum(v.numel() for v in sd.values()))
    g = torch.Generator().manual_seed(43)
    px = torch.randn(4, 64, 800, generator=g)
    py = torch.tensor([1, 4, 0, 2], dtype=torch.uint8)
    out["prod/x_fp"] = fingerprint(px)
    out["prod/y"] = py.numpy()
    # index kernels: reference ops, bit exact
    xw = px.unfold(-1, 250, 125)
    import einops
    xw = einops.rearrange(xw, 'B C N T -> (B N) C T')
    out["prod/unfold_sha"] = np.array(sha(xw))
    out["prod/unfold_row7"] = xw[7].numpy().copy()[:, :8]
    for a in Z:
        out[f"prod/gather_sha/{a}"] = np.array(sha(xw[:, pm.head.index_dict[a]]))
    # 795-sample input gives identical logits (samples 750.. never read)
    logits = pm(px)
    out["prod/logits"] = logits.detach().numpy()
    out["prod/logits_795"] = pm(px[:, :, :795].contiguous()).detach().numpy()
    out["prod/feat"] = pm.forward_head(px).detach().numpy()
    out["prod/train_head_logits"] = pm(px, forward_mode='train_head').detach().numpy()
    pm.zero_grad()
    pxr = px.clone().requires_grad_(True)
    loss = torch.nn.CrossEntropyLoss()(pm(pxr), py)
    loss.backward()
    out["prod/loss"] = loss.detach().numpy()
    out["prod/grad_fp"] = np.stack([fingerprint(p.grad) for _, p in pm.named_parameters()])
    out["prod/grad_keys"] = np.array([k for k, _ in pm.named_parameters()])
    for k, p in pm.named_parameters():
        if p.numel() <= 5200:
            out[f"prod/grad/{k}"] = p.grad.numpy().copy()
    out["prod/dx_fp"] = fingerprint(pxr.grad)
    out["prod/dx_b0_c3"] = pxr.grad[0, 3].numpy().copy()

    # head features scaled up (x * 30): makes the head dominate logits, the regime SURVEY 7.3-a warns about
    big = pm(px * 30.0)
    out["prod/logits_x30"] = big.detach().numpy()

    # ---------------------------------------------------------------- trainer arithmetic (bare torch)
    sys.modules.setdefault("torchmetrics", types.ModuleType("torchmetrics"))
    sched_src = open("/root/reference/src/fast/train/trainer.py").read().split("class EEG_Encoder_Module")[0]
    ns = {}
    exec(compile(sched_src.replace("import torchmetrics", "").replace("import lightning as pl", "")
                 .replace("from ..models import FAST", ""), "trainer_head", "exec"), ns)
    out["train/cosine_200x5_w10"] = ns["cosine_scheduler"](1, 0.1, 200, 5, warmup_epochs=10)
    out["train/cosine_5x5_w0"] = ns["cosine_scheduler"](1, 0.1, 5, 5, warmup_epochs=0)
    torch.manual_seed(3)
    p = torch.nn.Parameter(torch.randn(257))
    opt = torch.optim.AdamW([p], lr=5e-4)
    out["train/adamw_p0"] = p.detach().numpy().copy()
    gs = []
    for step in range(4):
        gq = torch.randn(257, generator=torch.Generator().manual_seed(100 + step))
        gs.append(gq.numpy().copy())
        p.grad = gq.clone()
        for grp in opt.param_groups:
            grp["lr"] = 5e-4 * [0.1, 0.0, 0.02, 1.0][step]
        opt.step()
        out[f"train/adamw_p{step + 1}"] = p.detach().numpy().copy()
    out["train/adamw_g"] = np.stack(gs)
    out["train/adamw_lr"] = 5e-4 * np.array([0.1, 0.0, 0.02, 1.0])

    # data side: transpose + cast + edge pad (preprocess.py:57-64 arithmetic on a synthetic .mat-layout array)
    rng = np.random.default_rng(5)
    xm = rng.standard_normal((795, 64, 6)) * 30.0                    # (T,C,N) float64, as loadmat returns
    xp_ = np.transpose(xm, (2, 1, 0)).astype(np.float32)
    xp_ = np.pad(xp_, ((0, 0), (0, 0), (0, 5)), 'edge')
    out["data/mat_sha"] = np.array(sha(xp_))
    out["data/mat_sample"] = xp_[2, 5, 780:].copy()
    yo = np.eye(5)[:, rng.integers(0, 5, 6)]
    out["data/onehot"] = yo
    out["data/labels"] = yo.argmax(0).astype(np.uint8)

    np.savez_compressed(os.path.join(HERE, "fast_golden.npz"), **out)

    # ---------------------------------------------------------------- filter goldens (scipy)
    from scipy import signal
    fo = {}
    sos = signal.butter(4, [4, 40], 'bandpass', fs=250, output='sos')
    fo["sos"] = sos
    rng = np.random.default_rng(1000)
    xf = (20 * rng.standard_normal((2, 4, 795)) + rng.uniform(-50, 50, (1, 4, 1))).astype(np.float32)
    fo["x"] = xf
    yf = signal.sosfiltfilt(sos, xf, axis=-1)
    fo["y_sosfiltfilt"] = yf
    fo["zi"] = signal.sosfilt_zi(sos)
    fo["y_resample_4_5"] = signal.resample_poly(yf, 4, 5, axis=-1)
    fo["y_resample_125_128"] = signal.resample_poly(yf, 125, 128, axis=-1)
    fo["y_resample_4_5_f32in"] = signal.resample_poly(xf, 4, 5, axis=-1)
    fo["h_4_5"] = signal.firwin(101, 1 / 5, window=('kaiser', 5.0)) * 4
    sos2 = signal.butter(2, 30, 'lowpass', fs=250, output='sos')
    fo["sos_lp"] = sos2
    fo["y_lp"] = signal.sosfiltfilt(sos2, xf[:, :, :120], axis=-1)
    np.savez_compressed(os.path.join(HERE, "filter_golden.npz"), **fo)
    print("wrote", {k: os.path.getsize(os.path.join(HERE, k)) for k in ("fast_golden.npz", "filter_golden.npz")})


if __name__ == "__main__":
    if "--alt" in sys.argv:
        alt_main()
    else:
        main()
