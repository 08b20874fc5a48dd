"""Host-side contract of the drop-in ``fast`` package that holds without a GPU: construction, names, state_dict wire
format, error behaviour.  Mirrors the structural pins of the reference's tests/test_model.py and tests/test_data.py."""
import numpy as np
import pytest
import torch

from conftest import make_prod_cfg, make_small_cfg


def test_package_exports():
    import fast
    import fast.data as D
    import fast.models as M
    import fast.train as T
    assert fast.__all__ == ["FAST", "seed_all", "green", "yellow", "Tick", "Tock"]
    assert M.__all__ == ["FAST", "AttentionBlock", "CVBlock", "Conv4Layers", "EEGNet_Encoder", "HeadConv_Paper_Version", "Head"]
    for n in ("BasicDataset", "load_standardized_h5", "Electrodes", "Zones", "SUBJECTS", "CLASSES", "load_training_set",
              "load_validation_set", "load_test_set", "load_test_set_per_subject", "load_subject_train_val"):
        assert n in D.__all__ and hasattr(D, n)
    for n in ("EEG_Encoder_Module", "cosine_scheduler", "inference_on_loader", "HistoryCallback"):
        assert n in T.__all__ and hasattr(T, n)


def test_dataset_constants(golden):
    from fast.data import CLASSES, SUBJECTS, Electrodes, Zones
    from fast.data.preprocess import TARGET_TIMEPOINTS
    assert SUBJECTS == [f"{i:02d}" for i in range(1, 16)] and len(CLASSES) == 5 and TARGET_TIMEPOINTS == 800
    assert Electrodes == list(golden["prod/electrodes"]) and len(set(Electrodes)) == 64
    assert list(Zones.keys()) == list(golden["prod/zone_names"])
    assert [len(v) for v in Zones.values()] == list(golden["prod/zone_sizes"])
    flat = [c for v in Zones.values() for c in v]
    assert sorted(flat) == sorted(Electrodes)                      # exact partition of the montage
    assert [Electrodes.index(c) for c in flat] == list(golden["prod/zone_index"])


def test_basic_dataset():
    from fast.data import BasicDataset
    x = np.random.randn(6, 8, 50).astype(np.float32)
    y = np.arange(6, dtype=np.uint8)
    ds = BasicDataset(x, y)
    assert len(ds) == 6 and ds[2][0].dtype == torch.float32 and ds[2][1].dtype == torch.uint8 and int(ds[2][1]) == 2
    ds4 = BasicDataset(x.reshape(2, 3, 8, 50), y.reshape(2, 3))
    assert len(ds4) == 6 and torch.equal(ds4[4][0], ds[4][0])


def test_model_structure_small():
    from fast.models import FAST
    m = FAST(make_small_cfg())
    assert m.name == 'FAST' and m.n_tokens == 3
    assert m.pos_embedding.shape == (1, 4, 16) and m.cls_token.shape == (1, 1, 16)
    assert m.head.index_dict['Central'].tolist() == [4, 5] and m.head.index_dict['Central'].dtype == torch.int64
    assert m.head.encoders['Central'].cnn2.weight.shape[2] == 2
    assert len(m.transformer) == 1 and m.last_layer.out_features == 3


def test_model_structure_prod(golden):
    from fast.models import FAST
    m = FAST(make_prod_cfg())
    assert m.n_tokens == 5
    assert sum(p.numel() for p in m.parameters()) == 191781
    assert [k for k, _ in m.named_parameters()] == list(golden["prod/grad_keys"])
    assert "head.encoders.Pre-frontal.cnn1.weight" in m.state_dict()


def test_state_dict_roundtrip_and_flattening():
    from fast.models import FAST
    torch.manual_seed(0)
    a, b = FAST(make_small_cfg()), FAST(make_small_cfg())
    b.load_state_dict(a.state_dict())
    flat, desc = b._runtime()                                         # flatten (CPU is fine for the bookkeeping)
    assert flat.numel() == sum(p.numel() for p in b.parameters()) == desc.n_params
    for (k, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.equal(pa, pb), k
        assert pb.data_ptr() == flat.data_ptr() + 4 * b._fp.offset[k]
    with torch.no_grad():
        flat.add_(1.0)                                                # an in-place optimiser update of the flat buffer ...
    assert torch.equal(b.cls_token, a.cls_token + 1.0)                # ... is visible through every Parameter
    b.load_state_dict(a.state_dict())                                 # copy_ into the views keeps them flat
    assert b._fp.ensure().data_ptr() == flat.data_ptr()
    assert torch.equal(b.cls_token, a.cls_token)


def test_unknown_mode_and_cpu_input_fail_loudly():
    from fast.models import FAST
    m = FAST(make_small_cfg())
    x = torch.randn(2, 8, 500)
    with pytest.raises(NotImplementedError):
        m(x, forward_mode='bogus')
    for mode in ('default', 'train_head', 'train_transformer'):
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            m(x, forward_mode=mode)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.forward_head(x)


def test_alternative_heads_construct():
    from fast.models import CVBlock, EEGNet_Encoder, FAST, HeadConv_Paper_Version
    assert CVBlock(4, 16).flat_dim == 256                 # reference: "flatten 256" for the 250-sample probe
    EEGNet_Encoder(4, 16, kernel_length=32)
    HeadConv_Paper_Version(4, 32)
    for head in ("CVBlock", "EEGNet_Encoder", "HeadConv_Paper_Version"):
        m = FAST(make_small_cfg(head=head))
        assert not m.head._fused and len(m.head.encoders) == 3


def test_alternative_heads_fresh_state_dict_equals_reference():
    """Same RNG stream AND the BatchNorm side effect of the reference's zero probe in CVBlock.__init__ (fast.py:66-74:
    running_var = 0.9, num_batches_tracked = 1), bit for bit against the real reference's fresh state_dict."""
    import os
    from conftest import ROOT
    from fast.models import FAST
    alt = np.load(os.path.join(ROOT, "tests", "golden", "alt_golden.npz"))
    for head in ("CVBlock", "EEGNet_Encoder", "HeadConv_Paper_Version"):
        torch.manual_seed(11)
        sd = FAST(make_small_cfg(head=head)).state_dict()
        ref_keys = [k[len(head) + 5:] for k in alt.files if k.startswith(head + "/sd0/")]
        assert list(sd.keys()) == ref_keys, head
        for k, v in sd.items():
            assert torch.equal(v, torch.from_numpy(alt[f"{head}/sd0/{k}"])), (head, k)
    from fast.models import CVBlock
    b = CVBlock(4, 16)
    assert float(b.bn1.running_var[0]) == float(torch.tensor(0.9)) and int(b.bn3.num_batches_tracked) == 1
    assert float(b.bn2.running_mean.abs().max()) == 0.0


def test_lr_factor_order():
    from fast.train import cosine_scheduler
    from fast.train.fused import lightning_lr_factors
    s = cosine_scheduler(1, 0.1, 200, 5, warmup_epochs=10)
    f = lightning_lr_factors(s)
    assert len(f) == len(s) and f[0] == s[-1] and f[1] == 0.0 and f[2] == s[1]


def test_lr_factor_order_pinned_by_lambdalr_mock():
    """tests/golden/make_golden.py --alt ran the reference's LambdaLR (trainer.py:50-53) through Lightning's
    optimiser.step -> global_step += 1 -> scheduler.step loop; lightning_lr_factors must give the lr of every step."""
    import os
    from conftest import ROOT
    from fast.train import cosine_scheduler
    from fast.train.fused import lightning_lr_factors
    alt = np.load(os.path.join(ROOT, "tests", "golden", "alt_golden.npz"))
    lst = cosine_scheduler(1, 0.1, 12, 3, warmup_epochs=10)
    np.testing.assert_array_equal(lst, alt["lr/list"])
    np.testing.assert_allclose(5e-4 * lightning_lr_factors(lst), alt["lr/used"], rtol=0, atol=1e-18)
    # and live, with the drop-in module's own scheduler object (no Lightning needed for the arithmetic)
    from fast.train import EEG_Encoder_Module
    mod = EEG_Encoder_Module(make_small_cfg(), max_epochs=12, niter_per_ep=3)
    if isinstance(getattr(type(mod), "global_step", None), property):
        return                                  # a real LightningModule owns global_step; the mock loop above already pinned it
    (opt,), (sch,) = mod.configure_optimizers()
    used = []
    for _ in range(len(lst)):
        used.append(opt.param_groups[0]["lr"])
        opt.step()
        mod.global_step += 1
        sch["scheduler"].step()
    np.testing.assert_allclose(used, alt["lr/used"], rtol=0, atol=1e-18)


def test_filter_design_helpers(filter_golden):
    from fast.data import filters as F
    np.testing.assert_allclose(F.bandpass_sos(4, 4, 40, 250), filter_golden["sos"], rtol=1e-12)
    assert F.default_padlen(filter_golden["sos"]) == 27
    np.testing.assert_allclose(F._kaiser_lowpass(4, 5), filter_golden["h_4_5"], rtol=1e-10, atol=1e-14)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        F.sosfiltfilt(filter_golden["sos"], torch.zeros(2, 100))


def test_alternative_head_chain_folding_matches_the_modules_on_cpu():
    """Host logic of the fused eval-mode alternative heads (fast/models/alt_fused.py): the stage chain handed to
    fastb200_alt_head_fwd -- BatchNorm folded into (scale, shift), EEGNet's spatial mix moved in front of the temporal filter,
    the paper head's cnn1_t . cnn1_s collapsed, weights laid out output-channel-fastest -- evaluated with plain torch ops in
    fp64 equals the module in eval mode (no kernel involved)."""
    import torch.nn.functional as F
    from fast.models import CVBlock, EEGNet_Encoder, HeadConv_Paper_Version
    from fast.models import alt_fused as A

    def run_chain(enc, x):
        stages, final = A._chain(enc)
        y = x
        for st in stages:
            in_ch = y.shape[1]
            if st["kind"] == A.DENSE:
                w = st["w"].reshape(in_ch, st["k"], st["out_ch"]).permute(2, 0, 1)
                y = F.conv1d(y, w, padding=st["pad"])
            else:
                w = st["w"].reshape(st["k"], st["out_ch"]).t().reshape(st["out_ch"], 1, st["k"])
                y = F.conv1d(y, w, padding=st["pad"], groups=in_ch)
            if "scale" in st:
                y = y * st["scale"][None, :, None] + st["shift"][None, :, None]
            y = F.elu(y) if st["act"] == A.ACT_ELU else F.gelu(y) if st["act"] == A.ACT_GELU else y
            if st["pool_kind"] == A.POOL_AVG:
                y = F.avg_pool1d(y, st["pool"])
            elif st["pool_kind"] == A.POOL_MAX:
                y = F.max_pool1d(y, st["pool"])
        if final["kind"] == A.FINAL_MEAN:
            return y.mean(-1)
        y = y.mean(-1) if final["kind"] == A.FINAL_MEAN_LINEAR else y.flatten(1)
        return y @ final["w"].reshape(final["lin_out"], final["lin_in"]).t() + final["b"]

    g = torch.Generator().manual_seed(5)
    for cls, args in ((CVBlock, (6, 32)), (EEGNet_Encoder, (6, 32)), (HeadConv_Paper_Version, (6, 32)), (EEGNet_Encoder, (4, 16, 32))):
        torch.manual_seed(1)
        enc = cls(*args).double()
        for mod in enc.modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.running_mean.copy_(torch.randn(mod.num_features, generator=g).double())
                mod.running_var.copy_(torch.rand(mod.num_features, generator=g).double() + 0.5)
                mod.weight.data.copy_(torch.randn(mod.num_features, generator=g).double())
                mod.bias.data.copy_(torch.randn(mod.num_features, generator=g).double())
        enc.eval()
        x = torch.randn(5, args[0], 250, generator=g).double()
        with torch.no_grad():
            ref, got = enc(x), run_chain(enc, x)
        assert float((ref - got).abs().max() / ref.abs().max()) < 1e-12, cls.__name__
