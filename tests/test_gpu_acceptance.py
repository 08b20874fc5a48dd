"""Acceptance suite of the drop-in on a CUDA device: what the reference's own unit tests pin about the model API
(/root/reference/tests/test_model.py:36-404, conftest.py:13-69 -- 45 tests, all on CPU tensors there), restated against this
package with the module and its inputs on the GPU (the kernels have no CPU path).  Values are covered by the parity tests;
this file is about shapes, dtypes, modes, autograd structure and the attribute surface a user of the reference relies on."""
import pytest
import torch

from conftest import make_prod_cfg, make_small_cfg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def prod():
    from fast.models import FAST
    torch.manual_seed(0)
    return FAST(make_prod_cfg()).cuda().eval()


@pytest.fixture()
def small():
    from fast.models import FAST
    torch.manual_seed(0)
    return FAST(make_small_cfg()).cuda()


def _batch(b=4, c=64, t=800, seed=1):
    return torch.randn(b, c, t, generator=torch.Generator().manual_seed(seed)).cuda()


# ---------------------------------------------------------------------------------------------- forward (ref :36-113)
def test_forward_modes_shapes_dtype_and_finiteness(prod, small):
    x = _batch()
    for mode in ("default", "train_head", "train_transformer"):
        out = prod(x, forward_mode=mode)
        assert out.shape == (4, 5) and out.dtype == torch.float32 and out.is_cuda
        assert bool(torch.isfinite(out).all())
    with pytest.raises(NotImplementedError):
        prod(x, forward_mode="invalid_mode")
    assert prod(_batch(b=1)).shape == (1, 5)                         # single-sample batch
    assert small(_batch(2, 8, 500)).shape == (2, 3)                  # 8 electrodes, 3 zones, 3 classes
    from fast.models import FAST
    assert FAST.name == "FAST" and prod.name == "FAST"


def test_token_bookkeeping(prod, small):
    assert prod.n_tokens == 5 and small.n_tokens == 3               # (800-250)//125+1, (500-250)//125+1
    assert prod.pos_embedding.shape == (1, 6, 32) and prod.cls_token.shape == (1, 1, 32)
    assert small.pos_embedding.shape == (1, 4, 16) and small.cls_token.shape == (1, 1, 16)
    assert prod.config.window_len == 250 and hasattr(prod, "head") and hasattr(prod, "input_layer")
    assert len(prod.transformer) == 4 and prod.last_layer.out_features == 5


# ---------------------------------------------------------------------------------------------- gradients (ref :141-175)
def test_gradient_structure_per_mode(small):
    x = _batch(2, 8, 500)
    small.train()
    small.zero_grad()
    small(x).sum().backward()
    for name, p in small.named_parameters():                          # default mode: every parameter receives a gradient
        assert p.grad is not None, name
        assert bool(torch.isfinite(p.grad).all()), name
    small.zero_grad(set_to_none=True)
    small(x, forward_mode="train_transformer").sum().backward()
    for name, p in small.named_parameters():                          # head runs under no_grad: its grads stay None
        if name.startswith("head."):
            assert p.grad is None, name
    assert small.last_layer.weight.grad is not None and small.transformer[0].attn.in_proj_weight.grad is not None
    small.zero_grad(set_to_none=True)
    small(x, forward_mode="train_head").sum().backward()
    assert any(p.grad is not None and float(p.grad.abs().sum()) > 0 for n, p in small.named_parameters() if n.startswith("head."))
    assert small.input_layer[0].weight.grad is not None and small.last_layer.weight.grad is not None


# ---------------------------------------------------------------------------------------------- forward_head (ref :177-215)
def test_forward_head_step_override_and_batched(small):
    x = _batch(2, 8, 500)
    small.eval()
    with torch.no_grad():
        f = small.forward_head(x)
        assert f.shape == (2, 3, 3, 16)                               # (B, n_windows, n_zones, dim_cnn)
        f50 = small.forward_head(x, step_override=50)
        assert f50.shape[1] == (500 - 250) // 50 + 1 and f50.shape[1] > f.shape[1]
        fb = small.batched_forward_head(x, small.config.slide_step, 1)
        torch.testing.assert_close(fb, f)


# ---------------------------------------------------------------------------------------------- AttentionBlock (ref :218-250)
def test_attention_block_standalone():
    from fast.models import AttentionBlock
    torch.manual_seed(0)
    blk = AttentionBlock(embed_dim=32, hidden_dim=64, num_heads=4, dropout=0.0).cuda()
    x = torch.randn(2, 10, 32, device="cuda")
    assert blk(x).shape == x.shape
    blk2 = AttentionBlock(embed_dim=16, hidden_dim=32, num_heads=2, dropout=0.0).cuda()
    x2 = torch.randn(1, 5, 16, device="cuda")
    out = blk2(x2)
    assert out.shape == x2.shape and not torch.allclose(out, x2)      # residual block, not the identity
    assert AttentionBlock(16, 32, 2).cuda()(torch.randn(2, 1, 16, device="cuda")).shape == (2, 1, 16)      # one token
    blk2.eval()
    x3 = torch.randn(1, 3, 16, device="cuda")
    torch.testing.assert_close(blk2(x3), blk2(x3))
    for n in ("layer_norm_1", "attn", "layer_norm_2", "linear"):
        assert hasattr(blk, n)


# ---------------------------------------------------------------------------------------------- head classes (ref :253-360)
def _oracle_conv4(enc, x):
    """The reference op sequence on the CPU in float64 (values of the stand-alone encoder, fused or not)."""
    import torch.nn.functional as F
    w = {k: v.detach().cpu().double() for k, v in enc.state_dict().items()}
    y = F.conv2d(x.cpu().double().unsqueeze(1), w["cnn1.weight"], w["cnn1.bias"])
    y = F.conv2d(y, w["cnn2.weight"])
    y = F.conv2d(y, w["cnn3.weight"], padding=(0, 2))
    y = F.conv2d(y, w["cnn4.weight"], padding=(0, 2))
    return F.gelu(y).mean(dim=(2, 3))


def test_conv4layers_standalone_any_dim_and_length():
    """dim 8/16/32 and windows that fit run as the fused kernels; dim 64 and long windows under autograd run as stock torch
    CUDA ops (the reference accepts every combination: tests/test_model.py:256-284)."""
    from fast.models import Conv4Layers
    torch.manual_seed(0)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        for channels, dim, T, B in ((6, 32, 250, 4), (1, 16, 250, 2), (4, 32, 100, 2), (4, 32, 500, 2), (4, 8, 250, 2),
                                    (4, 64, 250, 2), (4, 32, 2000, 2)):
            enc = Conv4Layers(channels=channels, dim=dim).cuda()
            x = torch.randn(B, channels, T, device="cuda")
            for grad in (False, True):
                with torch.set_grad_enabled(grad):
                    out = enc(x)
                assert out.shape == (B, dim) and out.dtype == torch.float32
                ref = _oracle_conv4(enc, x)
                err = float((out.detach().cpu().double() - ref).abs().max() / ref.abs().max())
                assert err < 1e-4, (channels, dim, T, grad, err)
            out.sum().backward()                                      # autograd works on both routes
            assert enc.cnn1.weight.grad is not None and enc.cnn4.weight.grad is not None
            assert enc.cnn2.weight.shape == (dim, dim, channels, 1)
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_alternative_head_classes_standalone():
    from fast.models import CVBlock, EEGNet_Encoder, HeadConv_Paper_Version
    assert HeadConv_Paper_Version(in_channels=6, feature_dim=32).cuda()(torch.randn(4, 6, 250, device="cuda")).shape == (4, 32)
    for dim in (16, 32, 64):
        assert HeadConv_Paper_Version(in_channels=4, feature_dim=dim).cuda()(torch.randn(2, 4, 250, device="cuda")).shape == (2, dim)
    cv = CVBlock(n_channels=6, dim_token=32).cuda().eval()
    x = torch.randn(4, 6, 250, device="cuda")
    o3, o4 = cv(x), cv(x.unsqueeze(1))                                # 3-D and 4-D input give the same result
    assert o3.shape == (4, 32) and o4.shape == (4, 32)
    torch.testing.assert_close(o3, o4)
    assert EEGNet_Encoder(in_channels=6, feature_dim=32).cuda()(torch.randn(4, 6, 250, device="cuda")).shape == (4, 32)
    for kl in (16, 32, 64):
        assert EEGNet_Encoder(4, 16, kernel_length=kl).cuda()(torch.randn(2, 4, 250, device="cuda")).shape == (2, 16)
    enc = EEGNet_Encoder(in_channels=4, feature_dim=16).cuda()
    for T in (125, 250, 500):                                         # adaptive pooling: any length
        assert enc(torch.randn(2, 4, T, device="cuda")).shape == (2, 16)


# ---------------------------------------------------------------------------------------------- Head (ref :362-404)
def test_head_module():
    from fast.data import Electrodes, Zones
    from fast.models import Head
    torch.manual_seed(0)
    h = Head("Conv4Layers", ["A", "B", "C", "D", "E", "F"], {"Z1": ["A", "B", "C"], "Z2": ["D", "E", "F"]}, feature_dim=32).cuda()
    assert h(torch.randn(4, 6, 250, device="cuda")).shape == (4, 2, 32)
    h2 = Head("Conv4Layers", ["A", "B", "C", "D"], {"Z1": ["B", "D"]}, feature_dim=16)
    assert h2.index_dict["Z1"].tolist() == [1, 3] and not h2.index_dict["Z1"].is_cuda      # a plain CPU LongTensor, as in the reference
    h3 = Head("Conv4Layers", ["A", "B", "C"], {"Z1": ["A"], "Z2": ["B", "C"]}, feature_dim=16).cuda()
    assert h3(torch.randn(2, 3, 250, device="cuda")).shape == (2, 2, 16)              # a single-channel zone
    h4 = Head("Conv4Layers", Electrodes, Zones, feature_dim=32).cuda()
    assert h4(torch.randn(2, len(Electrodes), 250, device="cuda")).shape == (2, len(Zones), 32)
    h5 = Head("Conv4Layers", ["A", "B", "C", "D", "E"], {"Z1": ["A", "B"], "Z2": ["C", "D", "E"]}, feature_dim=16)
    assert h5.encoders["Z1"].cnn2.weight.shape[2] == 2 and h5.encoders["Z2"].cnn2.weight.shape[2] == 3
    # dim 64 is outside the fused kernels: same module, stock torch CUDA ops, same shapes; and FAST builds on top of it
    h6 = Head("Conv4Layers", ["A", "B", "C"], {"Z1": ["A"], "Z2": ["B", "C"]}, feature_dim=64).cuda()
    assert h6(torch.randn(2, 3, 250, device="cuda")).shape == (2, 2, 64)
    from fast.models import FAST
    m = FAST(make_small_cfg(dim_cnn=64)).cuda()
    x = _batch(2, 8, 500)
    out = m(x)
    assert out.shape == (2, 3)
    out.sum().backward()
    assert all(p.grad is not None for p in m.parameters())
