"""CPU oracle for the FAST hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  Nothing under
``imagined-speech-decoding_b200/`` imports it; the product path is CUDA-only and fails
loudly without its extension.

It is a functional, from-first-principles restatement (plain fp32 PyTorch on the CPU,
state_dict in / tensors out) of the reference algorithm:

* window unfold + zone gather      -- /root/reference/src/fast/models/fast.py:242-252, 199-210
* Conv4Layers zone embedding       -- fast.py:103-119
* input layer, CLS, pos, blocks    -- fast.py:260-268, 10-29
* forward modes                    -- fast.py:270-284
* CE loss / AdamW / LR schedule    -- /root/reference/src/fast/train/trainer.py:15-27, 48-67
* edge pad / transpose / labels    -- /root/reference/src/fast/data/preprocess.py:57-64

Parity pinning: ``tests/golden/fast_golden.npz`` was produced by importing the real
reference module in the build container (``tests/golden/make_golden.py``);
``tests/test_oracle.py`` checks this oracle against it.  The band-pass / resample
restatement lives in ``oracle/filter_oracle.py`` (parity unpinned BY THE REFERENCE --
the reference has no such code; it is pinned against scipy 1.18.1 instead).
"""
from __future__ import annotations

import math
from typing import Dict, List, Mapping, Sequence

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# ----------------------------------------------------------------------------- config
class OracleConfig:
    """Attribute bag with the 12 fields FAST.__init__ reads (fast.py:219-230)."""

    def __init__(self, electrodes, zone_dict, head="Conv4Layers", dim_cnn=32, dim_token=32,
                 seq_len=800, window_len=250, slide_step=125, n_classes=5, num_heads=8,
                 num_layers=4, dropout=0.1):
        self.electrodes = list(electrodes)
        self.zone_dict = dict(zone_dict)
        self.head = head
        self.dim_cnn = dim_cnn
        self.dim_token = dim_token
        self.seq_len = seq_len
        self.window_len = window_len
        self.slide_step = slide_step
        self.n_classes = n_classes
        self.num_heads = num_heads
        self.num_layers = num_layers
        self.dropout = dropout

    @property
    def n_tokens(self):  # fast.py:232
        return (self.seq_len - self.window_len) // self.slide_step + 1


def n_tokens(cfg) -> int:
    """fast.py:232 -- works for any attribute bag with the reference's config fields."""
    return (cfg.seq_len - cfg.window_len) // cfg.slide_step + 1


def zone_index(cfg) -> Dict[str, List[int]]:
    """fast.py:205-206 -- electrode name -> row index, per zone, in zone order."""
    return {a: [cfg.electrodes.index(ch) for ch in chs] for a, chs in cfg.zone_dict.items()}


# ------------------------------------------------------------------ index kernels (M3, M5)
def window_unfold(x: Tensor, window_len: int, step: int) -> Tensor:
    """(B,C,T) -> (B*N, C, W): window n of trial b is row b*N+n (fast.py:247-249).

    Written as an explicit slice loop (not Tensor.unfold) so it is an independent
    statement of the indexing; bit-exact by construction.
    """
    B, C, T = x.shape
    n = (T - window_len) // step + 1
    out = x.new_empty(B, n, C, window_len)
    for i in range(n):
        out[:, i] = x[:, :, i * step:i * step + window_len]
    return out.reshape(B * n, C, window_len)


def zone_gather(xw: Tensor, idx: Sequence[int]) -> Tensor:
    """(R,C,W) -> (R,Cz,W) rows picked in zone order (fast.py:210)."""
    return torch.stack([xw[:, i] for i in idx], dim=1)


# ------------------------------------------------------------------------- head (M6)
def conv4layers(xz: Tensor, w1: Tensor, b1: Tensor, w2: Tensor, w3: Tensor, w4: Tensor,
                return_intermediates: bool = False):
    """Per-zone embedding (fast.py:111-119).

    xz (R,Cz,W); w1 (F,1,1,5) b1 (F); w2 (F,F,Cz,1); w3,w4 (F,F,1,5).
    temporal conv (valid) -> spatial collapse over the zone's electrodes -> two
    same-padded temporal convs -> exact GELU -> mean over time.
    """
    h = xz.unsqueeze(1)                              # (R,1,Cz,W)
    h = F.conv2d(h, w1, b1)                          # (R,F,Cz,W-4)
    y2 = F.conv2d(h, w2)                             # (R,F,1,W-4)
    y3 = F.conv2d(y2, w3, padding=(0, 2))
    y4 = F.conv2d(y3, w4, padding=(0, 2))
    feat = F.gelu(y4).mean(dim=(2, 3))               # (R,F)
    if return_intermediates:
        return feat, y2[:, :, 0], y3[:, :, 0], y4[:, :, 0]
    return feat


def forward_head(sd: Mapping[str, Tensor], cfg, x: Tensor, step_override=None) -> Tensor:
    """(B,C,T) -> (B,N,Z,F)  (fast.py:242-252)."""
    step = cfg.slide_step if step_override is None else step_override
    B = x.shape[0]
    xw = window_unfold(x, cfg.window_len, step)
    feats = []
    for area, idx in zone_index(cfg).items():
        p = f"head.encoders.{area}."
        feats.append(conv4layers(zone_gather(xw, idx), sd[p + "cnn1.weight"], sd[p + "cnn1.bias"],
                                 sd[p + "cnn2.weight"], sd[p + "cnn3.weight"], sd[p + "cnn4.weight"]))
    f = torch.stack(feats, dim=1)                    # (B*N,Z,F)
    return f.reshape(B, -1, f.shape[1], f.shape[2])


# ------------------------------------------------------------------ transformer (M7, M8)
def layer_norm(x: Tensor, w: Tensor, b: Tensor, eps: float = 1e-5) -> Tensor:
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * w + b


def gelu_exact(x: Tensor) -> Tensor:
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


def mha_self(x: Tensor, in_w: Tensor, in_b: Tensor, out_w: Tensor, out_b: Tensor, n_heads: int) -> Tensor:
    """nn.MultiheadAttention(batch_first) self-attention, eval mode (fast.py:26)."""
    B, S, D = x.shape
    hd = D // n_heads
    qkv = x @ in_w.t() + in_b
    q, k, v = qkv.split(D, dim=-1)
    sh = lambda t: t.reshape(B, S, n_heads, hd).permute(0, 2, 1, 3)
    q, k, v = sh(q), sh(k), sh(v)
    att = torch.softmax((q @ k.transpose(-1, -2)) / math.sqrt(hd), dim=-1)
    o = (att @ v).permute(0, 2, 1, 3).reshape(B, S, D)
    return o @ out_w.t() + out_b


def attention_block(sd, prefix: str, x: Tensor, n_heads: int) -> Tensor:
    """Pre-LN block (fast.py:24-29); dropout inactive (eval / p = 0)."""
    h = layer_norm(x, sd[prefix + "layer_norm_1.weight"], sd[prefix + "layer_norm_1.bias"])
    x = x + mha_self(h, sd[prefix + "attn.in_proj_weight"], sd[prefix + "attn.in_proj_bias"],
                     sd[prefix + "attn.out_proj.weight"], sd[prefix + "attn.out_proj.bias"], n_heads)
    h = layer_norm(x, sd[prefix + "layer_norm_2.weight"], sd[prefix + "layer_norm_2.bias"])
    u = gelu_exact(h @ sd[prefix + "linear.0.weight"].t() + sd[prefix + "linear.0.bias"])
    return x + (u @ sd[prefix + "linear.3.weight"].t() + sd[prefix + "linear.3.bias"])


def input_tokens(sd, feat: Tensor) -> Tensor:
    B, N = feat.shape[:2]
    return gelu_exact(feat.reshape(B, N, -1) @ sd["input_layer.0.weight"].t() + sd["input_layer.0.bias"])


def forward_transformer(sd, cfg, feat: Tensor) -> Tensor:
    """(B,N,Z,F) -> (B,n_classes)  (fast.py:260-268)."""
    tok = input_tokens(sd, feat)
    B = tok.shape[0]
    x = torch.cat([sd["cls_token"].expand(B, -1, -1), tok], dim=1)
    x = x + sd["pos_embedding"][:, :n_tokens(cfg) + 1]
    for l in range(cfg.num_layers):
        x = attention_block(sd, f"transformer.{l}.", x, cfg.num_heads)
    return x[:, 0] @ sd["last_layer.weight"].t() + sd["last_layer.bias"]


def forward(sd, cfg, x: Tensor, forward_mode: str = "default") -> Tensor:
    """fast.py:270-284 (dropout inactive)."""
    if forward_mode == "default":
        return forward_transformer(sd, cfg, forward_head(sd, cfg, x))
    if forward_mode == "train_head":
        tok = input_tokens(sd, forward_head(sd, cfg, x))
        return (tok @ sd["last_layer.weight"].t() + sd["last_layer.bias"]).mean(dim=1)
    if forward_mode == "train_transformer":
        with torch.no_grad():
            f = forward_head(sd, cfg, x)
        return forward_transformer(sd, cfg, f)
    raise NotImplementedError(forward_mode)


# ------------------------------------------------------------- loss / optimiser (T1, T2)
def cross_entropy_mean(logits: Tensor, y: Tensor) -> Tensor:
    """nn.CrossEntropyLoss() default: log-softmax + NLL, mean (trainer.py:37,59).  uint8 labels legal."""
    lse = torch.logsumexp(logits, dim=1)
    picked = logits.gather(1, y.long().view(-1, 1)).squeeze(1)
    return (lse - picked).mean()


def loss_and_grads(sd: Mapping[str, Tensor], cfg, x: Tensor, y: Tensor, forward_mode="default",
                   input_grad: bool = False):
    """CE loss, logits and d(loss)/d(param) for every state_dict entry (autograd over this oracle)."""
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    xin = x.detach().clone().requires_grad_(input_grad)
    logits = forward(leaves, cfg, xin, forward_mode)
    loss = cross_entropy_mean(logits, y)
    loss.backward()
    grads = {k: (v.grad if v.grad is not None else None) for k, v in leaves.items()}
    return loss.detach(), logits.detach(), grads, (xin.grad if input_grad else None)


def cosine_scheduler(base_value, final_value, epochs, niter_per_ep, warmup_epochs=0, start_warmup_value=0):
    """trainer.py:15-27: linear warm-up then half-cosine, one value per optimiser step."""
    warm = warmup_epochs * niter_per_ep
    total = epochs * niter_per_ep
    out = np.empty(total, dtype=np.float64)
    if warmup_epochs > 0:
        out[:warm] = np.linspace(start_warmup_value, base_value, warm)
    n = total - warm
    i = np.arange(n)
    out[warm:] = final_value + 0.5 * (base_value - final_value) * (1 + np.cos(np.pi * i / n))
    assert len(out) == total
    return out


def adamw_step(p: Tensor, g: Tensor, m: Tensor, v: Tensor, step: int, lr: float,
               beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01):
    """torch.optim.AdamW single-tensor update (trainer.py:49 defaults), in place; step is 1-based."""
    p.mul_(1.0 - lr * weight_decay)
    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)


# -------------------------------------------------------------------- data side (P1, P2)
def trials_from_mat_layout(x_tcn: np.ndarray, target_len: int = 800) -> np.ndarray:
    """(T,C,N) -> (N,C,target_len) float32, right edge-padded (preprocess.py:60-62)."""
    x = np.ascontiguousarray(np.transpose(x_tcn, (2, 1, 0))).astype(np.float32)
    pad = target_len - x.shape[-1]
    if pad > 0:
        x = np.concatenate([x, np.repeat(x[..., -1:], pad, axis=-1)], axis=-1)
    return x


def labels_from_onehot(y_onehot: np.ndarray) -> np.ndarray:
    """(n_classes, N) one-hot -> (N,) uint8 (preprocess.py:59,64)."""
    return y_onehot.argmax(0).astype(np.uint8)
