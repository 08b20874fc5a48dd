"""Eval-mode fused forward of the alternative zone heads (reference fast.py:32-100, 122-167, 170-196).

In eval mode BatchNorm is a per-channel affine on its running statistics and Dropout is the identity, so CVBlock,
EEGNet_Encoder and HeadConv_Paper_Version each reduce to a chain of "temporal conv -> (scale, shift) -> activation -> pool"
stages plus a final mean / Linear.  This module folds the parameters of every zone's encoder into that chain
(``fastb200_alt_head`` descriptor + one packed fp32 buffer, built with a handful of tiny device ops and cached until a
parameter or buffer changes) and runs ``fastb200_alt_head_fwd``: window unfold, zone gather and all stages of all zones in
ONE launch.  Training mode (batch statistics, running-stat updates, autograd) stays on the stock torch modules.
"""
import torch

from .. import _native as N

DENSE, DEPTHWISE = 0, 1
ACT_NONE, ACT_ELU, ACT_GELU = 0, 1, 2
POOL_NONE, POOL_AVG, POOL_MAX = 0, 1, 2
FINAL_MEAN, FINAL_MEAN_LINEAR, FINAL_FLATTEN_LINEAR = 0, 1, 2


def _bn_affine(bn):
    """BatchNorm2d in eval mode: y = x * scale + shift."""
    scale = bn.weight.detach() * torch.rsqrt(bn.running_var + bn.eps)
    return scale, bn.bias.detach() - bn.running_mean * scale


def _eegnet_front(conv_t, bn1, conv_s, bn2, pool):
    """Temporal conv (1 -> F1, k) + BN1 + depth-wise spatial conv (F1 -> F1 * D over the electrodes) + BN2 + ELU + avg-pool,
    re-ordered as: spatial mix z[o] = sum_c ws[o][c] x[c] (dense, k = 1), then the temporal filter of group o // D applied
    depth-wise; BN1's shift passes through the spatial sum as b1 * sum_c ws[o][c]."""
    wt = conv_t.weight.detach()[:, 0, 0, :]                       # (F1, k)
    ws = conv_s.weight.detach()[:, 0, :, 0]                       # (F2, Cz): channel o reads temporal filter o // D
    mult = ws.shape[0] // wt.shape[0]
    s1, b1 = _bn_affine(bn1)
    s2, b2 = _bn_affine(bn2)
    grp = torch.arange(ws.shape[0], device=ws.device) // mult
    scale = s2 * s1[grp]
    shift = s2 * b1[grp] * ws.sum(dim=1) + b2
    k = wt.shape[1]
    return [dict(kind=DENSE, out_ch=ws.shape[0], k=1, pad=0, act=ACT_NONE, pool_kind=POOL_NONE, pool=1, w=ws.t().reshape(-1)),
            dict(kind=DEPTHWISE, out_ch=ws.shape[0], k=k, pad=conv_t.padding[1], act=ACT_ELU, pool_kind=POOL_AVG, pool=pool,
                 w=wt[grp].t().reshape(-1), scale=scale, shift=shift)]


def _chain(enc):
    """(stages, final) of one zone encoder.  Stage 0 is the one that reads the zone's electrodes."""
    name = type(enc).__name__
    if name == "EEGNet_Encoder":
        conv_t, bn1 = enc.temporal_conv[0], enc.temporal_conv[1]
        conv_s, bn2, pool1 = enc.spatial_conv[0], enc.spatial_conv[1], enc.spatial_conv[3]
        dw, pw, bn3, pool2 = enc.separable_conv[0], enc.separable_conv[1], enc.separable_conv[2], enc.separable_conv[4]
        st = _eegnet_front(conv_t, bn1, conv_s, bn2, int(pool1.kernel_size[1]))
        st.append(dict(kind=DEPTHWISE, out_ch=dw.out_channels, k=dw.kernel_size[1], pad=dw.padding[1], act=ACT_NONE,
                       pool_kind=POOL_NONE, pool=1, w=dw.weight.detach()[:, 0, 0, :].t().reshape(-1)))
        s3, b3 = _bn_affine(bn3)
        st.append(dict(kind=DENSE, out_ch=pw.out_channels, k=1, pad=0, act=ACT_ELU, pool_kind=POOL_AVG,
                       pool=int(pool2.kernel_size[1]), w=pw.weight.detach()[:, :, 0, 0].t().reshape(-1), scale=s3, shift=b3))
        lin = enc.projector[2]
        return st, dict(kind=FINAL_MEAN_LINEAR, w=lin.weight.detach().reshape(-1), b=lin.bias.detach(), lin_in=lin.in_features,
                        lin_out=lin.out_features)
    if name == "CVBlock":
        st = _eegnet_front(enc.conv1, enc.bn1, enc.conv2, enc.bn2, int(enc.pool1.kernel_size[1]))
        s3, b3 = _bn_affine(enc.bn3)
        c3 = enc.conv3
        st.append(dict(kind=DENSE, out_ch=c3.out_channels, k=c3.kernel_size[1], pad=c3.padding[1], act=ACT_ELU, pool_kind=POOL_AVG,
                       pool=int(enc.pool2.kernel_size[1]), w=c3.weight.detach()[:, :, 0, :].permute(1, 2, 0).reshape(-1), scale=s3, shift=b3))
        lin = enc.projector
        return st, dict(kind=FINAL_FLATTEN_LINEAR, w=lin.weight.detach().reshape(-1), b=lin.bias.detach(), lin_in=lin.in_features,
                        lin_out=lin.out_features)
    if name == "HeadConv_Paper_Version":
        wt, bt = enc.cnn1_t.weight.detach()[:, 0, 0, :], enc.cnn1_t.bias.detach()      # (F1, 3), (F1)
        ws = enc.cnn1_s.weight.detach()[:, :, :, 0]                                      # (F1 out, F1 in, Cz)
        weff = torch.einsum("gfc,fk->gck", ws, wt)                                       # cnn1_s . cnn1_t, no nonlinearity between
        beff = torch.einsum("gfc,f->g", ws, bt)
        st = []
        convs = [(weff, enc.norm1, beff), (enc.cnn2.weight.detach()[:, :, 0, :], enc.norm2, None),
                 (enc.cnn3.weight.detach()[:, :, 0, :], enc.norm3, None), (enc.cnn4.weight.detach()[:, :, 0, :], enc.norm4, None)]
        for w, norm, bias in convs:
            s, b = _bn_affine(norm)
            if bias is not None:
                b = b + s * bias
            st.append(dict(kind=DENSE, out_ch=w.shape[0], k=w.shape[2], pad=0, act=ACT_GELU, pool_kind=POOL_MAX, pool=2,
                           w=w.permute(1, 2, 0).reshape(-1), scale=s, shift=b))
        return st, dict(kind=FINAL_MEAN, lin_in=0, lin_out=st[-1]["out_ch"])
    raise TypeError(f"no fused eval-mode chain for head class {name}")


def supported(head) -> bool:
    return all(type(e).__name__ in ("EEGNet_Encoder", "CVBlock", "HeadConv_Paper_Version") for e in head.encoders.values())


class FusedAltHead:
    """Descriptor + packed parameters of a ``Head`` built from alternative encoders; rebuilt when a tensor changes."""

    def __init__(self, head):
        self.head = head
        self.key = None
        self.desc = None
        self.packed = None

    def _state_key(self, window_len):
        ts = list(self.head.parameters()) + list(self.head.buffers())
        return (window_len,) + tuple((t.data_ptr(), t._version) for t in ts)

    def _build(self, window_len):
        encs = list(self.head.encoders.values())
        zone_index = self.head._zone_index()
        d = N.AltHeadDesc()
        d.window_len, d.n_zones = int(window_len), len(encs)
        if len(encs) > N.MAX_ZONES or sum(len(z) for z in zone_index) > N.MAX_CHANS:
            raise RuntimeError("too many zones / electrodes for the fused alternative head")
        pieces, off, pos = [], 0, 0
        for z, (enc, idx) in enumerate(zip(encs, zone_index)):
            stages, final = _chain(enc)
            if len(stages) > N.ALT_MAX_STAGES:
                raise RuntimeError("alternative head chain too long")
            d.zone_size[z], d.zone_off[z] = len(idx), pos
            for j, c in enumerate(idx):
                d.zone_chan[pos + j] = int(c)
            pos += len(idx)
            d.zone_base[z] = off
            rel = 0
            blocks = []

            def put(t):
                nonlocal rel
                at = rel
                t = t.reshape(-1).float()
                blocks.append(t)
                pad = (-t.numel()) % 4
                if pad:
                    blocks.append(t.new_zeros(pad))
                rel += t.numel() + pad
                return at
            # zone-independent layout first (stages >= 1, affines, Linear), the electrode-reading stage 0 last
            offs = []
            for i, st in enumerate(stages):
                o = dict(w=-1, scale=-1, shift=-1)
                if i > 0:
                    o["w"] = put(st["w"])
                if "scale" in st:
                    o["scale"], o["shift"] = put(st["scale"]), put(st["shift"])
                offs.append(o)
            lin_w = put(final["w"]) if final["kind"] != FINAL_MEAN else 0
            lin_b = put(final["b"]) if final["kind"] != FINAL_MEAN else 0
            offs[0]["w"] = put(stages[0]["w"])
            if z == 0:
                d.n_stages, d.final_kind = len(stages), final["kind"]
                d.lin_in, d.lin_out, d.lin_w_off, d.lin_b_off = final["lin_in"], final["lin_out"], lin_w, lin_b
                for i, (st, o) in enumerate(zip(stages, offs)):
                    s = d.stage[i]
                    s.kind, s.out_ch, s.k, s.pad = st["kind"], int(st["out_ch"]), int(st["k"]), int(st["pad"])
                    s.act, s.pool_kind, s.pool = st["act"], st["pool_kind"], int(st["pool"])
                    s.w_off, s.scale_off, s.shift_off = o["w"], o["scale"], o["shift"]
            else:                                                  # every zone must share the offsets of zone 0
                for i, o in enumerate(offs):
                    s = d.stage[i]
                    assert (s.w_off, s.scale_off, s.shift_off) == (o["w"], o["scale"], o["shift"]), "zone layouts differ"
            pieces += blocks
            off += rel
        self.desc, self.packed = d, torch.cat(pieces).contiguous()

    def forward(self, x, window_len, step):
        """x (B, C, T) -> (B, n_win, Z, F), eval mode."""
        N.require_cuda(x, "input")
        key = self._state_key(window_len)
        if key != self.key:
            with torch.no_grad():
                self._build(window_len)
            self.key = key
        x = x.contiguous().float()
        B, C, T = x.shape
        n_win = (T - window_len) // step + 1
        feat = torch.empty(B, n_win, self.desc.n_zones, self.desc.lin_out, device=x.device, dtype=torch.float32)
        N.check(N.lib().fastb200_alt_head_fwd(self.desc, N.ptr(self.packed), N.ptr(x), B, C, T, n_win, int(step), N.ptr(feat),
                                              N.stream_ptr(x.device)), "alt_head_fwd")
        return feat
