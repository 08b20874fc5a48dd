"""Dev: token-path kernel time: warp-per-trial kernels (G = 0) against the CTA-per-group kernels with G trials per CTA
(debug build: FASTB200_TR_GENERIC / FASTB200_TR_GROUP are read per call; GROUPS=0,1,4 selects).
usage: FASTB200_LIB=tools/variants/dbg/libfastb200.so python tools/tr_group_probe.py"""
import os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.models import FAST
from fast.data import Electrodes, Zones
cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.1)
torch.manual_seed(0)
model = FAST(cfg).cuda()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timed(fn, n=20):
    for _ in range(3): fn()
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for B in (64, 256, 1024, 8192):
    feat = torch.randn(B, 5, len(Zones), 32, device="cuda")
    for G in (os.environ.get("GROUPS", "0").split(",")):
        G = int(G)
        if G > 0: os.environ["FASTB200_TR_GROUP"] = str(G)
        os.environ["FASTB200_TR_GENERIC"] = "1" if G > 0 else "0"
        model.eval()
        with torch.no_grad():
            f = timed(lambda: model.forward_transformer(feat))
        if B <= 1024:
            model.train()
            fr = feat.clone().requires_grad_(True)
            def step():
                out = model.forward_transformer(fr)
                out.backward(torch.ones_like(out))
            fb = timed(step)
        else:
            fb = float("nan")
        print(f"B {B:5d} G {G:2d} infer_fwd {f:8.1f} us   train_fwd+bwd(+autograd) {fb:8.1f} us", flush=True)
