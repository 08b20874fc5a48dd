"""Dev: time the head forward / training step with the library FASTB200_LIB points at (A/B of kernel variants).
usage: FASTB200_LIB=path python tools/ab_time.py [tag]"""
import os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.models import FAST
from fast.models import fast as FM
from fast.train import FusedTrainer
from fast.data import Electrodes, Zones
tag = sys.argv[1] if len(sys.argv) > 1 else os.environ.get("FASTB200_LIB", "default")
cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.1)
torch.manual_seed(0)
model = FAST(cfg).cuda().eval()
x = torch.randn(8192, 64, 795, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
out = {}
with torch.no_grad():
    FM.set_head_precision("fp32")
    ref = model.forward_head(x[:256])
    for mode in ("bf16x2", "bf16x3"):
        FM.set_head_precision(mode)
        f = model.forward_head(x[:256])
        err = float((f - ref).abs().max() / ref.abs().max())
        for _ in range(3):
            model.forward_head(x)
        e0.record()
        for _ in range(10):
            model.forward_head(x)
        e1.record()
        torch.cuda.synchronize()
        out[f"fwd_{mode}_ms"] = round(e0.elapsed_time(e1) / 10, 3)
        out[f"fwd_{mode}_err"] = f"{err:.1e}"
FM.set_head_precision("bf16x2")
for b in (1024, 64):
    torch.manual_seed(1)
    tm = FAST(cfg).cuda().train()
    ft = FusedTrainer(tm, lr=5e-4, use_graph=True)
    xb = x[:b]
    yb = torch.randint(0, 5, (b,), device="cuda").to(torch.uint8)
    for _ in range(5):
        ft.train_step(xb, yb)
    e0.record()
    for _ in range(20):
        ft.train_step(xb, yb)
    e1.record()
    torch.cuda.synchronize()
    out[f"train_b{b}_ms"] = round(e0.elapsed_time(e1) / 20, 4)
    out[f"train_b{b}_loss"] = round(float(ft.loss), 5)
print(tag, out, flush=True)
