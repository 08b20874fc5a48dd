"""Dev: run the tensor-core head a few times (target for ncu).  usage: tc_run.py [mode] [batch] [iters]"""
import os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.models import FAST
from fast.models import fast as FM
from fast.data import Electrodes, Zones
mode = sys.argv[1] if len(sys.argv) > 1 else "bf16x2"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.0)
torch.manual_seed(0)
model = FAST(cfg).cuda().eval()
FM.set_head_precision(mode)
x = torch.randn(B, 64, 795, device="cuda")
with torch.no_grad():
    for _ in range(iters):
        y = model(x)
torch.cuda.synchronize()
print("ok", float(y.abs().sum()))
