"""Dev: run a few fused training steps (target for ncu).  usage: train_run.py [batch] [steps] [graph 0|1]"""
import os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.models import FAST
from fast.train import FusedTrainer
from fast.data import Electrodes, Zones
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
graph = bool(int(sys.argv[3])) if len(sys.argv) > 3 else False
cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.1)
torch.manual_seed(0)
model = FAST(cfg).cuda().train()
ft = FusedTrainer(model, lr=5e-4, use_graph=graph)
x = torch.randn(B, 64, 795, device="cuda")
y = torch.randint(0, 5, (B,), device="cuda").to(torch.uint8)
for _ in range(steps):
    loss, _ = ft.train_step(x, y)
torch.cuda.synchronize()
print("ok", float(loss))
