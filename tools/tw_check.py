"""Dev: the three token-path schedules on the same input (debug build: FASTB200_TR_GENERIC / FASTB200_TW_SPLIT per call)."""
import os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.models import FAST
from fast.data import Electrodes, Zones
cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.1)
torch.manual_seed(0)
model = FAST(cfg).cuda().eval()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 300
feat = torch.randn(B, 5, len(Zones), 32, device="cuda")
outs = {}
for name, env in (("generic", {"FASTB200_TR_GENERIC": "1"}), ("warp1", {"FASTB200_TR_GENERIC": "0", "FASTB200_TW_SPLIT": "1"}),
                  ("warp6", {"FASTB200_TR_GENERIC": "0", "FASTB200_TW_SPLIT": "6"})):
    os.environ.update(env)
    for mode in ("default", "train_head"):
        with torch.no_grad():
            outs[name, mode] = model._tokens_path(feat, 0 if mode == "default" else 1).clone()
for mode in ("default", "train_head"):
    g = outs["generic", mode]
    for name in ("warp1", "warp6"):
        d = (outs[name, mode] - g).abs().max().item()
        print(mode, name, "vs generic max abs diff", d, "equal to warp1:", torch.equal(outs[name, mode], outs["warp1", mode]))
