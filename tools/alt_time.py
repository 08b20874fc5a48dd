"""Dev: eval-mode forward_head of the alternative heads, fused kernel vs stock torch modules (ms per batch)."""
import os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.models import FAST
from fast.data import Electrodes, Zones
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
x = torch.randn(B, 64, 800, device="cuda")


def timed(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for head in ("CVBlock", "EEGNet_Encoder", "HeadConv_Paper_Version"):
    cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                                slide_step=125, head=head, n_classes=5, num_layers=4, num_heads=8, dropout=0.1)
    m = FAST(cfg).cuda().eval()
    with torch.no_grad():
        fused = timed(lambda: m.forward_head(x))
        m.head.alt_fused_ok = lambda: False
        stock = timed(lambda: m.forward_head(x))
    print(f"{head:24s} batch {B}: fused {fused:8.3f} ms   stock torch {stock:8.3f} ms   x{stock / fused:.1f}")
