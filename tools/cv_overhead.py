"""Dev: host-side fixed cost of one CV job (model build, trainer set-up, graph capture, best-state download)."""
import cProfile, os, pstats, sys, time, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.data import Electrodes, Zones
from fast.train import cv as CV
cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.1)
(X, Y), _ = CV.synthetic_subject(0)
X, Y = torch.as_tensor(X).cuda(), torch.as_tensor(Y).cuda()
tr, va = CV.kfold_indices(350, 5, True, 42)[0]
CV.train_fold(cfg, X, Y, tr, va, max_epochs=3, seed=1)
torch.cuda.synchronize()
def run(ep):
    t0 = time.perf_counter(); CV.train_fold(cfg, X, Y, tr, va, max_epochs=ep, seed=2); torch.cuda.synchronize(); return time.perf_counter() - t0
t2, t22 = min(run(2) for _ in range(3)), min(run(22) for _ in range(3))
per = (t22 - t2) / 20
print(f"per epoch {per*1e3:.2f} ms; fixed cost per job {1e3*(t2 - 2*per):.1f} ms")
pr = cProfile.Profile(); pr.enable(); run(2); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
