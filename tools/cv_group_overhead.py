"""Dev: fixed cost of one concurrent group of CV jobs (train_folds, J = 5) and of producing one synthetic subject."""
import cProfile, os, pstats, sys, time, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.data import Electrodes, Zones
from fast.train import cv as CV
cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.1)
t0 = time.perf_counter(); (X, Y), _ = CV.synthetic_subject(0); print(f"synthetic_subject: {time.perf_counter() - t0:.3f} s")
t0 = time.perf_counter(); X, Y = torch.as_tensor(X).cuda(), torch.as_tensor(Y).cuda(); torch.cuda.synchronize(); print(f"upload: {time.perf_counter() - t0:.3f} s")
J = int(sys.argv[1]) if len(sys.argv) > 1 else 5
folds = [dict(x=X, y=Y, train_idx=tr, val_idx=va, seed=10 + k) for k, (tr, va) in enumerate(CV.kfold_indices(350, 5, True, 42))][:J]
CV.train_folds(cfg, folds, max_epochs=3)
torch.cuda.synchronize()
def run(ep):
    t0 = time.perf_counter(); CV.train_folds(cfg, folds, max_epochs=ep); torch.cuda.synchronize(); return time.perf_counter() - t0
t2, t42 = min(run(2) for _ in range(3)), min(run(42) for _ in range(3))
per = (t42 - t2) / 40
print(f"J={J}: per epoch (all folds) {per*1e3:.2f} ms; fixed cost per group {1e3*(t2 - 2*per):.1f} ms; 200 epochs = {per*200:.2f} s")
pr = cProfile.Profile(); pr.enable(); run(2); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
