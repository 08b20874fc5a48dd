"""Dev: where does a CV fold's wall-clock go?  usage: cv_probe.py [epochs]"""
import os, sys, time, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.data import Electrodes, Zones
from fast.train import cv as CV
epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 20
cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.1)
(X, Y), _ = CV.synthetic_subject(0)
X, Y = torch.as_tensor(X).cuda(), torch.as_tensor(Y).cuda()
tr, va = CV.kfold_indices(350, 5, True, 42)[0]
CV.train_fold(cfg, X, Y, tr, va, max_epochs=2, seed=1)          # warm-up
torch.cuda.synchronize()
t0 = time.perf_counter()
CV.train_fold(cfg, X, Y, tr, va, max_epochs=epochs, seed=2)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print(f"fold: {epochs} epochs in {dt:.3f} s -> {1e3 * dt / epochs:.2f} ms/epoch, {280 * epochs / dt:.0f} train trials/s")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    CV.train_fold(cfg, X, Y, tr, va, max_epochs=5, seed=3)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=18, max_name_column_width=60))
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=14, max_name_column_width=60))
