"""Warm per-kernel timing of the fused training step: monkey-patches the C-ABI calls FusedTrainer makes with CUDA-event
brackets (no graph), so each kernel's steady-state duration is seen with L2-resident weights -- what an ncu launch list
(cold caches, serialised) cannot show.  Usage: python tools/step_probe.py [batch] [iters]"""
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "imagined-speech-decoding_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

from bench import prod_cfg  # noqa: E402
from fast import _native as N  # noqa: E402
from fast.models import FAST  # noqa: E402
from fast.train import FusedTrainer  # noqa: E402


class TimedLib:
    def __init__(self, lib):
        self._lib = lib
        self.events = collections.defaultdict(list)

    def __getattr__(self, name):
        fn = getattr(self._lib, name)
        if not name.startswith("fastb200_") or name in ("fastb200_last_error",):
            return fn

        def call(*a):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn(*a)
            e1.record()
            self.events[name].append((e0, e1))
            return r
        return call


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    torch.manual_seed(0)
    model = FAST(prod_cfg()).cuda().train()
    ft = FusedTrainer(model, lr=5e-4, use_graph=False)
    x = torch.randn(B, 64, 795, device="cuda")
    y = torch.randint(0, 5, (B,), device="cuda").to(torch.uint8)
    for _ in range(5):
        ft.train_step(x, y)
    timed = TimedLib(ft.lib)
    ft.lib = timed
    for _ in range(iters):
        ft.train_step(x, y)
    torch.cuda.synchronize()
    tot = 0.0
    for name, ev in timed.events.items():
        per_step = len(ev) // iters
        ms = sorted(a.elapsed_time(b) for a, b in ev)
        med = ms[len(ms) // 2] * per_step
        tot += med
        print(f"{name:36s} x{per_step}  median {med * 1e3:8.1f} us")
    print(f"{'sum of medians':36s}     {tot * 1e3:8.1f} us   (batch {B})")


if __name__ == "__main__":
    main()
