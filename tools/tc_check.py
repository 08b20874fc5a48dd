"""Dev check (GPU): tensor-core head vs FFMA head vs the CPU oracle; timing at batch 8192.
Needs a developer build of the library: FASTB200_NVCC_EXTRA="-DFASTB200_PROFILE" python imagined-speech-decoding_b200/build.py --force"""
import os, sys, types, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.models import FAST
from fast.models import fast as FM
from fast.data import Electrodes, Zones
from oracle import fast_oracle as O

cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.0)
torch.manual_seed(0)
model = FAST(cfg).cuda().eval()
sd = {k: v.detach().cpu().double() for k, v in model.state_dict().items()}


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).abs().max() / b.abs().max())


for B in (1, 3, 64):
    x = torch.randn(B, 64, 795, generator=torch.Generator().manual_seed(B)) * 5.0 + 2.0
    with torch.no_grad():
        ref = O.forward_head(sd, cfg, x.double()) if hasattr(O, "forward_head") else None
        out = {}
        for mode in ("fp32", "bf16x3", "bf16x2"):
            FM.set_head_precision(mode)
            out[mode] = model.forward_head(x.cuda())
            torch.cuda.synchronize()
        msg = f"B={B}: " + "  ".join(f"{m} vs fp32 {rel(out[m], out['fp32']):.2e}" for m in ("bf16x3", "bf16x2"))
        if ref is not None:
            msg += "  | vs fp64 oracle: " + "  ".join(f"{m} {rel(out[m], ref):.2e}" for m in out)
        print(msg, flush=True)

B = 8192
x = torch.randn(B, 64, 795, device="cuda")
for mode in ("fp32", "bf16x3", "bf16x2"):
    FM.set_head_precision(mode)
    with torch.no_grad():
        for _ in range(3):
            model.forward_head(x)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            model.forward_head(x)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"{mode}: {ms:.3f} ms per batch of {B} -> {B / ms * 1e3:.0f} trials/s", flush=True)

# ---- wait-cycle profile of the tensor-core kernel (instrumented variant)
import ctypes
from fast import _native as N
lib = N.lib()
lib.fastb200_debug_set_tc_profile.argtypes = [ctypes.c_void_p]
lib.fastb200_debug_set_tc_profile.restype = None
for mode, dbg in (("bf16x3", 0), ("bf16x2", 0), ("bf16x2", 7)):
    os.environ["FASTB200_TC_DBG"] = str(dbg)
    FM.set_head_precision(mode)
    prof = torch.zeros(148, 16, dtype=torch.int64, device="cuda")
    lib.fastb200_debug_set_tc_profile(prof.data_ptr())
    with torch.no_grad():
        model.forward_head(x)
    torch.cuda.synchronize()
    lib.fastb200_debug_set_tc_profile(None)
    p = prof.double().cpu()
    n_it = p[:, 5]
    names = ["mma total", "wait x_full", "wait d_free", "wait y_ready(E epi)", "wait y_ready(C3 epi)", "items", "epiA wait d_full", "epiB wait d_full",
             "epi embed (incl wait)", "epi cnn3 (incl wait)", "epi cnn4 after wait"]
    print(mode, "dbg", dbg, "per-item cycles (mean over CTAs):", {n: round(float((p[:, i] / n_it).mean()), 1) for i, n in enumerate(names) if i != 5},
          "items/CTA", float(n_it.mean()), "max total Mcyc", float(p[:, 0].max()) / 1e6, flush=True)
