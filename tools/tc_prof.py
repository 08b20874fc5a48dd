"""Dev: wait-cycle profile of the tensor-core head forward (needs a library built with -DFASTB200_PROFILE:
FASTB200_BUILD_DIR=tools/variants/prof FASTB200_NVCC_EXTRA=-DFASTB200_PROFILE python imagined-speech-decoding_b200/build.py;
run with FASTB200_LIB=tools/variants/prof/libfastb200.so)."""
import ctypes, os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast.models import FAST
from fast.models import fast as FM
from fast.data import Electrodes, Zones
from fast import _native as N
cfg = types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
                            slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.0)
torch.manual_seed(0)
model = FAST(cfg).cuda().eval()
x = torch.randn(8192, 64, 795, device="cuda")
handle = ctypes.CDLL(N.LIB_PATH)
handle.fastb200_debug_set_tc_profile.argtypes = [ctypes.c_void_p]
handle.fastb200_debug_set_tc_profile.restype = None
names = ["mma total", "wait x_full", "wait d_free", "wait ye_ready", "wait yc_ready", "items", "epiA wait", "epiB wait",
         "epi embed (incl wait)", "epi cnn3 (incl wait)", "epi cnn4 after wait", "e4: drain+gelu+reduce", "e4: red_free wait", "conv: wait stg_full", "conv: wait x_empty", "conv: convert"]
for mode, dbg in (("bf16x2", 0), ("bf16x2", 0), ("bf16x2", 2), ("bf16x2", 4), ("bf16x2", 1)):
    os.environ["FASTB200_TC_DBG"] = str(dbg)
    FM.set_head_precision(mode)
    with torch.no_grad():
        model.forward_head(x)
    prof = torch.zeros(148, 16, dtype=torch.int64, device="cuda")
    handle.fastb200_debug_set_tc_profile(prof.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.no_grad():
        e0.record(); model.forward_head(x); e1.record()
    torch.cuda.synchronize()
    handle.fastb200_debug_set_tc_profile(None)
    p = prof.double().cpu()
    n_it = p[:, 5]
    print(mode, "dbg", dbg, f"{e0.elapsed_time(e1):.3f} ms", {n: round(float((p[:, i] / n_it).mean()), 1) for i, n in enumerate(names) if i != 5},
          "items/CTA", float(n_it.mean()), flush=True)
