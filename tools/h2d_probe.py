"""Dev: pinned host -> device bandwidth, contiguous vs the row-strided copy FAST.predict uses (750 of 795 samples per row)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "imagined-speech-decoding_b200"), ROOT]
import torch
from fast import _native as N
lib = N.lib()
rows, T, W = 1024 * 64, 795, 750
for chunks in (1, 2, 4, 8):
    R = rows * chunks
    h = torch.empty(R, T, dtype=torch.float32).pin_memory()
    d = torch.empty(R, T, dtype=torch.float32, device="cuda")
    d2 = torch.empty(R, W, dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream()

    def timed(fn, n=5):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e-3
    t_c = timed(lambda: d.copy_(h, non_blocking=True))
    t_s = timed(lambda: N.check(lib.fastb200_copy_rows_async(d2.data_ptr(), W * 4, h.data_ptr(), T * 4, W * 4, R, st.cuda_stream), "copy"))
    print(f"{R} rows: contiguous {R * T * 4 / t_c / 1e9:6.1f} GB/s ({t_c * 1e3:.2f} ms)   strided 750/795 {R * W * 4 / t_s / 1e9:6.1f} GB/s ({t_s * 1e3:.2f} ms)")
