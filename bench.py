"""Benchmark of the FAST hot path on B200 (contract: see the task statement / DESIGN.md section 7).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch B] [--no-extras]

Workload (BASELINE.json configs[1]): FAST inference, production config, batch 8192 synthetic BCIC2020-Track3-shaped
trials (64 ch x 795 samples, fp32) per GPU -- 1.67 GB of input per step, larger than the 126 MB L2, so consecutive
steps cannot serve each other from cache.  One "step" = the whole hot path over one batch: window unfold + zone gather
+ 8 x Conv4Layers (one kernel), input layer + CLS/pos + 4 transformer blocks + classifier (one kernel), argmax.

  value   trials/s, input already resident in HBM, K steps timed with CUDA events, max over ranks
  e2e     trials/s through the public drop-in API (``model.predict(x_host)``: the argmax-of-logits op of
          ``inference_on_loader``) with the batch in pinned HOST memory: the chunked H2D copy (overlapped with the kernels
          on a side stream), the kernels and the D2H read of the predictions are all inside the timed region
  extra   training throughput (FusedTrainer, fwd+bwd+AdamW) and a small batch sweep, outside the headline

Multi-GPU (torchrun): trials are independent, so ranks shard them with NO data-path collective (weak scaling: every
rank runs its own batch of 8192); the only collectives are the timing barrier and a MAX-reduce of the elapsed time.

--impl reference: the reference algorithm on the host cores -- the CPU oracle port (oracle/fast_oracle.py; the
reference is pure Python and /root/reference does not exist on the GPU box), all threads, bounded sample per step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
import types

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "imagined-speech-decoding_b200")
for _p in (PKG, ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np
import torch

# algorithmic work per trial: direct form as the reference computes it, and what the kernels
# execute after the exact cnn1 o cnn2 collapse (window sharing is not used by the current head kernel).
FLOP_INFER_DIRECT = 388_426_048
HEAD_MAC_DIRECT_PER_WINDOW = 38_793_216
HEAD_MAC_EXEC_PER_WINDOW = 246 * 32 * 5 * 64 + 8 * 2 * 246 * 32 * 160       # collapsed embed (sum Cz = 64) + 8 zones x (cnn3 + cnn4)
# tensor-core head: MMA flops actually issued per window (M = 128 rows x 2 streams; per K16 step one MMA per activation
# term with N = 32 * (terms - i); embed: 3 K16 steps for zones of <= 8 electrodes, 5 otherwise; cnn3 / cnn4: 10 each)
PROD_ZONE_SIZES = (6, 9, 6, 7, 7, 10, 15, 4)


def tc_mma_flop_per_window(n_terms):
    n_sum = sum(32 * (n_terms - i) for i in range(n_terms))
    steps = sum((3 if cz <= 8 else 5) + 20 for cz in PROD_ZONE_SIZES)
    return 2 * steps * 2.0 * 128 * 16 * n_sum
FLOP_TRAIN_DIRECT = 1_140_087_744
BYTES_PER_TRIAL = 64 * 795 * 4


def prod_cfg(dropout=0.1):
    from fast.data import Electrodes, Zones
    return types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800,
                                 window_len=250, slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4,
                                 num_heads=8, dropout=dropout)


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi sampled every 200 ms while the timed region runs (B200_PROFILING.md 'clocks' line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_forward_trials_per_s(n_trials, chunk, repeats, threads):
    """The oracle port of the reference forward on the host cores: eval-mode logits for `n_trials` synthetic trials."""
    from fast.models import FAST
    from oracle import fast_oracle as O
    torch.set_num_threads(threads)
    cfg = prod_cfg()
    torch.manual_seed(42)
    sd = {k: v.detach().clone() for k, v in FAST(cfg).state_dict().items()}
    x = torch.randn(n_trials, 64, 795, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        O.forward(sd, cfg, x[:min(chunk, 8)])                                  # warm-up (oneDNN primitive cache)
        best = float("inf")
        for _ in range(repeats):
            t0 = time.perf_counter()
            for c in x.split(chunk):
                O.forward(sd, cfg, c).argmax(1)
            best = min(best, time.perf_counter() - t0)
    return n_trials / best, best


def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n = 64
    times = []
    from fast.models import FAST
    from oracle import fast_oracle as O
    torch.set_num_threads(threads)
    cfg = prod_cfg()
    torch.manual_seed(42)
    sd = {k: v.detach().clone() for k, v in FAST(cfg).state_dict().items()}
    x = torch.randn(n, 64, 795, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            O.forward(sd, cfg, x).argmax(1)
            if i >= args.warmup:
                times.append(time.perf_counter() - t0)
    total = sum(times)
    value = n * len(times) / total
    sample = f"{n} of the {args.batch} trials of one batch per step (bounded sample), {threads} torch threads, oracle port of the reference forward"
    line = {"impl": "reference", "metric": "FAST inference trials/s", "value": value, "unit": "trials/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args),
            "cpu_baseline": {"value": value, "unit": "trials/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(args):
    return {"workload": "FAST inference, BASELINE configs[1] (batch sweep top end)", "batch_per_gpu": args.batch, "channels": 64,
            "head_precision": getattr(args, "head_precision", None),
            "samples": 795, "window_len": 250, "slide_step": 125, "zones": 8, "dim_cnn": 32, "dim_token": 32, "layers": 4,
            "heads": 8, "classes": 5, "l2_policy": "input per step (1.67 GB at batch 8192) larger than the 126 MB L2",
            "parallelism": f"trial-sharded x{args.gpus}, no data-path collective"}


# ------------------------------------------------------------------------------------------------ B200 arm
def run_b200(args, rank, world, local_rank):
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback for the hot path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    from fast import _native as N
    from fast.models import FAST
    from fast.train import FusedTrainer
    from fast.models import fast as FM
    N.lib()                                                                     # fail loudly if the extension is missing
    FM.set_head_precision(args.head_precision)
    n_terms = FM._TERMS[args.head_precision]
    B, K, W = args.batch, args.steps, args.warmup
    cfg = prod_cfg()
    torch.manual_seed(42)
    model = FAST(cfg).to(dev).eval()
    gen = torch.Generator(device=dev).manual_seed(rank)
    x = torch.randn(B, 64, 795, device=dev, generator=gen)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    head_ev = []

    def step(record):
        with torch.no_grad():
            if record:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            feat = model.forward_head(x)
            if record:
                e1.record()
                head_ev.append((e0, e1))
            return model.forward_transformer(feat).argmax(1)

    for _ in range(W):
        step(False)
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        t0.record()
        for _ in range(K):
            pred = step(True)
        t1.record()
        barrier()
    ms = t0.elapsed_time(t1)
    head_ms = float(np.mean([a.elapsed_time(b) for a, b in head_ev]))
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    value = world * B * K / (ms * 1e-3)

    # ---- end to end through the public API, batch in pinned host memory, predictions read back to the host
    xh = torch.empty(B, 64, 795, pin_memory=True)
    xh.copy_(x)
    ph = torch.empty(B, dtype=torch.int64, pin_memory=True)

    def e2e_step():
        ph.copy_(model.predict(xh, chunk=args.e2e_chunk), non_blocking=True)

    for _ in range(max(3, W // 2)):
        e2e_step()
    barrier()
    Ke = max(3, K // 2)
    t0.record()
    for _ in range(Ke):
        e2e_step()
    t1.record()
    barrier()
    e2e_ms = t0.elapsed_time(t1)
    if dist is not None:
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t)
    e2e_value = world * B * Ke / (e2e_ms * 1e-3)
    assert torch.equal(ph.to(dev), pred), "end-to-end predictions differ from the device-resident run"

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (the zone-embedding head: 99.9 % of the forward arithmetic)
    peaks = {}
    pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk_path):
        peaks = json.load(open(pk_path))
    n_win = 5
    alg_flop = 2.0 * HEAD_MAC_DIRECT_PER_WINDOW * n_win * B                      # direct form, as the reference computes it
    exe_flop = 2.0 * HEAD_MAC_EXEC_PER_WINDOW * n_win * B                        # fp32-equivalent flops after the exact collapse
    ach = alg_flop / (head_ms * 1e-3) / 1e12
    exe = exe_flop / (head_ms * 1e-3) / 1e12
    alg_bytes = B * 64 * 750 * 4 + B * n_win * 8 * 32 * 4
    traffic = None
    tr_path = os.path.join(ROOT, "profiles", "head_fwd_traffic.json")           # dram bytes per launch from ncu --set full
    if os.path.exists(tr_path):
        t = json.load(open(tr_path)).get(f"{args.head_precision}@{B}")
        traffic = t["dram_bytes"] if t else None
    if n_terms:
        tensor_peak = peaks.get("bf16_tflops_sustained", 1400.0)                # timed inside a long step -> sustained figure
        mma = tc_mma_flop_per_window(n_terms) * n_win * B / (head_ms * 1e-3) / 1e12
        roofline = {"kernel": f"head_fwd_tc_kernel<{n_terms}>", "bound": "tensor", "achieved": ach, "peak": tensor_peak,
                    "unit": "TFLOP/s", "frac": ach / tensor_peak, "fp32_equivalent_executed": exe,
                    "bf16_mma_issued": mma, "frac_mma_issued": mma / tensor_peak,
                    "peak_source": ("bf16_tflops_sustained of measured (MEASURED_PEAKS.json)" if peaks else "1.4 PFLOP/s sustained of fallback"),
                    "note": "achieved = direct-form (reference) head flops / CUDA-event time of the kernel; each fp32 operand is "
                            f"{n_terms} bf16 terms, so the tensor pipe issues bf16_mma_issued; the SS-mode N<=96 MMA shape is "
                            "shared-memory-operand bound at 47-57 cycles per M128 K16 MMA (tools/umma_probe.cu), see DESIGN.md",
                    "ms_per_launch": head_ms, "traffic": traffic, "algorithmic_bytes_per_launch": alg_bytes}
    else:
        fp32_peak = 148 * 128 * 2 * (peaks.get("sm_max_mhz", 1965.0) * 1e6) / 1e12  # FFMA lanes x 2 flop x clock
        roofline = {"kernel": "head_fwd_kernel<32>", "bound": "fp32-fma (CUDA-core kernel)", "achieved": ach, "peak": fp32_peak,
                    "unit": "TFLOP/s", "frac": ach / fp32_peak, "fp32_equivalent_executed": exe, "frac_executed": exe / fp32_peak,
                    "peak_source": "FP32 FMA pipe = 148 SM x 128 lanes x 2 x sm_max_mhz (MEASURED_PEAKS.json clock)",
                    "ms_per_launch": head_ms, "traffic": traffic, "algorithmic_bytes_per_launch": alg_bytes}

    extra = {}
    if not args.no_extras:
        extra = extras(model, cfg, dev, x)

    threads = os.cpu_count() or 1
    n_cpu = args.cpu_trials
    cpu_v, cpu_s = cpu_forward_trials_per_s(n_cpu, 256, 1, threads)
    line = {"metric": "FAST inference trials/s", "value": value, "unit": "trials/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": (f"bf16x{n_terms} split operands on tcgen05, f32 accumulate (f32 in / f32 out)" if n_terms else "f32"),
            "data": "synthetic", "config": workload_config(args), "clocks": clocks.summary(),
            "e2e": {"value": e2e_value, "unit": "trials/s", "h2d_bytes_per_step": B * BYTES_PER_TRIAL, "d2h_bytes_per_step": B * 8,
                    "steps": Ke, "ms_per_step": e2e_ms / Ke},
            "gpu_launches": 3 * K, "roofline": roofline,
            "cpu_baseline": {"value": cpu_v, "unit": "trials/s", "cores": threads, "kind": "port",
                             "sample": f"{n_cpu} of the {B} trials of one batch, in chunks of 256 ({cpu_s:.1f} s), oracle port of the reference forward"},
            "extra": extra}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def extras(model, cfg, dev, x):
    """Secondary numbers: inference batch sweep and the fused training step (fwd + bwd + AdamW)."""
    from fast.models import FAST
    from fast.train import FusedTrainer
    out = {"infer_sweep_trials_per_s": {}, "train_trials_per_s": {}}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.no_grad():
        for b in (64, 256, 1024, 4096):
            xb = x[:b]
            for _ in range(5):
                model(xb)
            n = 20
            e0.record()
            for _ in range(n):
                model(xb).argmax(1)
            e1.record()
            torch.cuda.synchronize(dev)
            out["infer_sweep_trials_per_s"][str(b)] = b * n / (e0.elapsed_time(e1) * 1e-3)
        from fast.models import fast as FM
        cur = FM.get_head_precision()
        out["infer_full_batch_by_head_precision"] = {}
        for mode in ("bf16x3", "bf16x2", "fp32"):                      # same batch, the three head arithmetics
            FM.set_head_precision(mode)
            for _ in range(3):
                model(x)
            n = 5
            e0.record()
            for _ in range(n):
                model(x).argmax(1)
            e1.record()
            torch.cuda.synchronize(dev)
            out["infer_full_batch_by_head_precision"][mode] = x.shape[0] * n / (e0.elapsed_time(e1) * 1e-3)
        FM.set_head_precision(cur)
    torch.manual_seed(42)
    tm = FAST(cfg).to(dev).train()
    ft = FusedTrainer(tm, lr=5e-4, use_graph=True)
    for b in (64, 1024):
        xb = x[:b]
        yb = torch.randint(0, 5, (b,), device=dev).to(torch.uint8)
        for _ in range(5):
            ft.train_step(xb, yb)
        n = 20
        e0.record()
        for _ in range(n):
            ft.train_step(xb, yb)
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / n
        out["train_trials_per_s"][str(b)] = b / (ms * 1e-3)
        out.setdefault("train_tflops_direct", {})[str(b)] = FLOP_TRAIN_DIRECT * b / (ms * 1e-3) / 1e12
    out["train_note"] = "FusedTrainer: fwd + CE + bwd + AdamW, dropout 0.1, CUDA graph replay, batch resident on device"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--head-precision", default=os.environ.get("FASTB200_HEAD_PRECISION", "bf16x3"), choices=["bf16x3", "bf16x2", "fp32"])
    ap.add_argument("--e2e-chunk", type=int, default=1024, help="trials per H2D piece of the end-to-end run")
    ap.add_argument("--cpu-trials", type=int, default=4096, help="trials of the bounded CPU-baseline sample (~10 s)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
