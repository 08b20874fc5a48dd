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


# ------------------------------------------------------------------------------------------------ other workloads
def _dist_setup(world, local_rank):
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    return dev, dist


def _max_over_ranks(ms, dist, dev):
    if dist is None:
        return ms
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def _emit(line, dist, rank):
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_train(args, rank, world, local_rank):
    """FAST training step (fwd + CE + bwd + AdamW) -- BASELINE metric 'train fwd+bwd' and configs[0]'s batch 64.
    Weak scaling: every rank trains its own replica on its own batch (the CV regime: independent jobs)."""
    dev, dist = _dist_setup(world, local_rank)
    from fast.models import FAST
    from fast.train import FusedTrainer
    from fast.models import fast as FM
    FM.set_head_precision(args.head_precision)
    cfg = prod_cfg()
    B, K, W = args.batch, args.steps, args.warmup
    torch.manual_seed(42 + rank)
    model = FAST(cfg).to(dev).train()
    ft = FusedTrainer(model, lr=5e-4, use_graph=True)
    nb = max(2, -(-200_000_000 // (B * BYTES_PER_TRIAL)))          # rotate over > L2 worth of distinct batches
    xs = torch.randn(nb, B, 64, 795, device=dev)
    ys = torch.randint(0, 5, (nb, B), device=dev).to(torch.uint8)
    for i in range(W):
        ft.train_step(xs[i % nb], ys[i % nb])
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        e0.record()
        for i in range(K):
            ft.train_step(xs[i % nb], ys[i % nb])
        e1.record()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)
    ms = _max_over_ranks(e0.elapsed_time(e1), dist, dev)
    value = world * B * K / (ms * 1e-3)
    # end to end: batch in pinned host memory, loss read back
    xh = torch.empty(B, 64, 795).pin_memory(); xh.copy_(xs[0])
    yh = ys[0].cpu().pin_memory()
    lh = torch.empty(1).pin_memory()
    for _ in range(3):
        ft.train_step(xh.to(dev, non_blocking=True), yh.to(dev, non_blocking=True))
    torch.cuda.synchronize(dev)
    Ke = max(3, K // 2)
    e0.record()
    for _ in range(Ke):
        loss, _ = ft.train_step(xh.to(dev, non_blocking=True), yh.to(dev, non_blocking=True))
        lh.copy_(loss, non_blocking=True)
    e1.record()
    torch.cuda.synchronize(dev)
    e2e_ms = _max_over_ranks(e0.elapsed_time(e1), dist, dev)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    fp32_peak = 148 * 128 * 2 * (peaks.get("sm_max_mhz", 1965.0) * 1e6) / 1e12
    ach = FLOP_TRAIN_DIRECT * B * K / (ms * 1e-3) / 1e12 / 1.0
    cpu = None
    if rank == 0 and not args.no_extras:
        cpu = cpu_train_trials_per_s(64, os.cpu_count() or 1)
    line = {"metric": "FAST training trials/s (fwd+bwd+AdamW)", "value": value, "unit": "trials/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": _train_dtype(FM),
            "data": "synthetic", "clocks": clocks.summary(),
            "config": {"workload": "FAST training step, production config, dropout 0.1 (BASELINE metric 'train fwd+bwd'; configs[0] batch = 64)",
                       "batch_per_gpu": B, "channels": 64, "samples": 795, "l2_policy": f"{nb} distinct resident batches rotated (> 126 MB)",
                       "parallelism": f"independent replicas x{world} (CV regime), no collective"},
            "e2e": {"value": world * B * Ke / (e2e_ms * 1e-3), "unit": "trials/s", "h2d_bytes_per_step": B * BYTES_PER_TRIAL + B,
                    "d2h_bytes_per_step": 4, "steps": Ke},
            "gpu_launches": ft.LAUNCHES_TRAIN * K,
            "roofline": _train_roofline(FM, ach / world, fp32_peak, peaks, B, ms / K),
            "cpu_baseline": cpu}
    _emit(line, dist, rank)


def _train_dtype(FM):
    p = FM.get_head_precision()
    if FM.get_head_backward() == "tc" and p != "fp32":
        return f"{p} split operands on tcgen05, f32 accumulate (f32 parameters, gradients and AdamW state)"
    return "f32"


def _train_roofline(FM, ach, fp32_peak, peaks, B, step_ms):
    """Roofline of the whole training step: with the tensor-core head (forward, input-gradient chain, weight gradients all
    tcgen05, split-bf16 operands) the bound is the tensor pipe; the step also moves ~16 MB of HBM per trial (saved
    activations written by the forward and re-read by both backward kernels), so the HBM fraction is reported beside it."""
    tc = FM.get_head_backward() == "tc" and FM.get_head_precision() != "fp32"
    if not tc:
        return {"kernel": "whole step (head_fwd + head_bwd dominate)", "bound": "fp32-fma (CUDA-core kernels)", "achieved": ach,
                "peak": fp32_peak, "unit": "TFLOP/s", "frac": ach / fp32_peak, "traffic": None,
                "note": "direct-form training flops (SURVEY 8d: 1,140,087,744 per trial) / step time, per GPU"}
    tp = peaks.get("bf16_tflops_sustained", 1400.0)
    hbm = peaks.get("hbm_gbs", 6500.0)
    traffic = 16.4e9 * B / 1024                  # ncu --set full at batch 1024 (profiles/r01_train_tc_ncu_full.txt), linear in B
    return {"kernel": "whole step (head_fwd_tc<SAVE> + head_dgrad_tc + head_wgrad_tc = 80 % of it)", "bound": "tensor",
            "achieved": ach, "peak": tp, "unit": "TFLOP/s", "frac": ach / tp, "traffic": traffic,
            "hbm_frac": traffic / (step_ms * 1e-3) / 1e9 / hbm,
            "peak_source": ("bf16_tflops_sustained / hbm_gbs of MEASURED_PEAKS.json" if peaks else "fallback 1400 TFLOP/s, 6500 GB/s"),
            "note": "achieved = direct-form training flops (SURVEY 8d: 1,140,087,744 per trial) / step time, per GPU; operands are "
                    "2-3 bf16 terms each, so the tensor pipe issues 2-3x that; traffic = dram bytes of one step scaled from the "
                    "batch-1024 ncu capture"}


def cpu_train_trials_per_s(batch, threads):
    """Oracle port of the reference's training step (CE + autograd + AdamW, trainer.py:48-67) on the host cores."""
    from fast.models import FAST
    from oracle import fast_oracle as O
    torch.set_num_threads(threads)
    cfg = prod_cfg(dropout=0.0)
    torch.manual_seed(42)
    sd = {k: v.detach().clone() for k, v in FAST(cfg).state_dict().items()}
    m = {k: torch.zeros_like(v) for k, v in sd.items()}
    v = {k: torch.zeros_like(t) for k, t in sd.items()}
    g = torch.Generator().manual_seed(0)
    x = torch.randn(batch, 64, 795, generator=g)
    y = torch.randint(0, 5, (batch,), generator=g)
    ts = []
    for step in range(1, 4):
        t0 = time.perf_counter()
        _, _, grads, _ = O.loss_and_grads(sd, cfg, x, y)
        for k in sd:
            O.adamw_step(sd[k], grads[k], m[k], v[k], step, 5e-4)
        ts.append(time.perf_counter() - t0)
    best = min(ts[1:])
    return {"value": batch / best, "unit": "trials/s", "cores": threads, "kind": "port",
            "sample": f"one batch of {batch} trials per step, best of 2 after 1 warm-up ({best:.1f} s/step), oracle port of the training step"}


def run_preprocess(args, rank, world, local_rank):
    """BASELINE configs[2]: zero-phase Butterworth band-pass + 4/5 polyphase resample of 15 subjects x 400 trials
    (SURVEY 8d-D3), subjects sharded over ranks (no collective)."""
    dev, dist = _dist_setup(world, local_rank)
    from fast.data import filters as F
    from oracle import filter_oracle as FO
    subs = list(range(15))[rank::world]
    n_tr = 400
    xs = []
    for s_ in subs:
        rng = np.random.default_rng(1000 + s_)
        x = (20 * rng.standard_normal((n_tr, 64, 795)) + rng.uniform(-50, 50, (1, 64, 1))).astype(np.float32)
        xs.append(torch.from_numpy(x).to(dev))
    sos = F.bandpass_sos()

    def step():
        out = None
        for x in xs:
            out = F.sosfiltfilt_resample(sos, x, 4, 5, out_dtype=torch.float32)    # one fused kernel per subject block
        return out

    K, W = args.steps, args.warmup
    for _ in range(W):
        step()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        e0.record()
        for _ in range(K):
            step()
        e1.record()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)
    ms = _max_over_ranks(e0.elapsed_time(e1), dist, dev)
    # the step IS the dominant kernel (one fused launch per subject block); the unfused two-kernel chain beside it
    filt_ms = e0.elapsed_time(e1) / (K * max(1, len(xs)))
    f0.record()
    for _ in range(K):
        for x in xs:
            F.resample_poly(F.sosfiltfilt(sos, x, out_dtype=torch.float64), 4, 5, out_dtype=torch.float32)
    f1.record()
    torch.cuda.synchronize(dev)
    chain_ms = f0.elapsed_time(f1) / (K * max(1, len(xs)))
    total_trials = 15 * n_tr
    value = total_trials * K / (ms * 1e-3)
    # end to end from pinned host memory
    xh = [x.cpu().pin_memory() for x in xs]
    for _ in F.bandpass_resample_host(xh, depth=3):       # warm-up: pinned output blocks enter the host allocator's cache
        pass
    torch.cuda.synchronize(dev)
    e0.record()
    last = None
    for oi in F.bandpass_resample_host(xh, depth=3):      # the public host-to-host call: copies overlap the kernel on side streams
        last = oi
    e1.record()
    torch.cuda.synchronize(dev)
    e2e_ms = _max_over_ranks(e0.elapsed_time(e1), dist, dev)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = peaks.get("hbm_gbs", 6500.0)
    alg = n_tr * 64 * (795 + 636) * 4                                   # f32 read + f32 resampled write per launch
    # fp64 work per row as the reference arithmetic defines it (2 passes x 4 sections x 9 flop per extended sample)
    # plus the 26-tap polyphase dot products; the chunked scan executes ~2.6x the recursion flops (two sweeps + the scan)
    fp64_alg = n_tr * 64 * (72.0 * (795 + 54) + 2.0 * 26 * 636)
    fp64_peak = 148 * 64 * 2 * (peaks.get("sm_max_mhz", 1965.0) * 1e6) / 1e12      # 16 DFMA lanes x 4 sub-partitions per SM
    traffic = None
    tr_path = os.path.join(ROOT, "profiles", "filter_traffic.json")
    if os.path.exists(tr_path):
        traffic = json.load(open(tr_path)).get("sosfiltfilt_resample@400x64x795", {}).get("dram_bytes")
    cpu = None
    if rank == 0:
        nc = 2
        xc = xs[0][:nc].cpu().numpy().reshape(-1, 795)
        t0 = time.perf_counter()
        yc = FO.sosfiltfilt(sos, xc)
        FO.resample_poly(yc, 4, 5)
        dt = time.perf_counter() - t0
        cpu = {"value": nc / dt, "unit": "trials/s", "cores": 1, "kind": "port",
               "sample": f"{nc} trials x 64 channels ({dt:.1f} s), NumPy float64 oracle port of scipy sosfiltfilt + resample_poly"}
    line = {"metric": "preprocessing trials/s (sosfiltfilt 4-40 Hz order 4 + resample_poly 4/5)", "value": value, "unit": "trials/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64 (f32 in, f32 out)", "data": "synthetic", "clocks": clocks.summary(),
            "config": {"workload": "BASELINE configs[2]: 15 subjects x 400 trials x 64 ch x 795 samples", "subjects_this_rank": len(subs),
                       "l2_policy": "every subject's block is 81 MB in / 163 MB intermediate; 15 blocks per step (> 126 MB L2)",
                       "parallelism": f"subjects sharded x{world}, no collective"},
            "e2e": {"value": total_trials / (e2e_ms * 1e-3),
                    "unit": "trials/s", "h2d_bytes_per_step": len(subs) * n_tr * 64 * 795 * 4, "d2h_bytes_per_step": len(subs) * n_tr * 64 * 636 * 4},
            "gpu_launches": len(xs) * K,
            "roofline": {"kernel": "sosfiltfilt_resample_kernel<4,4,5,26,2,float,float>", "bound": "hbm",
                         "achieved": alg / (filt_ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                         "frac": alg / (filt_ms * 1e-3) / 1e9 / hbm, "traffic": traffic, "ms_per_launch": filt_ms,
                         "algorithmic_bytes_per_launch": alg,
                         "fp64_pipe": {"achieved": fp64_alg / (filt_ms * 1e-3) / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
                                       "frac": fp64_alg / (filt_ms * 1e-3) / 1e12 / fp64_peak},
                         "two_kernel_chain_ms": chain_ms,
                         "note": "the 1e-5 parity bar forces an fp64 recursion (SURVEY 7.3-h), so the kernel sits on the FP64 pipe, "
                                 "not on HBM: ncu shows sm__pipe_fp64_cycles_active ~70 % (profiles/), dram < 15 %"},
            "cpu_baseline": cpu}
    _emit(line, dist, rank)


def run_cv(args, rank, world, local_rank):
    """BASELINE configs[3]: 15 subjects x 5 folds of FAST training, jobs sharded over ranks (strong scaling, no data-path
    collective), synthetic subjects.  value = wall-clock seconds from first dispatch to the last result."""
    import shutil
    import tempfile
    dev, dist = _dist_setup(world, local_rank)
    from fast.data import SUBJECTS
    from fast.models import fast as FM
    from fast.train import cv as CV
    FM.set_head_precision(args.head_precision)
    cfg = prod_cfg()
    subjects = SUBJECTS[:args.subjects]
    jobs = CV.cv_jobs(subjects, 5)
    mine = CV.shard_jobs(jobs, rank, world)
    save = args.cv_dir or os.path.join(tempfile.gettempdir(), "fastb200_cv")
    if rank == 0:
        shutil.rmtree(save, ignore_errors=True)
        os.makedirs(save, exist_ok=True)
    if dist is not None:
        dist.barrier()                                  # nobody writes into the tree before rank 0 has cleaned it
    data = {}

    def load(sid):
        if sid not in data:
            data.clear()
            data[sid] = CV.synthetic_subject(int(sid) - 1)
        return data[sid][0]

    # warm-up: one short job compiles nothing but builds graphs / primes allocators (not part of the 75 jobs)
    CV.run_cv_jobs(cfg, [(0, "warm", 0)], lambda s: CV.synthetic_subject(99)[0], os.path.join(save, f"_warm{rank}"), max_epochs=2,
                   batch_size=args.cv_batch, seed=1)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(dev)
    with ClockSampler(local_rank) as clocks:
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        done = CV.run_cv_jobs(cfg, mine, load, save, max_epochs=args.epochs, batch_size=args.cv_batch, seed=42, resume=False)
        e1.record()
        torch.cuda.synchronize(dev)
        dev_s = e0.elapsed_time(e1) * 1e-3
        if dist is not None:
            dist.barrier()
        wall = time.perf_counter() - t0
    dev_s = _max_over_ranks(dev_s, dist, dev)
    wall = _max_over_ranks(wall, dist, dev)
    rows = []
    if rank == 0:
        tests = {sid: CV.synthetic_subject(int(sid) - 1)[1] for sid in subjects}
        rows = CV.finalize_cv(cfg, subjects, save, n_folds=5, test_per_subject=tests)
    n_train_trials = len(jobs) * args.epochs * 280
    steps = len(jobs) * args.epochs * 5
    line = {"metric": "15-subject x 5-fold CV wall-clock", "value": dev_s, "unit": "s", "n_gpus": world, "steps": steps, "warmup": 1,
            "ms_per_step": 1e3 * dev_s / max(1, len(mine) * args.epochs * 5), "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": _train_dtype(FM), "data": "synthetic", "clocks": clocks.summary(),
            "config": {"workload": "BASELINE configs[3]: per-subject 5-fold CV of FAST, batch 64", "jobs": len(jobs), "jobs_this_rank": len(mine),
                       "epochs": args.epochs, "subjects": len(subjects), "trials_per_subject": 350, "batch": args.cv_batch,
                       "parallelism": f"subject x fold jobs sharded x{world}, no collective", "host_wall_s": wall,
                       "ideal_rounds": -(-len(jobs) // world)},
            "train_trials_per_s": n_train_trials / dev_s, "subjects_done": len(rows),
            "mean_test_acc": (float(np.mean([r[2] for r in rows])) if rows else None),
            "e2e": {"value": wall, "unit": "s", "h2d_bytes_per_step": 350 * 64 * 800 * 4 // (5 * args.epochs * 5), "d2h_bytes_per_step": 0},
            "gpu_launches": len(mine) * args.epochs * (5 * 9 + 4)}
    _emit(line, dist, rank)


def run_stress(args, rank, world, local_rank):
    """BASELINE configs[4]: 64 ch x 4000-sample trials (31 windows, 32 tokens), data-parallel training with the flat
    gradient exchanged every step.  Weak scaling: args.batch trials per GPU (512 x 8 = 4096 at 8 GPUs)."""
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(local_rank)
    import torch.distributed as dist
    if world == 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29533")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    else:
        dist.init_process_group("nccl", device_id=dev)
    from fast.models import FAST
    from fast.models import fast as FM
    from fast.train import DataParallelTrainer
    FM.set_head_precision(args.head_precision)
    cfg = prod_cfg()
    cfg.seq_len = 4000
    B, K, W = args.batch, args.steps, args.warmup
    torch.manual_seed(42)
    model = FAST(cfg).to(dev).train()
    ft = DataParallelTrainer(model, reduce=args.reduce, lr=5e-4, use_graph=True)
    gen = torch.Generator(device=dev).manual_seed(rank)
    x = torch.randn(B, 64, 4000, device=dev, generator=gen)
    y = torch.randint(0, 5, (B,), device=dev, generator=gen).to(torch.uint8)
    for _ in range(W):
        ft.train_step(x, y)
    dist.barrier(); torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        e0.record()
        for _ in range(K):
            ft.train_step(x, y)
        e1.record()
        dist.barrier(); torch.cuda.synchronize(dev)
    ms = _max_over_ranks(e0.elapsed_time(e1), dist, dev)
    # exposed exchange time: the same K steps of the update phase alone (barrier + fused reduce/AdamW or NCCL + AdamW)
    ws = next(w for k, w in ft._ws.items() if isinstance(w, dict) and w.get("graph"))
    upd = [fn for fn in ws["graph"][1:]] if world > 1 else []
    e0.record()
    for _ in range(K):
        for fn in upd:
            fn()
    e1.record()
    torch.cuda.synchronize(dev)
    upd_ms = e0.elapsed_time(e1) / K if upd else 0.0
    loss, _ = ft.global_loss_and_correct()
    flop = 2.0 * (31 * (38_793_216 + 75_067_392) + 3 * 1_564_832) * 1.0
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    tensor_peak = (json.load(open(pk)) if os.path.exists(pk) else {}).get("bf16_tflops_sustained", 1400.0)
    line = {"metric": "FAST stress training trials/s (64 ch x 4000 samples, data-parallel)", "value": world * B * K / (ms * 1e-3),
            "unit": "trials/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": _train_dtype(FM), "data": "synthetic", "clocks": clocks.summary(),
            "config": {"workload": "BASELINE configs[4]: seq_len 4000 (31 windows, 32 tokens), DP", "batch_per_gpu": B,
                       "global_batch": B * world, "reduce": args.reduce if world > 1 else "none",
                       "l2_policy": f"input {B * 64 * 4000 * 4 / 1e9:.2f} GB per step > 126 MB L2",
                       "parallelism": f"dp{world}: one flat 767 KB gradient exchange per step"},
            "exchange_plus_update_ms": upd_ms, "loss": loss,
            "e2e": None, "gpu_launches": 9 * K,
            "roofline": {"kernel": "whole step (tcgen05 head kernels)", "bound": "tensor", "achieved": flop * B * K / (ms * 1e-3) / 1e12,
                         "peak": tensor_peak, "unit": "TFLOP/s", "frac": flop * B * K / (ms * 1e-3) / 1e12 / tensor_peak, "traffic": None,
                         "note": "direct-form training flops per trial / step time, per GPU, against bf16_tflops_sustained"}}
    if rank == 0:
        print(json.dumps(line), flush=True)
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--head-precision", default=os.environ.get("FASTB200_HEAD_PRECISION", "bf16x2"), choices=["bf16x3", "bf16x2", "fp32"])
    ap.add_argument("--e2e-chunk", type=int, default=1024, help="trials per H2D piece of the end-to-end run")
    ap.add_argument("--cpu-trials", type=int, default=4096, help="trials of the bounded CPU-baseline sample (~10 s)")
    ap.add_argument("--workload", default="infer", choices=["infer", "train", "preprocess", "cv", "stress"],
                    help="infer = the headline (BASELINE configs[1]); the others measure configs[0]/[2]/[3]/[4]")
    ap.add_argument("--epochs", type=int, default=200, help="cv: epochs per fold (reference: 200)")
    ap.add_argument("--subjects", type=int, default=15, help="cv: number of synthetic subjects")
    ap.add_argument("--cv-batch", type=int, default=64)
    ap.add_argument("--cv-dir", default=None)
    ap.add_argument("--reduce", default="p2p", choices=["p2p", "nccl"], help="stress: gradient exchange")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
    elif args.workload != "infer":
        if "--batch" not in sys.argv:
            args.batch = {"train": 1024, "stress": 512}.get(args.workload, args.batch)
        {"train": run_train, "preprocess": run_preprocess, "cv": run_cv, "stress": run_stress}[args.workload](args, rank, world, local_rank)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
