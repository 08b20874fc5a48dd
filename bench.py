"""Benchmark of the FAST hot path on B200 (contract: see the task statement / DESIGN.md section 5).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload all|infer|train|cv|preprocess|stress]

The JSON line's top-level keys are the headline workload (BASELINE.json configs[1]): FAST inference, production config,
batch 8192 synthetic BCIC2020-Track3-shaped trials (64 ch x 795 samples, fp32) per GPU -- 1.67 GB of input per step, larger
than the 126 MB L2.  One "step" = the whole hot path over one batch: window unfold + zone gather + 8 x Conv4Layers (one
kernel), input layer + CLS/pos + 4 transformer blocks + classifier (one kernel), argmax.

  value   trials/s, input already resident in HBM, K steps timed with CUDA events, max over ranks
  e2e     trials/s through the public drop-in API (``model.predict(x_host)``) with the batch in pinned HOST memory: the
          chunked H2D copy, the kernels and the D2H read of the predictions are all inside the timed region

BASELINE.json's metric has three legs -- "trials/s (train fwd+bwd, infer) at 1 GPU; 75-job CV time at 1/2/4/8 GPU" -- so
the default run (``--workload all``) measures ALL of them at this N and nests the other legs, each a complete line of the
same contract (value / e2e / roofline / cpu_baseline / gpu_launches), under ``extra``:

  extra.train_b1024, extra.train_b64   FusedTrainer step (fwd + CE + bwd + AdamW), independent replicas per rank
  extra.cv                             15 subjects x 5 folds x 200 epochs through run_cv_jobs, jobs sharded over the N ranks
  extra.preprocess                     fused zero-phase band-pass + 4/5 resample, subjects sharded over the N ranks
  extra.stress_p2p / extra.stress_nccl seq_len-4000 data-parallel step, flat-gradient exchange over NVLink (N >= 2: both
                                       reduce modes + the exposed exchange time); extra.dp_check = replica-equality evidence
  extra.eager_port_gpu                 the SAME model through stock torch ops (cuDNN / cuBLAS eager = what the reference
                                       runs on a GPU) on this B200: the same-box library baseline

``--workload X`` prints leg X alone as the top-level line.  Multi-GPU (torchrun): ranks shard trials / jobs / subjects
with NO data-path collective except the stress leg's one gradient exchange per step; timing = max over ranks.

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
# executed training MACs per window: forward + (dgrad of cnn4, cnn3) + (wgrad of cnn4, cnn3, collapsed embed)
HEAD_MAC_EXEC_TRAIN_PER_WINDOW = HEAD_MAC_EXEC_PER_WINDOW + 8 * 2 * 246 * 32 * 160 + 8 * 2 * 246 * 32 * 160 + 246 * 32 * 5 * 64
# tensor-core head: MMA flops actually issued per window (M = 128 rows x 2 streams; per K16 step one MMA per activation
# term with N = 32 * (terms - i); embed: 3 K16 steps for zones of <= 8 electrodes, 5 otherwise; cnn3 / cnn4: 10 each)
PROD_ZONE_SIZES = (6, 9, 6, 7, 7, 10, 15, 4)
FLOP_TRAIN_DIRECT = 1_140_087_744
BYTES_PER_TRIAL = 64 * 795 * 4
# ncu --set full dram bytes of one training step at batch 1024 (profiles/; see _train_roofline)
TRAIN_TRAFFIC_FILE = os.path.join(ROOT, "profiles", "train_traffic.json")


def tc_mma_flop_per_window(n_terms):
    n_sum = sum(32 * (n_terms - i) for i in range(n_terms))
    steps = sum((3 if cz <= 8 else 5) + 20 for cz in PROD_ZONE_SIZES)
    return 2 * steps * 2.0 * 128 * 16 * n_sum


def prod_cfg(dropout=0.1):
    from fast.data import Electrodes, Zones
    return types.SimpleNamespace(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800,
                                 window_len=250, slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4,
                                 num_heads=8, dropout=dropout)


def load_peaks():
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return json.load(open(pk)) if os.path.exists(pk) else {}


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


# ------------------------------------------------------------------------------------------------ process context
class Ctx:
    """One process per GPU.  NCCL is initialised when world > 1 (torchrun); the stress leg at world == 1 initialises a
    1-rank group lazily (DataParallelTrainer needs one)."""

    def __init__(self):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback for the hot path)"
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        self.dist = None
        if self.world > 1:
            self.ensure_dist()

    def ensure_dist(self):
        if self.dist is None:
            import torch.distributed as dist
            if self.world == 1:
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                os.environ.setdefault("MASTER_PORT", "29533")
                dist.init_process_group("nccl", rank=0, world_size=1, device_id=self.dev)
            else:
                dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        return self.dist

    def barrier(self):
        if self.dist is not None and self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, v):
        if self.dist is None or self.world == 1:
            return float(v)
        t = torch.tensor([v], device=self.dev, dtype=torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t)

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()
            self.dist = None


def launches():
    from fast import _native as N
    return N.launch_count()


# The contract is ONE JSON line on stdout.  NCCL (banner under NCCL_DEBUG=VERSION/INFO) and other native libraries write to
# file descriptor 1 directly, so for the B200 arm fd 1 is pointed at stderr for the whole run and the line goes to the saved
# descriptor.
_JSON_FD = None


def capture_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


# ------------------------------------------------------------------------------------------------ CPU arm (oracle port)
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


def run_reference(args, rank):
    """Reference arm: the CPU oracle port of the reference forward, all host threads, a bounded sample of the headline batch
    per step (--ref-trials of its 8192 trials -- the metric is per trial, so the rate carries over)."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n = args.ref_trials
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
    sample = (f"{n} of the {args.batch} trials of one batch per step (bounded sample: the full batch is ~20 s of CPU work per step), "
              f"{threads} torch threads, oracle port of the reference forward")
    cfg_line = workload_config(args)
    cfg_line["reference_sample_trials_per_step"] = n
    line = {"impl": "reference", "metric": "FAST inference trials/s", "value": value, "unit": "trials/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg_line,
            "cpu_baseline": {"value": value, "unit": "trials/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(args):
    return {"workload": "FAST inference, BASELINE configs[1] (batch sweep top end)", "batch_per_gpu": args.batch, "channels": 64,
            "head_precision": getattr(args, "head_precision", None),
            "samples": 795, "window_len": 250, "slide_step": 125, "zones": 8, "dim_cnn": 32, "dim_token": 32, "layers": 4,
            "heads": 8, "classes": 5, "l2_policy": "input per step (1.67 GB at batch 8192) larger than the 126 MB L2",
            "parallelism": f"trial-sharded x{args.gpus}, no data-path collective"}


# ------------------------------------------------------------------------------------------------ leg: inference (headline)
def measure_infer(args, ctx):
    from fast import _native as N
    from fast.models import FAST
    from fast.models import fast as FM
    N.lib()                                                                     # fail loudly if the extension is missing
    FM.set_head_precision(args.head_precision)
    n_terms = FM._TERMS[args.head_precision]
    B, K, W = args.batch, args.steps, args.warmup
    dev, world = ctx.dev, ctx.world
    cfg = prod_cfg()
    torch.manual_seed(42)
    model = FAST(cfg).to(dev).eval()
    gen = torch.Generator(device=dev).manual_seed(ctx.rank)
    x = torch.randn(B, 64, 795, device=dev, generator=gen)
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
    ctx.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(ctx.local_rank) as clocks:
        l0 = launches()
        t0.record()
        for _ in range(K):
            pred = step(True)
        t1.record()
        n_launch = launches() - l0
        ctx.barrier()
    ms = ctx.max_over_ranks(t0.elapsed_time(t1))
    head_ms = float(np.mean([a.elapsed_time(b) for a, b in head_ev]))      # includes the (tiny) head_pack_tc launch
    value = world * B * K / (ms * 1e-3)

    # ---- end to end through the public API, batch in pinned host memory, predictions read back to the host
    xh = torch.empty(B, 64, 795, pin_memory=True)
    xh.copy_(x)
    ph = torch.empty(B, dtype=torch.int64, pin_memory=True)

    def e2e_step():
        ph.copy_(model.predict(xh, chunk=args.e2e_chunk), non_blocking=True)

    for _ in range(max(3, W // 2)):
        e2e_step()
    ctx.barrier()
    Ke = max(3, K // 2)
    t0.record()
    for _ in range(Ke):
        e2e_step()
    t1.record()
    ctx.barrier()
    e2e_ms = ctx.max_over_ranks(t0.elapsed_time(t1))
    e2e_value = world * B * Ke / (e2e_ms * 1e-3)
    assert torch.equal(ph.to(dev), pred), "end-to-end predictions differ from the device-resident run"
    h2d = int(getattr(model, "_last_h2d_bytes", B * BYTES_PER_TRIAL))

    # ---- roofline of the dominant kernel (the zone-embedding head: 99.9 % of the forward arithmetic)
    peaks = load_peaks()
    n_win = 5
    alg_flop = 2.0 * HEAD_MAC_DIRECT_PER_WINDOW * n_win * B                      # direct form, as the reference computes it
    exe_flop = 2.0 * HEAD_MAC_EXEC_PER_WINDOW * n_win * B                        # fp32-equivalent flops after the exact collapse
    ach = alg_flop / (head_ms * 1e-3) / 1e12
    exe = exe_flop / (head_ms * 1e-3) / 1e12
    alg_bytes = B * 64 * 750 * 4 + B * n_win * 8 * 32 * 4
    traffic, traffic_src = None, None
    tr_path = os.path.join(ROOT, "profiles", "head_fwd_traffic.json")           # dram bytes per launch from ncu --set full
    if os.path.exists(tr_path):
        t = json.load(open(tr_path)).get(f"{args.head_precision}@{B}")
        if t:
            traffic = t["dram_bytes"]
            traffic_src = f"profiles/head_fwd_traffic.json ({t.get('source', 'ncu --set full capture')}); not re-measured in this run"
    if n_terms:
        # the timed region is K x ~8 ms at full clocks with no power cap active: burst conditions -> the burst bf16 peak
        tensor_peak = peaks.get("bf16_tflops", 1590.0)
        mma = tc_mma_flop_per_window(n_terms) * n_win * B / (head_ms * 1e-3) / 1e12
        roofline = {"kernel": f"head_fwd_tc_kernel<{n_terms}>", "bound": "tensor", "achieved": ach, "peak": tensor_peak,
                    "unit": "TFLOP/s", "frac": ach / tensor_peak, "fp32_equivalent_executed": exe,
                    "frac_executed": exe / tensor_peak,
                    "bf16_mma_issued": mma, "frac_mma_issued": mma / tensor_peak,
                    "frac_of_sustained_peak": ach / peaks.get("bf16_tflops_sustained", 1400.0),
                    "peak_source": ("bf16_tflops (burst) of measured (MEASURED_PEAKS.json)" if peaks else "1.59 PFLOP/s burst of fallback"),
                    "note": "achieved = direct-form (reference) head flops / CUDA-event time of the kernel; each fp32 operand is "
                            f"{n_terms} bf16 terms, so the tensor pipe issues bf16_mma_issued for them; DESIGN.md section 4.1 derives "
                            "why 3 bf16 products per fp32 product cap the direct-form fraction near 0.55 even at a 100 % busy tensor pipe",
                    "ms_per_launch": head_ms, "traffic": traffic, "traffic_source": traffic_src,
                    "algorithmic_bytes_per_launch": alg_bytes}
    else:
        fp32_peak = 148 * 128 * 2 * (peaks.get("sm_max_mhz", 1965.0) * 1e6) / 1e12  # FFMA lanes x 2 flop x clock
        roofline = {"kernel": "head_fwd_kernel<32>", "bound": "fp32-fma (CUDA-core kernel)", "achieved": ach, "peak": fp32_peak,
                    "unit": "TFLOP/s", "frac": ach / fp32_peak, "fp32_equivalent_executed": exe, "frac_executed": exe / fp32_peak,
                    "peak_source": "FP32 FMA pipe = 148 SM x 128 lanes x 2 x sm_max_mhz (MEASURED_PEAKS.json clock)",
                    "ms_per_launch": head_ms, "traffic": traffic, "traffic_source": traffic_src,
                    "algorithmic_bytes_per_launch": alg_bytes}
    line = {"metric": "FAST inference trials/s", "value": value, "unit": "trials/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": (f"bf16x{n_terms} split operands on tcgen05, f32 accumulate (f32 in / f32 out)" if n_terms else "f32"),
            "data": "synthetic", "config": workload_config(args), "clocks": clocks.summary(),
            "e2e": {"value": e2e_value, "unit": "trials/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": B * 8,
                    "steps": Ke, "ms_per_step": e2e_ms / Ke,
                    "note": "model.predict(x_host): pinned host batch streamed in chunks on a side stream; only the samples the "
                            "windows read (750 of 795 per row) are copied, as one 2-D async copy per chunk"},
            "gpu_launches": n_launch, "roofline": roofline, "cpu_baseline": None, "extra": {}}
    return line, model, x


def infer_sweeps(model, x, dev):
    """Secondary inference numbers: batch sweep (BASELINE configs[1] is a sweep 64-8192) and the three head arithmetics."""
    from fast.models import fast as FM
    out = {"infer_sweep_trials_per_s": {}, "infer_full_batch_by_head_precision": {}}
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
        cur = FM.get_head_precision()
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
    return out


# ------------------------------------------------------------------------------------------------ leg: training step
def _train_dtype(FM):
    p = FM.get_head_precision()
    if FM.get_head_backward() == "tc" and p != "fp32":
        return f"{p} split operands on tcgen05, f32 accumulate (f32 parameters, gradients and AdamW state)"
    return "f32"


def _train_roofline(FM, ach, exe, peaks, B, step_ms):
    """Roofline of the whole training step (per GPU).  With the tensor-core head (forward, input-gradient chain, weight
    gradients all tcgen05, split-bf16 operands) the bound is the tensor pipe; the step also streams the saved activations
    through HBM, so the HBM fraction is reported beside it (dram bytes from the committed ncu capture, scaled by batch)."""
    tc = FM.get_head_backward() == "tc" and FM.get_head_precision() != "fp32"
    fp32_peak = 148 * 128 * 2 * (peaks.get("sm_max_mhz", 1965.0) * 1e6) / 1e12
    if not tc:
        return {"kernel": "whole step (head_fwd + head_bwd dominate)", "bound": "fp32-fma (CUDA-core kernels)", "achieved": ach,
                "peak": fp32_peak, "unit": "TFLOP/s", "frac": ach / fp32_peak, "traffic": None,
                "note": "direct-form training flops (SURVEY 8d: 1,140,087,744 per trial) / step time, per GPU"}
    tp = peaks.get("bf16_tflops", 1590.0)
    hbm = peaks.get("hbm_gbs", 6650.0)
    traffic, src, per_kernel = None, None, None
    if os.path.exists(TRAIN_TRAFFIC_FILE):
        t = json.load(open(TRAIN_TRAFFIC_FILE))
        traffic = t["dram_bytes_per_1024_trials"] * B / 1024.0
        per_kernel = t.get("per_kernel")
        src = f"{t.get('source', 'ncu --set full at batch 1024')}, scaled linearly to batch {B}; not re-measured in this run"
    return {"kernel": "whole step (head_fwd_tc<SAVE> + head_dgrad_tc + head_wgrad_tc = 80-85 % of it)", "bound": "tensor",
            "achieved": ach, "peak": tp, "unit": "TFLOP/s", "frac": ach / tp, "fp32_equivalent_executed": exe,
            "frac_executed": exe / tp, "frac_of_sustained_peak": ach / peaks.get("bf16_tflops_sustained", 1400.0),
            "traffic": traffic, "traffic_source": src, "traffic_per_kernel_at_1024": per_kernel,
            "hbm_gbs_achieved": (traffic / (step_ms * 1e-3) / 1e9 if traffic else None),
            "hbm_frac": (traffic / (step_ms * 1e-3) / 1e9 / hbm if traffic else None),
            "algorithmic_bytes_per_step": B * 64 * 750 * 4 + 5 * 767_124,
            "peak_source": ("bf16_tflops (burst) / hbm_gbs of measured (MEASURED_PEAKS.json)" if peaks else "fallback 1590 TFLOP/s, 6650 GB/s"),
            "note": "achieved = direct-form training flops (SURVEY 8d: 1,140,087,744 per trial) / step time, per GPU; executed = "
                    "fp32-equivalent flops after the cnn1 o cnn2 collapse; operands are 2 bf16 terms each, so the tensor pipe issues "
                    "3-4x the executed flops"}


def measure_train(args, ctx, B):
    """FAST training step (fwd + CE + bwd + AdamW) -- BASELINE metric 'train fwd+bwd'; batch 64 is configs[0]'s batch.
    Weak scaling: every rank trains its own replica on its own batches (the CV regime: independent jobs)."""
    from fast.models import FAST
    from fast.train import FusedTrainer
    from fast.models import fast as FM
    FM.set_head_precision(args.head_precision)
    dev, world = ctx.dev, ctx.world
    cfg = prod_cfg()
    K, W = args.steps, max(args.warmup, 3)
    torch.manual_seed(42 + ctx.rank)
    model = FAST(cfg).to(dev).train()
    ft = FusedTrainer(model, lr=5e-4, use_graph=True)
    nb = max(2, -(-200_000_000 // (B * BYTES_PER_TRIAL)))          # rotate over > L2 worth of distinct batches
    xs = torch.randn(nb, B, 64, 795, device=dev)
    ys = torch.randint(0, 5, (nb, B), device=dev).to(torch.uint8)
    for i in range(W):
        ft.train_step(xs[i % nb], ys[i % nb])
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(ctx.local_rank) as clocks:
        l0 = launches()
        e0.record()
        for i in range(K):
            ft.train_step(xs[i % nb], ys[i % nb])
        e1.record()
        n_launch = launches() - l0
        ctx.barrier()
    ms = ctx.max_over_ranks(e0.elapsed_time(e1))
    value = world * B * K / (ms * 1e-3)
    # end to end: batch in pinned host memory, loss read back
    xh = torch.empty(B, 64, 795).pin_memory(); xh.copy_(xs[0])
    yh = ys[0].cpu().pin_memory()
    lh = torch.empty(1).pin_memory()
    for _ in range(3):
        ft.train_step(xh, yh)
    ctx.barrier()
    Ke = max(3, K // 2)
    e0.record()
    for _ in range(Ke):
        loss, _ = ft.train_step(xh, yh)                 # host batch: uploaded on the trainer's side stream, overlapping the previous step
        lh.copy_(loss, non_blocking=True)
    e1.record()
    ctx.barrier()
    e2e_ms = ctx.max_over_ranks(e0.elapsed_time(e1))
    peaks = load_peaks()
    ach = FLOP_TRAIN_DIRECT * B * K / (ms * 1e-3) / 1e12
    exe = (2.0 * HEAD_MAC_EXEC_TRAIN_PER_WINDOW * 5) * B * K / (ms * 1e-3) / 1e12
    del xs, ys
    return {"metric": "FAST training trials/s (fwd+bwd+AdamW)", "value": value, "unit": "trials/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": _train_dtype(FM), "data": "synthetic", "clocks": clocks.summary(),
            "config": {"workload": "FAST training step, production config, dropout 0.1 (BASELINE metric 'train fwd+bwd'; configs[0] batch = 64)",
                       "batch_per_gpu": B, "channels": 64, "samples": 795, "l2_policy": f"{nb} distinct resident batches rotated (> 126 MB)",
                       "parallelism": f"independent replicas x{world} (CV regime), no collective"},
            "e2e": {"value": world * B * Ke / (e2e_ms * 1e-3), "unit": "trials/s", "h2d_bytes_per_step": B * BYTES_PER_TRIAL + B,
                    "d2h_bytes_per_step": 4, "steps": Ke},
            "gpu_launches": n_launch,
            "roofline": _train_roofline(FM, ach, exe, peaks, B, ms / K),
            "cpu_baseline": None}


# ------------------------------------------------------------------------------------------------ leg: preprocessing
def measure_preprocess(args, ctx):
    """BASELINE configs[2]: zero-phase Butterworth band-pass + 4/5 polyphase resample of 15 subjects x 400 trials
    (SURVEY 8d-D3), subjects sharded over ranks (no collective)."""
    from fast.data import filters as F
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
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

    K, W = args.steps, max(args.warmup, 3)
    for _ in range(W):
        step()
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(ctx.local_rank) as clocks:
        l0 = launches()
        e0.record()
        for _ in range(K):
            step()
        e1.record()
        n_launch = launches() - l0
        ctx.barrier()
    ms = ctx.max_over_ranks(e0.elapsed_time(e1))
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
    ctx.barrier()
    e0.record()
    for _ in F.bandpass_resample_host(xh, depth=3):       # the public host-to-host call: copies overlap the kernel on side streams
        pass
    e1.record()
    ctx.barrier()
    e2e_ms = ctx.max_over_ranks(e0.elapsed_time(e1))
    peaks = load_peaks()
    hbm = peaks.get("hbm_gbs", 6650.0)
    alg = n_tr * 64 * (795 + 636) * 4                                   # f32 read + f32 resampled write per launch
    # fp64 work per row as the reference arithmetic defines it (2 passes x 4 sections x 9 flop per extended sample)
    # plus the 26-tap polyphase dot products; the chunked scan executes ~2.6x the recursion flops (two sweeps + the scan)
    fp64_alg = n_tr * 64 * (72.0 * (795 + 54) + 2.0 * 26 * 636)
    fp64_peak = 148 * 64 * 2 * (peaks.get("sm_max_mhz", 1965.0) * 1e6) / 1e12      # 16 DFMA lanes x 4 sub-partitions per SM
    traffic, traffic_src = None, None
    tr_path = os.path.join(ROOT, "profiles", "filter_traffic.json")
    if os.path.exists(tr_path):
        traffic = json.load(open(tr_path)).get("sosfiltfilt_resample@400x64x795", {}).get("dram_bytes")
        traffic_src = "profiles/filter_traffic.json (ncu --set full capture); not re-measured in this run"
    del xs
    return {"metric": "preprocessing trials/s (sosfiltfilt 4-40 Hz order 4 + resample_poly 4/5)", "value": value, "unit": "trials/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64 (f32 in, f32 out)", "data": "synthetic", "clocks": clocks.summary(),
            "config": {"workload": "BASELINE configs[2]: 15 subjects x 400 trials x 64 ch x 795 samples", "subjects_this_rank": len(subs),
                       "l2_policy": "every subject's block is 81 MB in / 65 MB out; 15 blocks per step (> 126 MB L2)",
                       "parallelism": f"subjects sharded x{world}, no collective"},
            "e2e": {"value": total_trials / (e2e_ms * 1e-3),
                    "unit": "trials/s", "h2d_bytes_per_step": len(subs) * n_tr * 64 * 795 * 4, "d2h_bytes_per_step": len(subs) * n_tr * 64 * 636 * 4},
            "gpu_launches": n_launch,
            "roofline": {"kernel": "sosfiltfilt_resample_kernel<4,4,5,26,2,float,float>", "bound": "hbm",
                         "achieved": alg / (filt_ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                         "frac": alg / (filt_ms * 1e-3) / 1e9 / hbm, "traffic": traffic, "traffic_source": traffic_src,
                         "ms_per_launch": filt_ms, "algorithmic_bytes_per_launch": alg,
                         "fp64_pipe": {"achieved": fp64_alg / (filt_ms * 1e-3) / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
                                       "frac": fp64_alg / (filt_ms * 1e-3) / 1e12 / fp64_peak},
                         "two_kernel_chain_ms": chain_ms,
                         "note": "the 1e-5 parity bar forces an fp64 recursion (SURVEY 7.3-h), so the kernel sits on the FP64 pipe, "
                                 "not on HBM: ncu shows sm__pipe_fp64_cycles_active ~62 % (profiles/), dram < 15 %"},
            "cpu_baseline": None}


def cpu_preprocess_baseline():
    from fast.data import filters as F
    from oracle import filter_oracle as FO
    nc = 2
    rng = np.random.default_rng(1000)
    xc = (20 * rng.standard_normal((nc, 64, 795)) + rng.uniform(-50, 50, (1, 64, 1))).astype(np.float32).reshape(-1, 795)
    sos = F.bandpass_sos()
    t0 = time.perf_counter()
    yc = FO.sosfiltfilt(sos, xc)
    FO.resample_poly(yc, 4, 5)
    dt = time.perf_counter() - t0
    return {"value": nc / dt, "unit": "trials/s", "cores": 1, "kind": "port",
            "sample": f"{nc} trials x 64 channels ({dt:.1f} s), NumPy float64 oracle port of scipy sosfiltfilt + resample_poly"}


# ------------------------------------------------------------------------------------------------ leg: cross-validation
def measure_cv(args, ctx):
    """BASELINE configs[3]: 15 subjects x 5 folds of FAST training, jobs sharded over ranks (strong scaling, no data-path
    collective), synthetic subjects.  value = seconds from first dispatch to the last result file."""
    import shutil
    import tempfile
    from fast.data import SUBJECTS
    from fast.models import fast as FM
    from fast.train import cv as CV
    FM.set_head_precision(args.head_precision)
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    cfg = prod_cfg()
    subjects = SUBJECTS[:args.subjects]
    jobs = CV.cv_jobs(subjects, 5)
    J = max(1, args.cv_concurrent)
    mine = CV.shard_jobs_contiguous(jobs, rank, world)      # subject-major blocks: a rank loads ~15 / N subjects, not one per job
    save = args.cv_dir or os.path.join(tempfile.gettempdir(), "fastb200_cv")
    if rank == 0:
        shutil.rmtree(save, ignore_errors=True)
        os.makedirs(save, exist_ok=True)
    ctx.barrier()                                       # nobody writes into the tree before rank 0 has cleaned it
    # The rank's subjects are generated BEFORE the clock starts and handed over as host arrays (the reference's jobs read
    # preprocessed files, scripts/train_fast.py:100-110; drawing 18 M normal deviates per subject is the bench's stand-in for
    # that and took 0.2 s each); the host -> device upload of every subject stays inside the timed region.
    data = {sid: CV.synthetic_subject(int(sid) - 1)[0] for sid in sorted({j[1] for j in mine})}

    def load(sid):
        return data[sid]

    # warm-up: one short group of jobs (builds graphs, primes the allocators; not part of the 75 jobs)
    CV.run_cv_jobs(cfg, [(0, "warm", k) for k in range(min(J, 5))], lambda s: CV.synthetic_subject(99)[0],
                   os.path.join(save, f"_warm{rank}"), max_epochs=2, batch_size=args.cv_batch, seed=1, concurrent=J)
    ctx.barrier()
    with ClockSampler(ctx.local_rank) as clocks:
        l0 = launches()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        CV.run_cv_jobs(cfg, mine, load, save, max_epochs=args.epochs, batch_size=args.cv_batch, seed=42, resume=False, concurrent=J)
        e1.record()
        torch.cuda.synchronize(dev)
        dev_s = e0.elapsed_time(e1) * 1e-3
        n_launch = launches() - l0
        ctx.barrier()
        wall = time.perf_counter() - t0
    dev_s = ctx.max_over_ranks(dev_s)
    wall = ctx.max_over_ranks(wall)
    rows = []
    if rank == 0:
        tests = {sid: CV.synthetic_subject(int(sid) - 1)[1] for sid in subjects}
        rows = CV.finalize_cv(cfg, subjects, save, n_folds=5, test_per_subject=tests)
    n_train_trials = len(jobs) * args.epochs * 280
    steps = len(jobs) * args.epochs * 5
    tp = load_peaks().get("bf16_tflops", 1590.0)
    ach = FLOP_TRAIN_DIRECT * n_train_trials / dev_s / 1e12 / world
    return {"metric": "15-subject x 5-fold CV wall-clock", "value": dev_s, "unit": "s", "n_gpus": world, "steps": steps, "warmup": 1,
            "ms_per_step": 1e3 * dev_s / max(1, len(mine) * args.epochs * 5), "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": _train_dtype(FM), "data": "synthetic", "clocks": clocks.summary(),
            "config": {"workload": "BASELINE configs[3]: per-subject 5-fold CV of FAST, batch 64", "jobs": len(jobs), "jobs_this_rank": len(mine),
                       "epochs": args.epochs, "subjects": len(subjects), "trials_per_subject": 350, "batch": args.cv_batch,
                       "concurrent_jobs_per_gpu": J,
                       "parallelism": f"subject x fold jobs sharded x{world}, no collective", "host_wall_s": wall,
                       "ideal_rounds": -(-len(jobs) // world)},
            "train_trials_per_s": n_train_trials / dev_s, "subjects_done": len(rows),
            "mean_test_acc": (float(np.mean([r[2] for r in rows])) if rows else None),
            "e2e": {"value": wall, "unit": "s", "h2d_bytes_per_step": 350 * 64 * 800 * 4 // (5 * args.epochs * 5), "d2h_bytes_per_step": 0,
                    "note": "host wall-clock incl. subject upload, result files and the join barrier"},
            "gpu_launches": n_launch,
            "roofline": {"kernel": "whole job stream (batch-64 steps: latency / occupancy bound, see extra.train_b64)", "bound": "tensor",
                         "achieved": ach, "peak": tp, "unit": "TFLOP/s", "frac": ach / tp, "traffic": None,
                         "note": "direct-form training flops of the 4.2 M trial fwd+bwd (validation passes not counted) / wall, per GPU"},
            "cpu_baseline": None}


# ------------------------------------------------------------------------------------------------ leg: stress (data parallel)
def measure_stress(args, ctx, reduce, B):
    """BASELINE configs[4]: 64 ch x 4000-sample trials (31 windows, 32 tokens), data-parallel training with the flat
    gradient exchanged every step.  Weak scaling: B trials per GPU (512 x 8 = 4096 at 8 GPUs)."""
    ctx.ensure_dist()
    from fast.models import FAST
    from fast.models import fast as FM
    from fast.train import DataParallelTrainer
    FM.set_head_precision(args.head_precision)
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    cfg = prod_cfg()
    cfg.seq_len = 4000
    K, W = args.steps, max(args.warmup, 3)
    torch.manual_seed(42)
    model = FAST(cfg).to(dev).train()
    ft = DataParallelTrainer(model, reduce=reduce, lr=5e-4, use_graph=True)
    gen = torch.Generator(device=dev).manual_seed(rank)
    x = torch.randn(B, 64, 4000, device=dev, generator=gen)
    y = torch.randint(0, 5, (B,), device=dev, generator=gen).to(torch.uint8)
    for _ in range(W):
        ft.train_step(x, y)
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(ctx.local_rank) as clocks:
        l0 = launches()
        e0.record()
        for _ in range(K):
            ft.train_step(x, y)
        e1.record()
        n_launch = launches() - l0
        ctx.barrier()
    ms = ctx.max_over_ranks(e0.elapsed_time(e1))
    # exposed exchange time: the same K steps of the update phase alone (barrier + fused reduce/AdamW, or NCCL + AdamW)
    ws = next(w for k, w in ft._ws.items() if isinstance(w, dict) and w.get("graph"))
    upd = [fn for fn in ws["graph"][1:]] if world > 1 else []
    ctx.barrier()
    e0.record()
    for _ in range(K):
        for fn in upd:
            fn()
    e1.record()
    ctx.barrier()
    upd_ms = ctx.max_over_ranks(e0.elapsed_time(e1) / K) if upd else 0.0
    loss, _ = ft.global_loss_and_correct()
    flop = 2.0 * (31 * (38_793_216 + 75_067_392) + 3 * 1_564_832) * 1.0
    tensor_peak = load_peaks().get("bf16_tflops", 1590.0)
    del x, ft, model
    return {"metric": "FAST stress training trials/s (64 ch x 4000 samples, data-parallel)", "value": world * B * K / (ms * 1e-3),
            "unit": "trials/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": _train_dtype(FM), "data": "synthetic", "clocks": clocks.summary(),
            "config": {"workload": "BASELINE configs[4]: seq_len 4000 (31 windows, 32 tokens), DP", "batch_per_gpu": B,
                       "global_batch": B * world, "reduce": reduce if world > 1 else "none",
                       "l2_policy": f"input {B * 64 * 4000 * 4 / 1e9:.2f} GB per step > 126 MB L2",
                       "parallelism": f"dp{world}: one flat 767 KB gradient exchange per step"},
            "exchange_plus_update_ms": upd_ms, "exchange_share_of_step": (upd_ms / (ms / K) if upd else 0.0), "loss": loss,
            "e2e": None, "gpu_launches": n_launch,
            "roofline": {"kernel": "whole step (tcgen05 head kernels)", "bound": "tensor", "achieved": flop * B * K / (ms * 1e-3) / 1e12,
                         "peak": tensor_peak, "unit": "TFLOP/s", "frac": flop * B * K / (ms * 1e-3) / 1e12 / tensor_peak, "traffic": None,
                         "note": "direct-form training flops per trial / step time, per GPU, against the measured burst bf16 peak; the "
                                 "exchange moves (world-1) x 767 KB per rank per step over NVLink: microseconds of a 770 GB/s link against "
                                 "a ~20 ms step, so the tensor pipe, not the link, is the bound"},
            "cpu_baseline": None}


# ------------------------------------------------------------------------------------------------ same-box library baseline
def measure_eager(ctx, quick=False):
    """The reference model through STOCK torch ops on this B200 (cuDNN convolutions, cuBLAS GEMMs, ATen element-wise = what
    the unmodified reference runs on a GPU; the functional oracle port is that same op sequence, and /root/reference does
    not exist on the box).  Three arithmetic settings: strict fp32 (TF32 off); the reference's
    ``seed_all`` globals (float32_matmul_precision('medium') + cudnn.deterministic, src/fast/utils.py:112-114); and
    ``torch.autocast(bfloat16)`` (scripts/train_fast.py:277 'bf16-mixed').  Baseline only: outside every timed region."""
    from fast.models import FAST
    from oracle import fast_oracle as O
    dev = ctx.dev
    cfg = prod_cfg(dropout=0.0)
    torch.manual_seed(42)
    sd = {k: v.detach().to(dev) for k, v in FAST(cfg).state_dict().items()}
    g = torch.Generator(device=dev).manual_seed(0)
    x_all = torch.randn(8192, 64, 795, device=dev, generator=g)
    y_all = torch.randint(0, 5, (8192,), device=dev, generator=g)
    saved = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.deterministic,
             torch.get_float32_matmul_precision(), torch.backends.cudnn.benchmark)

    def timed(fn, budget_s=0.6):
        fn(); fn()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize(dev)
        one = max(e0.elapsed_time(e1) * 1e-3, 1e-5)
        n = int(min(50, max(2, budget_s / one)))
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) * 1e-3 / n

    def setting(name):
        torch.backends.cudnn.benchmark = False
        if name == "fp32_strict":
            torch.backends.cudnn.allow_tf32 = False
            torch.backends.cuda.matmul.allow_tf32 = False
            torch.backends.cudnn.deterministic = False
            torch.set_float32_matmul_precision("highest")
        else:                                         # what seed_all leaves behind; autocast runs on top of it in the reference
            torch.backends.cudnn.allow_tf32 = True
            torch.set_float32_matmul_precision("medium")
            torch.backends.cudnn.deterministic = True

    out = {"infer_trials_per_s": {}, "train_trials_per_s": {},
           "note": "functional oracle port of the reference forward / training step (CE + autograd + torch.optim.AdamW) on CUDA "
                   "tensors: stock cuDNN / cuBLAS / ATen kernels, eager, no CUDA graph -- the path the reference takes on a GPU"}
    try:
        for name in ("fp32_strict", "seed_all_medium_tf32", "bf16_autocast"):
            setting(name)
            amp = name == "bf16_autocast"
            inf, trn = {}, {}
            for b in ((64, 1024) if quick else (64, 1024, 8192)):
                xb = x_all[:b]

                def f_inf():
                    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                        for c in xb.split(2048):                       # batched_forward_head-style cap on the cnn1 intermediate
                            O.forward(sd, cfg, c).argmax(1)
                inf[str(b)] = b / timed(f_inf)
            for b in (64, 1024):
                xb, yb = x_all[:b], y_all[:b]
                leaves = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
                opt = torch.optim.AdamW(list(leaves.values()), lr=5e-4)

                def f_trn():
                    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                        logits = O.forward(leaves, cfg, xb)
                        loss = torch.nn.functional.cross_entropy(logits.float(), yb)
                    loss.backward()
                    opt.step()
                    opt.zero_grad(set_to_none=True)
                trn[str(b)] = b / timed(f_trn)
                del leaves, opt
            out["infer_trials_per_s"][name] = inf
            out["train_trials_per_s"][name] = trn
    finally:
        (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.deterministic) = saved[:3]
        torch.set_float32_matmul_precision(saved[3])
        torch.backends.cudnn.benchmark = saved[4]
    del x_all
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------ drivers
def run_all(args, ctx):
    """Headline inference line + every other leg of BASELINE.json's metric nested under ``extra`` (see the module docstring)."""
    rank, world = ctx.rank, ctx.world
    line, model, x = measure_infer(args, ctx)
    extra = line["extra"]
    if not args.no_extras:
        extra.update(infer_sweeps(model, x, ctx.dev))
    del model, x
    torch.cuda.empty_cache()
    if not args.no_extras:
        extra["train_b1024"] = measure_train(args, ctx, 1024)
        extra["train_b64"] = measure_train(args, ctx, 64)
        extra["train_trials_per_s"] = {"64": extra["train_b64"]["value"], "1024": extra["train_b1024"]["value"]}
        torch.cuda.empty_cache()
        extra["preprocess"] = measure_preprocess(args, ctx)
        torch.cuda.empty_cache()
        extra["cv"] = measure_cv(args, ctx)
        torch.cuda.empty_cache()
        if world > 1:
            from fast.train.dp import dp_selfcheck
            extra["dp_check"] = [dp_selfcheck(prod_cfg(dropout=0.0), reduce=r) for r in ("p2p", "nccl")]
            extra["stress_p2p"] = measure_stress(args, ctx, "p2p", args.stress_batch)
            extra["stress_nccl"] = measure_stress(args, ctx, "nccl", args.stress_batch)
        else:
            extra["stress_p2p"] = measure_stress(args, ctx, "p2p", args.stress_batch)      # world 1: no exchange, the plain fused step
        torch.cuda.empty_cache()
        ctx.barrier()
        if rank == 0:
            e = extra["eager_port_gpu"] = measure_eager(ctx)
            extra["speedup_vs_eager_port_gpu"] = {
                "infer_b8192": {k: line["value"] / world / v["8192"] for k, v in e["infer_trials_per_s"].items()},
                "train_b1024": {k: extra["train_b1024"]["value"] / world / v["1024"] for k, v in e["train_trials_per_s"].items()},
                "train_b64": {k: extra["train_b64"]["value"] / world / v["64"] for k, v in e["train_trials_per_s"].items()}}
        ctx.barrier()
    ctx.close()
    if rank != 0:
        return
    # CPU baselines last: the other ranks have exited, the host cores are free
    threads = os.cpu_count() or 1
    cpu_v, cpu_s = cpu_forward_trials_per_s(args.cpu_trials, 256, 1, threads)
    line["cpu_baseline"] = {"value": cpu_v, "unit": "trials/s", "cores": threads, "kind": "port",
                            "sample": f"{args.cpu_trials} of the {args.batch} trials of one batch, in chunks of 256 ({cpu_s:.1f} s), oracle port of the reference forward"}
    if not args.no_extras:
        cpu_t = cpu_train_trials_per_s(64, threads)
        extra["train_b64"]["cpu_baseline"] = cpu_t
        extra["train_b1024"]["cpu_baseline"] = dict(cpu_t, sample=cpu_t["sample"] + " (batch 64: a 1024-trial CPU step is minutes; the rate is per trial)")
        extra["preprocess"]["cpu_baseline"] = cpu_preprocess_baseline()
        n_tr, n_va = 75 * args.epochs * 280, 75 * args.epochs * 70
        est = n_tr / cpu_t["value"] + n_va / cpu_v
        extra["cv"]["cpu_baseline"] = {"value": est, "unit": "s", "cores": threads, "kind": "port",
                                       "sample": f"extrapolated, not run: {n_tr} training trials at the measured CPU training rate + {n_va} "
                                                 "validation forwards at the measured CPU inference rate (the oracle port of the 75 jobs is hours)"}
        for k in ("stress_p2p", "stress_nccl"):
            if k in extra:
                extra[k]["cpu_baseline"] = dict(cpu_t, sample=cpu_t["sample"] + " (production seq_len 800; a 4000-sample trial costs ~6.2x per trial)")
    emit(line)


def run_single(args, ctx):
    w = args.workload
    line = None
    if w == "infer":
        line, model, x = measure_infer(args, ctx)
        if not args.no_extras:
            line["extra"].update(infer_sweeps(model, x, ctx.dev))
    elif w == "train":
        line = measure_train(args, ctx, args.batch)
    elif w == "preprocess":
        line = measure_preprocess(args, ctx)
    elif w == "cv":
        line = measure_cv(args, ctx)
    elif w == "stress":
        line = measure_stress(args, ctx, args.reduce, args.batch)
    elif w == "eager":
        if ctx.rank == 0:
            line = {"metric": "eager_port_gpu", "extra": measure_eager(ctx, quick=args.no_extras)}
        ctx.barrier()
    ctx.close()
    if ctx.rank != 0:
        return
    if not args.no_extras:
        threads = os.cpu_count() or 1
        if w == "infer":
            cpu_v, cpu_s = cpu_forward_trials_per_s(args.cpu_trials, 256, 1, threads)
            line["cpu_baseline"] = {"value": cpu_v, "unit": "trials/s", "cores": threads, "kind": "port",
                                    "sample": f"{args.cpu_trials} trials in chunks of 256 ({cpu_s:.1f} s), oracle port of the reference forward"}
        elif w in ("train", "stress"):
            line["cpu_baseline"] = cpu_train_trials_per_s(64, threads)
        elif w == "preprocess":
            line["cpu_baseline"] = cpu_preprocess_baseline()
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--no-extras", action="store_true", help="headline leg only (no nested legs, no sweeps)")
    ap.add_argument("--head-precision", default=os.environ.get("FASTB200_HEAD_PRECISION", "bf16x2"), choices=["bf16x3", "bf16x2", "fp32"])
    ap.add_argument("--e2e-chunk", type=int, default=1024, help="trials per H2D piece of the end-to-end run")
    ap.add_argument("--cpu-trials", type=int, default=4096, help="trials of the bounded CPU-baseline sample (~10 s)")
    ap.add_argument("--ref-trials", type=int, default=256, help="--impl reference: trials of the bounded sample per step")
    ap.add_argument("--workload", default="all", choices=["all", "infer", "train", "preprocess", "cv", "stress", "eager"],
                    help="all = the headline (BASELINE configs[1]) with every other leg nested under extra; else one leg alone")
    ap.add_argument("--epochs", type=int, default=200, help="cv: epochs per fold (reference: 200)")
    ap.add_argument("--subjects", type=int, default=15, help="cv: number of synthetic subjects")
    ap.add_argument("--cv-batch", type=int, default=64)
    ap.add_argument("--cv-concurrent", type=int, default=5, help="cv: folds trained concurrently per GPU (one CUDA graph per epoch over all of them)")
    ap.add_argument("--cv-dir", default=None)
    ap.add_argument("--stress-batch", type=int, default=512, help="stress leg: trials per GPU")
    ap.add_argument("--reduce", default="p2p", choices=["p2p", "nccl"], help="--workload stress: gradient exchange")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    capture_stdout()
    ctx = Ctx()
    if args.workload == "all":
        run_all(args, ctx)
    else:
        if "--batch" not in sys.argv:
            args.batch = {"train": 1024, "stress": args.stress_batch}.get(args.workload, args.batch)
        run_single(args, ctx)


if __name__ == "__main__":
    main()
