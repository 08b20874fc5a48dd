"""Per-subject k-fold cross-validation of FAST, sharded over GPUs as independent jobs
(the throughput path behind ``finetune_per_subject_cv``, reference scripts/train_fast.py:68-262).

The 15 subjects x 5 folds are 75 jobs that share nothing (each builds a fresh model, train_fast.py:113), so ranks take
jobs and there is NO data-path collective; the only join is on the host: per subject "best of the folds -> that
subject's test set" (train_fast.py:182-212) and the summary CSVs (:226-240).

One job = the reference's ``trainer.fit`` for one fold, with the Lightning loop replaced by a device-resident loop:
the subject's trials stay in HBM (72 MB), batches are index gathers on the device, and a WHOLE EPOCH -- the batch
gathers, the fused training steps, the validation pass, the metric row and the predicated best-``val_acc`` snapshot --
is one CUDA graph replayed once per epoch (the host only draws the next permutation), so the host neither synchronises
inside a job nor limits it when several ranks share one host.  One batch-64 job cannot fill a B200, so a process runs
several jobs concurrently (``run_cv_jobs(concurrent=J)`` / ``train_folds``): the epoch graph then has J parallel branches.  Artefacts keep the reference's names and formats
(``sub-XX/fold-k_history.csv``, ``fold_metrics.csv``, ``best_subject.pth``, ``test_predictions.csv``,
``summary_per_subject.csv``, ``global_test_predictions.csv``) so ``scripts/benchmark.py`` reads them unchanged.

Documented deviation: the reference seeds once before all loops, so job k's initial weights and
shuffles depend on every job before it; sharded jobs cannot reproduce that stream and use per-job seeds
``seed + n_folds * subject_index + fold`` instead.
"""
from __future__ import annotations

import csv
import json
import math
import os

import numpy as np
import torch

from .. import _native as N
from .fused import FusedTrainer, lightning_lr_factors
from .trainer import cosine_scheduler


# ---------------------------------------------------------------------------------------------------- host logic
def kfold_indices(n_samples: int, n_splits: int, shuffle: bool = True, random_state=None):
    """[(train_idx, val_idx)] exactly as ``sklearn.model_selection.KFold(n_splits, shuffle, random_state).split``
    (used at train_fast.py:94): a ``RandomState(seed).shuffle`` of ``arange(n)``, then contiguous folds whose first
    ``n % k`` members have one extra sample; the train indices keep ascending order."""
    if n_splits < 2 or n_splits > n_samples:
        raise ValueError(f"n_splits={n_splits} must be in [2, n_samples={n_samples}]")
    idx = np.arange(n_samples)
    if shuffle:
        np.random.RandomState(random_state).shuffle(idx)
    sizes = np.full(n_splits, n_samples // n_splits, dtype=int)
    sizes[:n_samples % n_splits] += 1
    out, cur = [], 0
    for s in sizes:
        mask = np.zeros(n_samples, dtype=bool)
        mask[idx[cur:cur + s]] = True
        out.append((np.flatnonzero(~mask), np.flatnonzero(mask)))
        cur += s
    return out


def cv_jobs(subjects, n_folds):
    """All (subject_index, subject_id, fold) jobs in the reference's loop order (train_fast.py:86,100)."""
    return [(si, sid, k) for si, sid in enumerate(subjects) for k in range(n_folds)]


def shard_jobs(jobs, rank: int, world: int):
    """Static interleaved partition: rank r takes jobs r, r + world, ...  Every job has the same cost (same trial count,
    epochs and model), so this is balanced to within one job: ceil(75 / g) rounds at g GPUs."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world {world}")
    return list(jobs[rank::world])


def shard_jobs_contiguous(jobs, rank: int, world: int):
    """Balanced CONTIGUOUS partition of the subject-major job list: rank r takes jobs [r n / g, (r + 1) n / g), sizes differing
    by at most one (the same ceil(75 / g) rounds as ``shard_jobs``).  A rank then touches ~n_subjects / g subjects instead of
    min(n_jobs / g, n_subjects): each subject is read from disk and uploaded by one or two ranks only, and the concurrent
    groups of ``run_cv_jobs(concurrent=J)`` are the folds of ONE subject sharing its resident trials."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world {world}")
    n = len(jobs)
    lo, hi = (rank * n) // world, ((rank + 1) * n) // world
    return list(jobs[lo:hi])


class JobQueue:
    """Dynamic alternative to ``shard_jobs``: an atomic counter in the torch.distributed key-value store hands out job
    indices, so a rank that finishes early takes the next job (no collective, works before/without NCCL)."""

    def __init__(self, jobs, store=None, key="fastb200/cv/next"):
        self.jobs, self.store, self.key, self._local = list(jobs), store, key, 0

    def __iter__(self):
        return self

    def __next__(self):
        i = (self.store.add(self.key, 1) - 1) if self.store is not None else self._local
        self._local += 1
        if i >= len(self.jobs):
            raise StopIteration
        return self.jobs[i]


def job_seed(seed: int, n_folds: int, subject_index: int, fold: int) -> int:
    return int(seed) + n_folds * int(subject_index) + int(fold)


def pick_best_fold(fold_best):
    """Index of the fold whose checkpoint the reference evaluates: strict ``>`` scan in fold order
    (train_fast.py:182-185), so ties go to the earliest fold."""
    best, best_k = -np.inf, None
    for k, v in fold_best:
        if v > best:
            best, best_k = v, k
    return best_k, best


def macro_f1(y_true, y_pred, n_classes=None):
    """``sklearn.metrics.f1_score(average='macro')`` over the labels present in y_true or y_pred (train_fast.py:214)."""
    y_true, y_pred = np.asarray(y_true).astype(int), np.asarray(y_pred).astype(int)
    labels = np.union1d(y_true, y_pred) if n_classes is None else np.arange(n_classes)
    f = []
    for c in labels:
        tp = np.sum((y_true == c) & (y_pred == c))
        d = np.sum(y_true == c) + np.sum(y_pred == c)
        f.append(2.0 * tp / d if d else 0.0)
    return float(np.mean(f)) if f else 0.0


# ---------------------------------------------------------------------------------------------------- folds on the device
class _FoldJob:
    """Device-resident state of one fold: model, fused trainer, metric accumulators, best-epoch snapshot.  ``epoch_body``
    is one whole epoch as stream-ordered launches only (no host reads, fixed buffers), so it can be captured -- alone or
    side by side with other folds' bodies -- into a CUDA graph."""

    def __init__(self, config, x, y, train_idx, val_idx, max_epochs, batch_size, seed, lr, warmup_epochs, model=None):
        from ..models import FAST
        dev = x.device
        if not x.is_cuda:
            raise RuntimeError("train_fold keeps the subject's trials resident on a CUDA device (there is no CPU fallback)")
        torch.manual_seed(seed)
        if model is None:
            model = FAST(config)
        self.model = model.to(dev).train()
        self.dev, self.x = dev, x
        self.tr = torch.as_tensor(np.asarray(train_idx), device=dev, dtype=torch.int64)
        va = torch.as_tensor(np.asarray(val_idx), device=dev, dtype=torch.int64)
        self.n_tr, self.n_va = self.tr.numel(), va.numel()
        self.n_it = math.ceil(self.n_tr / batch_size)             # len(train_loader)
        if warmup_epochs > max_epochs:                            # the stock warmup_epochs=10 asserts for short runs
            warmup_epochs = 0
        sched = cosine_scheduler(1, 0.1, max_epochs, self.n_it, warmup_epochs=warmup_epochs)
        # the graph is per EPOCH (below), so the trainer itself launches eagerly
        self.ft = FusedTrainer(self.model, lr=lr, lr_factors=lightning_lr_factors(sched), use_graph=False, seed=seed)
        self.gen = torch.Generator(device=dev).manual_seed(seed)
        if y.dtype not in (torch.uint8, torch.int64):
            y = y.long()
        self.y = y
        self.xv, self.yv = x[va].contiguous(), y[va].contiguous()
        self.hist = torch.zeros(max_epochs, 4, device=dev, dtype=torch.float64)
        self.best_correct = torch.full((1,), -1, device=dev, dtype=torch.int32)
        self.best_epoch = torch.zeros(1, device=dev, dtype=torch.int32)
        self.best_flat = self.ft.flat.clone()
        self.loss_sum = torch.zeros(1, device=dev, dtype=torch.float64)
        self.corr_sum = torch.zeros(1, device=dev, dtype=torch.int64)
        self.perm = torch.empty(self.n_tr, device=dev, dtype=torch.int64)   # this epoch's shuffled trial indices (fixed address: graph input)
        self.ep_dev = torch.zeros(1, device=dev, dtype=torch.int64)         # epoch counter, advanced inside the epoch graph
        self.T = x.shape[2]
        self.sizes = [min(batch_size, self.n_tr - lo) for lo in range(0, self.n_tr, batch_size)]

    def draw_perm(self):
        self.perm.copy_(self.tr[torch.randperm(self.n_tr, device=self.dev, generator=self.gen)])

    def epoch_body(self):
        """Batch gathers, the fused training steps, the validation pass, the metric row and the best-epoch snapshot."""
        ft, x, y = self.ft, self.x, self.y
        self.loss_sum.zero_(); self.corr_sum.zero_()
        lo = 0
        for b in self.sizes:
            ws = ft._workspace(b, self.T, y.dtype, True)
            idx = self.perm[lo:lo + b]
            torch.index_select(x, 0, idx, out=ws["x"])
            torch.index_select(y, 0, idx, out=ws["y"])
            ft._launch_train(ws)
            self.loss_sum.add_(ft.loss.double() * b)              # Lightning's on_epoch mean is batch-size weighted
            self.corr_sum.add_(ft.n_correct)
            lo += b
        wsv = ft._workspace(self.n_va, self.T, y.dtype, False)
        wsv["x"].copy_(self.xv); wsv["y"].copy_(self.yv)
        ft._launch_eval(wsv)
        vcorr = ft.n_correct
        row = torch.cat([self.loss_sum / self.n_tr, self.corr_sum.double() / self.n_tr, ft.loss.double(),
                         vcorr.double() / self.n_va]).unsqueeze(0)
        self.hist.index_copy_(0, self.ep_dev, row)
        better = vcorr > self.best_correct                        # (1,) bool, stays on the device
        self.best_flat.copy_(torch.where(better, ft.flat, self.best_flat))
        self.best_epoch.copy_(torch.where(better, self.ep_dev.to(torch.int32), self.best_epoch))
        self.best_correct.copy_(torch.maximum(self.best_correct, vcorr))
        self.ep_dev.add_(1)

    def finish(self):
        """Host values of the finished fold (call after a device synchronisation point: ``hist.cpu()`` is one)."""
        ft, model = self.ft, self.model
        h = self.hist.cpu().numpy()
        with torch.no_grad():
            final = ft.flat.clone()
            ft.flat.copy_(self.best_flat)
            # one device-to-host copy of the flat buffer; the state_dict entries are views into it (94 separate .cpu() calls cost
            # 94 synchronisations)
            flat_cpu, base, esz = ft.flat.cpu(), ft.flat.data_ptr(), ft.flat.element_size()
            best_sd = {}
            for k, v in model.state_dict().items():
                off = (v.data_ptr() - base) // esz
                if v.dtype == ft.flat.dtype and v.is_contiguous() and 0 <= off and off + v.numel() <= flat_cpu.numel():
                    best_sd[k] = flat_cpu[off:off + v.numel()].view(v.shape).clone()
                else:
                    best_sd[k] = v.detach().cpu().clone()
            ft.flat.copy_(final)
        return {"history": {"loss": h[:, 0].tolist(), "acc": h[:, 1].tolist(), "val_loss": h[:, 2].tolist(),
                            "val_acc": h[:, 3].tolist()},
                "best_val_acc": float(self.best_correct.item()) / self.n_va, "best_epoch": int(self.best_epoch.item()),
                "train_len": self.n_it, "steps": ft.steps_done, "state_dict": best_sd}


def train_folds(config, folds, max_epochs=200, batch_size=64, lr=5e-4, warmup_epochs=10, use_graph=True):
    """``trainer.fit`` of several folds AT ONCE on one GPU (train_fast.py:100-140 without Lightning).

    ``folds``: list of dicts ``x, y, train_idx, val_idx, seed`` (CUDA tensors holding a subject's train+val trials, the
    fold's index arrays, the job seed); folds of one subject share ``x``.  A batch-64 step cannot fill a B200 (64 of 148 SMs
    in the token kernels, mostly launch / drain latency in the head kernels), so the folds run CONCURRENTLY: every epoch is
    ONE CUDA graph whose parallel branches -- one capture stream per fold -- are the folds' epoch bodies; the host only
    draws the next permutations and replays.  Each fold's arithmetic is exactly that of running it alone (own model, own
    buffers, own dropout / shuffle streams).  Returns one result dict per fold (see ``train_fold``)."""
    if not folds:
        return []
    jobs = [_FoldJob(config, f["x"], f["y"], f["train_idx"], f["val_idx"], max_epochs, batch_size, f["seed"], lr, warmup_epochs,
                     model=f.get("model")) for f in folds]
    dev = jobs[0].dev
    if any(j.dev != dev for j in jobs):
        raise RuntimeError("concurrent folds must live on one device")
    graph, n_launch = None, 0
    lib = jobs[0].ft.lib
    sides = [torch.cuda.Stream(dev) for _ in jobs] if len(jobs) > 1 else []
    for ep in range(max_epochs):
        for j in jobs:
            if j.ft.lr_by_step and j.ft.steps_done + j.n_it > j.ft.n_sched:
                raise RuntimeError("LR schedule shorter than the run")
            j.draw_perm()
        if not use_graph or ep == 0:
            for j in jobs:
                j.epoch_body()                                # epoch 0 runs eagerly (kernel attributes, allocator warm-up)
        else:
            if graph is None:
                graph = torch.cuda.CUDAGraph()
                cap = torch.cuda.Stream(dev)
                cap.wait_stream(torch.cuda.current_stream(dev))
                c0 = lib.fastb200_launch_count()
                # capture_begin / capture_end directly: the torch.cuda.graph context also runs gc.collect() and empties the
                # device and pinned-host caches on entry, tens of ms per job that 75 jobs pay 75 times
                with torch.cuda.stream(cap):
                    graph.capture_begin()
                    try:
                        if len(jobs) == 1:
                            jobs[0].epoch_body()
                        else:
                            for j, side in zip(jobs, sides):      # fork: one branch per fold
                                side.wait_stream(cap)
                                with torch.cuda.stream(side):
                                    j.epoch_body()
                            for side in sides:                    # join
                                cap.wait_stream(side)
                    finally:
                        graph.capture_end()
                torch.cuda.current_stream(dev).wait_stream(cap)
                n_launch = lib.fastb200_launch_count() - c0
            graph.replay()
            N.note_replayed(n_launch)
        for j in jobs:
            j.ft.steps_done += j.n_it
    return [j.finish() for j in jobs]


def train_fold(config, x, y, train_idx, val_idx, max_epochs=200, batch_size=64, seed=42, lr=5e-4, warmup_epochs=10,
               use_graph=True, model=None):
    """``trainer.fit`` of one fold (train_fast.py:100-140) without Lightning.

    x (n, C, T) float32 and y (n,) uint8/int64 are CUDA tensors holding the subject's train+val trials; train_idx /
    val_idx index them.  Returns a dict of host values read ONCE at the end of the job:
    ``history`` {'loss','acc','val_loss','val_acc'} per epoch (the HistoryCallback keys, callbacks.py:12),
    ``best_val_acc``, ``best_epoch`` and ``state_dict`` of the best epoch (ModelCheckpoint monitor='val_acc' mode='max'
    top-1: replaced only on a strictly better epoch, train_fast.py:118-125)."""
    return train_folds(config, [dict(x=x, y=y, train_idx=train_idx, val_idx=val_idx, seed=seed, model=model)], max_epochs,
                       batch_size, lr, warmup_epochs, use_graph)[0]


# ---------------------------------------------------------------------------------------------------- the sharded run
def _write_history(path, history):
    keys = list(history.keys())
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["Epoch"] + keys)                          # pandas to_csv(index_label='Epoch'), train_fast.py:150
        for i in range(min(len(v) for v in history.values())):
            w.writerow([i] + [history[k][i] for k in keys])


def run_cv_jobs(config, jobs, load_subject, save_dir, max_epochs=200, batch_size=64, seed=42, n_folds=5, device=None,
                resume=True, concurrent=1, **fold_kw):
    """Run ``jobs`` (an iterable of (subject_index, sid, fold): ``shard_jobs(...)`` or a ``JobQueue``) on this process's
    GPU, ``concurrent`` of them at a time (``train_folds``: one CUDA graph per epoch over all of them).
    ``load_subject(sid) -> (X (n,C,T) float32, Y (n,))`` NumPy arrays or tensors (CUDA tensors are used in place:
    ``fast.data.subject_train_val_device``); each subject is uploaded once and stays resident while its folds run.  Every
    finished job leaves ``sub-SID/fold-k-best.pth`` + ``fold-k_history.csv`` + ``fold-k_result.json``; with ``resume`` a job
    whose result file exists is skipped, so a crashed worker's jobs can simply be re-queued.  Returns the result dicts."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    concurrent = max(1, int(concurrent))
    cache, done = {}, []
    # a static shard knows its next subjects: load them on a worker thread while the current group trains (the loaders
    # release the GIL in NumPy / file I/O); a JobQueue hands jobs out one by one, so nothing to look ahead to
    pending, pool, todo = {}, None, None
    if isinstance(jobs, (list, tuple)):
        from concurrent.futures import ThreadPoolExecutor
        todo = []
        for _, sid, k in jobs:
            if not (resume and os.path.exists(os.path.join(save_dir, f"sub-{sid}", f"fold-{k}_result.json"))) and sid not in todo:
                todo.append(sid)                                   # subjects in the order they are first needed
        pool = ThreadPoolExecutor(max_workers=1)

    def subject(sid):
        if sid not in cache:
            while len(cache) > concurrent:                         # resident subjects: this group's + one looked ahead
                cache.pop(next(iter(cache)))
            X, Y = pending.pop(sid).result() if sid in pending else load_subject(sid)
            cache[sid] = (torch.as_tensor(X).to(dev, non_blocking=True), torch.as_tensor(Y).to(dev, non_blocking=True))
        return cache[sid]

    def prefetch(after_sid):
        if pool is None or after_sid not in todo:
            return
        for nxt in todo[todo.index(after_sid) + 1:todo.index(after_sid) + 1 + concurrent]:
            if nxt not in pending and nxt not in cache:
                pending[nxt] = pool.submit(load_subject, nxt)
                return                                             # one look-ahead load at a time

    def run_group(group):
        folds = []
        for si, sid, k in group:
            X, Y = subject(sid)
            tr_idx, va_idx = kfold_indices(X.shape[0], n_folds, True, seed)[k]
            folds.append(dict(x=X, y=Y, train_idx=tr_idx, val_idx=va_idx, seed=job_seed(seed, n_folds, si, k)))
        prefetch(group[-1][1])
        outs = train_folds(config, folds, max_epochs, batch_size, **fold_kw)
        for (si, sid, k), out in zip(group, outs):
            sub_dir = os.path.join(save_dir, f"sub-{sid}")
            res_path = os.path.join(sub_dir, f"fold-{k}_result.json")
            torch.save(out.pop("state_dict"), os.path.join(sub_dir, f"fold-{k}-best.pth"))
            _write_history(os.path.join(sub_dir, f"fold-{k}_history.csv"), out.pop("history"))
            res = {"subject": sid, "subject_index": si, "fold": k, **out}
            tmp = res_path + f".tmp{os.getpid()}"
            with open(tmp, "w") as f:
                json.dump(res, f)
            os.replace(tmp, res_path)                              # the result file appears atomically = the job's commit
            done.append(res)

    group = []
    for si, sid, k in jobs:
        sub_dir = os.path.join(save_dir, f"sub-{sid}")
        os.makedirs(sub_dir, exist_ok=True)
        res_path = os.path.join(sub_dir, f"fold-{k}_result.json")
        if resume and os.path.exists(res_path):
            done.append(json.load(open(res_path)))
            continue
        group.append((si, sid, k))
        if len(group) == concurrent:
            run_group(group)
            group = []
    if group:
        run_group(group)
    if pool is not None:
        pool.shutdown(wait=True)
    return done


def finalize_cv(config, subjects, save_dir, n_folds=5, test_per_subject=None, device=None):
    """Host-side join after all ranks finished (train_fast.py:182-240): per subject pick the best fold, save
    ``best_subject.pth``, predict that subject's test set, write the per-subject and global CSVs.  Returns the
    summary rows ``[sid, best_val_acc, test_acc, test_f1]``."""
    from ..models import FAST
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    rows, gp, gt = [], [], []
    model = None
    for sid in subjects:
        sub_dir = os.path.join(save_dir, f"sub-{sid}")
        folds = []
        for k in range(n_folds):
            p = os.path.join(sub_dir, f"fold-{k}_result.json")
            if os.path.exists(p):
                folds.append((k, json.load(open(p))["best_val_acc"]))
        if not folds:
            continue                                           # "No checkpoint found ... skipping" (train_fast.py:199-201)
        best_k, best_acc = pick_best_fold(folds)
        sd = torch.load(os.path.join(sub_dir, f"fold-{best_k}-best.pth"), weights_only=True)
        torch.save(sd, os.path.join(sub_dir, "best_subject.pth"))
        test_acc = test_f1 = float("nan")
        if test_per_subject is not None and sid in test_per_subject:
            if model is None:
                model = FAST(config).to(dev)
            model.load_state_dict(sd)
            Xt, Yt = test_per_subject[sid]
            pred = model.predict(torch.as_tensor(Xt).to(dev)).cpu().numpy()
            true = np.asarray(torch.as_tensor(Yt).cpu()).astype(np.int64)
            test_acc = float(np.mean(pred == true))
            test_f1 = macro_f1(true, pred)
            np.savetxt(os.path.join(sub_dir, "test_predictions.csv"), np.array([pred, true]).T, delimiter=',', fmt='%d',
                       header='Predicted,True')
            gp.append(pred); gt.append(true)
        rows.append([sid, best_acc, test_acc, test_f1])
        with open(os.path.join(sub_dir, "fold_metrics.csv"), "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["Fold", "Best_Val_Acc"])
            w.writerows([[k, v] for k, v in folds])
    with open(os.path.join(save_dir, "summary_per_subject.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["Subject", "Best_Val_Acc", "Test_Acc", "Test_F1"])
        w.writerows(rows)
    if gp:
        np.savetxt(os.path.join(save_dir, "global_test_predictions.csv"), np.array([np.concatenate(gp), np.concatenate(gt)]).T,
                   delimiter=',', fmt='%d', header='Predicted,True')
    return rows


def synthetic_subject(subject_index: int, n_trials=350, n_test=50, n_chan=64, n_samples=800, n_classes=5):
    """SURVEY.md 8d-D4 synthetic subject: ``randn(350,64,800)`` train/val with balanced labels shuffled with seed
    2000+s, and ``randn(50,64,800)`` test.  Returns ((X, Y), (X_test, Y_test)) as NumPy float32 / uint8."""
    rng = np.random.default_rng(3000 + subject_index)
    X = rng.standard_normal((n_trials, n_chan, n_samples), dtype=np.float32)
    Y = np.repeat(np.arange(n_classes), math.ceil(n_trials / n_classes))[:n_trials].astype(np.uint8)
    np.random.default_rng(2000 + subject_index).shuffle(Y)
    Xt = rng.standard_normal((n_test, n_chan, n_samples), dtype=np.float32)
    Yt = np.repeat(np.arange(n_classes), math.ceil(n_test / n_classes))[:n_test].astype(np.uint8)
    return (X, Y), (Xt, Yt)
