"""Host logic of the multi-GPU paths, on CPU: job sharding of the cross-validation (no collective) and the
flat-gradient exchange of the data-parallel step, with world_size-2 ``gloo`` process groups."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, PKG, make_small_cfg, rel_err, small_state_dict


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, world, *args):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_entry, args=(fn, r, world, port, q, args)) for r in range(world)]
    for p in procs:
        p.start()
    out = {}
    try:
        for _ in range(world):                      # drain BEFORE join: a big payload blocks the child until it is read
            r, v = q.get(timeout=180)
            out[r] = v
    finally:
        for p in procs:
            p.join(30)
            if p.is_alive():
                p.kill()
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert len(out) == world
    return [out[r] for r in range(world)]


def _entry(fn, rank, world, port, q, args):
    import sys
    for p in (PKG, ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q.put((rank, fn(rank, world, *args)))
    finally:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ single process
def test_kfold_matches_sklearn():
    from sklearn.model_selection import KFold
    from fast.train.cv import kfold_indices
    for n, k, seed in ((350, 5, 42), (23, 4, 7), (10, 10, 0)):
        ours = kfold_indices(n, k, True, seed)
        ref = list(KFold(n_splits=k, shuffle=True, random_state=seed).split(np.zeros(n)))
        assert len(ours) == k
        for (a, b), (c, d) in zip(ours, ref):
            assert np.array_equal(a, c) and np.array_equal(b, d)
    tr, va = kfold_indices(350, 5, True, 42)[0]
    assert len(tr) == 280 and len(va) == 70
    with pytest.raises(ValueError):
        kfold_indices(3, 5)


def test_job_partition_and_selection_rules():
    from fast.data import SUBJECTS
    from fast.train.cv import cv_jobs, job_seed, macro_f1, pick_best_fold, shard_jobs, JobQueue
    from sklearn.metrics import f1_score
    jobs = cv_jobs(SUBJECTS, 5)
    assert len(jobs) == 75 and jobs[0] == (0, "01", 0) and jobs[-1] == (14, "15", 4)
    for world in (1, 2, 4, 8):
        parts = [shard_jobs(jobs, r, world) for r in range(world)]
        assert sorted(j for p in parts for j in p) == sorted(jobs)
        assert max(map(len, parts)) == -(-75 // world) and max(map(len, parts)) - min(map(len, parts)) <= 1
    with pytest.raises(ValueError):
        shard_jobs(jobs, 2, 2)
    assert list(JobQueue(jobs[:3])) == jobs[:3]
    assert len({job_seed(42, 5, s, k) for s in range(15) for k in range(5)}) == 75
    assert pick_best_fold([(0, 0.3), (1, 0.5), (2, 0.5)]) == (1, 0.5)          # strict >: ties keep the earlier fold
    rng = np.random.default_rng(0)
    yt, yp = rng.integers(0, 5, 200), rng.integers(0, 4, 200)
    assert abs(macro_f1(yt, yp) - f1_score(yt, yp, average="macro")) < 1e-12


def test_contiguous_shards_are_balanced_disjoint_and_subject_local():
    from fast.train.cv import cv_jobs, shard_jobs_contiguous
    jobs = cv_jobs([f"{i:02d}" for i in range(1, 16)], 5)
    for world in (1, 2, 3, 4, 8, 75, 80):
        parts = [shard_jobs_contiguous(jobs, r, world) for r in range(world)]
        assert sum(parts, []) == jobs                                   # disjoint, complete, order preserved
        sizes = [len(p) for p in parts]
        assert max(sizes) - min(sizes) <= 1 and max(sizes) == -(-75 // world)
        if world <= 8:
            assert max(len({sid for _, sid, _ in p}) for p in parts) <= -(-15 // world) + 1
    with pytest.raises(ValueError):
        shard_jobs_contiguous(jobs, 3, 3)


def test_synthetic_subject_is_balanced_and_seeded():
    from fast.train.cv import synthetic_subject
    (X, Y), (Xt, Yt) = synthetic_subject(3, n_trials=35, n_test=10, n_chan=4, n_samples=16)
    assert X.shape == (35, 4, 16) and X.dtype == np.float32 and Y.dtype == np.uint8 and Xt.shape == (10, 4, 16)
    assert np.bincount(Y).tolist() == [7] * 5 and np.bincount(Yt).tolist() == [2] * 5
    (X2, Y2), _ = synthetic_subject(3, n_trials=35, n_test=10, n_chan=4, n_samples=16)
    assert np.array_equal(X, X2) and np.array_equal(Y, Y2)


def test_finalize_cv_join(tmp_path):
    """The per-subject join reads only result files: best fold by strict >, reference CSV names and headers."""
    import json
    from fast.models import FAST
    from fast.train.cv import finalize_cv
    cfg = make_small_cfg()
    torch.manual_seed(0)
    sd = FAST(cfg).state_dict()
    for sid, accs in (("01", [0.4, 0.6, 0.6]), ("02", [0.7, 0.1, 0.2])):
        d = tmp_path / f"sub-{sid}"
        d.mkdir()
        for k, a in enumerate(accs):
            torch.save({n: v + k for n, v in sd.items()}, d / f"fold-{k}-best.pth")
            json.dump({"subject": sid, "fold": k, "best_val_acc": a}, open(d / f"fold-{k}_result.json", "w"))
    rows = finalize_cv(cfg, ["01", "02", "03"], str(tmp_path), n_folds=3, test_per_subject=None, device="cpu")
    assert [r[0] for r in rows] == ["01", "02"] and rows[0][1] == 0.6 and rows[1][1] == 0.7
    best = torch.load(tmp_path / "sub-01" / "best_subject.pth", weights_only=True)
    assert torch.equal(best["cls_token"], sd["cls_token"] + 1)                  # fold 1, not the tied fold 2
    assert open(tmp_path / "sub-02" / "fold_metrics.csv").read().splitlines()[0] == "Fold,Best_Val_Acc"
    assert open(tmp_path / "summary_per_subject.csv").read().splitlines()[0] == "Subject,Best_Val_Acc,Test_Acc,Test_F1"


# ------------------------------------------------------------------------------------------------ world_size 2, gloo
def _w_shard(rank, world):
    from fast.data import SUBJECTS
    from fast.train.cv import cv_jobs, shard_jobs, JobQueue
    jobs = cv_jobs(SUBJECTS, 5)
    mine = shard_jobs(jobs, rank, world)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    # dynamic queue through the process group's own key-value store: no collective, ranks race for indices
    store = dist.distributed_c10d._get_default_store()
    taken = list(JobQueue(jobs, store=store))
    dyn = [None] * world
    dist.all_gather_object(dyn, taken)
    return gathered, dyn


def test_cv_jobs_shard_over_two_ranks():
    from fast.data import SUBJECTS
    from fast.train.cv import cv_jobs
    jobs = [tuple(j) for j in cv_jobs(SUBJECTS, 5)]
    (static0, dyn0), (static1, dyn1) = _spawn(_w_shard, 2)
    assert static0 == static1
    parts = [[tuple(j) for j in p] for p in static0]
    assert len(parts[0]) == 38 and len(parts[1]) == 37 and sorted(parts[0] + parts[1]) == sorted(jobs)
    assert not set(parts[0]) & set(parts[1])
    d = [[tuple(j) for j in p] for p in dyn0]
    assert sorted(d[0] + d[1]) == sorted(jobs) and not set(d[0]) & set(d[1])


def _w_dp(rank, world, seed):
    """Each rank: oracle gradients of ITS half of the batch, scaled 1/world, SUM all-reduced as one flat buffer."""
    import numpy as _np
    from oracle import fast_oracle as O
    from fast.train.dp import allreduce_flat_, shard_batch
    golden = _np.load(os.path.join(ROOT, "tests", "golden", "fast_golden.npz"))
    cfg = make_small_cfg()
    sd = small_state_dict(golden)
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(8, 8, 500, generator=g)
    y = torch.randint(0, 3, (8,), generator=g)
    sl = shard_batch(8, rank, world)
    loss, _, grads, _ = O.loss_and_grads(sd, cfg, x[sl], y[sl])
    flat = torch.cat([grads[k].reshape(-1) for k in sd]) / world
    allreduce_flat_(flat)
    lt = loss.clone() / world
    dist.all_reduce(lt)
    return flat.numpy(), float(lt)


def test_dp_flat_gradient_exchange_equals_full_batch(golden):
    from oracle import fast_oracle as O
    from fast.train.dp import shard_batch
    (f0, l0), (f1, l1) = _spawn(_w_dp, 2, 5)
    assert np.array_equal(f0, f1) and l0 == l1                                  # replicas stay in lock-step
    cfg = make_small_cfg()
    sd = small_state_dict(golden)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(8, 8, 500, generator=g)
    y = torch.randint(0, 3, (8,), generator=g)
    loss, _, grads, _ = O.loss_and_grads(sd, cfg, x, y)
    full = torch.cat([grads[k].reshape(-1) for k in sd])
    assert rel_err(f0, full) < 1e-5 and abs(l0 - float(loss)) < 1e-6            # mean of equal-size shard means
    with pytest.raises(ValueError):
        shard_batch(7, 0, 2)


def test_dp_trainer_requires_process_group():
    from fast.models import FAST
    from fast.train import DataParallelTrainer
    with pytest.raises(RuntimeError, match="process group"):
        DataParallelTrainer(FAST(make_small_cfg()))
