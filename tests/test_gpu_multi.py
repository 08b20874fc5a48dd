"""GPU tests of the cross-validation job runner and the data-parallel step.

One-GPU tests always run; the 2-rank NCCL / peer-memory tests run when the box has >= 2 GPUs."""
import json
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist

from conftest import make_small_cfg, rel_err
from oracle import fast_oracle as O

pytestmark = pytest.mark.gpu


def _small_subject(n=40, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 8, 500, generator=g)
    y = (torch.arange(n) % 3).to(torch.uint8)
    x[:, 0, :] += (y.float() - 1.0)[:, None] * 2.0            # a learnable offset so accuracy moves
    return x, y


def test_train_fold_matches_oracle_loop_and_keeps_best_epoch():
    """Two epochs of ``train_fold`` reproduce, batch for batch, the oracle's CE + AdamW + LambdaLR loop fed with the same
    device-side shuffles; the returned best state is the first epoch reaching the maximal val_acc."""
    from fast.models import FAST
    from fast.train.cv import kfold_indices, train_fold
    from fast.train.fused import lightning_lr_factors
    cfg = make_small_cfg()
    x, y = _small_subject()
    tr, va = kfold_indices(40, 4, True, 1)[0]
    epochs, bs, seed = 3, 16, 9
    out = train_fold(cfg, x.cuda(), y.cuda(), tr, va, max_epochs=epochs, batch_size=bs, seed=seed)
    h = out["history"]
    assert [len(h[k]) for k in ("loss", "acc", "val_loss", "val_acc")] == [epochs] * 4
    assert out["train_len"] == 2 and out["steps"] == epochs * 2
    assert out["best_val_acc"] == max(h["val_acc"]) and out["best_epoch"] == int(np.argmax(h["val_acc"]))

    # oracle replay with the same initial weights and the same permutations
    torch.manual_seed(seed)
    sd = {k: v.detach().clone() for k, v in FAST(cfg).state_dict().items()}
    m = {k: torch.zeros_like(v) for k, v in sd.items()}
    v2 = {k: torch.zeros_like(v) for k, v in sd.items()}
    gen = torch.Generator(device="cuda").manual_seed(seed)
    fac = lightning_lr_factors(O.cosine_scheduler(1, 0.1, epochs, 2, warmup_epochs=0))
    trt = torch.as_tensor(tr)
    step, ref_loss, ref_vacc, states = 0, [], [], []
    for ep in range(epochs):
        perm = trt[torch.randperm(len(tr), device="cuda", generator=gen).cpu()]
        tot = 0.0
        for lo in range(0, len(tr), bs):
            idx = perm[lo:lo + bs]
            loss, _, grads, _ = O.loss_and_grads(sd, cfg, x[idx], y[idx])
            tot += float(loss) * len(idx)
            step += 1
            for k in sd:
                O.adamw_step(sd[k], grads[k], m[k], v2[k], step, 5e-4 * fac[step - 1])
        ref_loss.append(tot / len(tr))
        with torch.no_grad():
            ref_vacc.append(float((O.forward(sd, cfg, x[va]).argmax(1) == y[va].long()).double().mean()))
        states.append({k: t.clone() for k, t in sd.items()})
    np.testing.assert_allclose(h["loss"], ref_loss, rtol=2e-5)
    assert h["val_acc"] == pytest.approx(ref_vacc, abs=1e-9)
    best = states[out["best_epoch"]]
    for k, t in out["state_dict"].items():
        if k.endswith("in_proj_bias"):
            continue        # the key-bias third has zero true gradient: Adam turns rounding noise into +-lr updates in ANY implementation
        assert rel_err(t, best[k]) < 2e-3, k                    # a few AdamW steps: compare loosely, exactness is tested elsewhere


def test_concurrent_folds_equal_folds_run_alone(tmp_path):
    """``train_folds`` (one epoch graph with a parallel branch per fold) gives every fold the result it has when trained
    alone -- production config (tensor-core kernels), three folds of one subject sharing x -- and ``run_cv_jobs(concurrent=3)``
    writes the same artefacts as the sequential runner."""
    from conftest import make_prod_cfg
    from fast.train.cv import cv_jobs, kfold_indices, run_cv_jobs, train_fold, train_folds
    cfg = make_prod_cfg()
    g = torch.Generator().manual_seed(4)
    x = torch.randn(36, 64, 800, generator=g)
    y = (torch.arange(36) % 5).to(torch.uint8)
    x[:, 3, :] += (y.float() - 2.0)[:, None]
    xd, yd = x.cuda(), y.cuda()
    folds = kfold_indices(36, 3, True, 1)
    epochs, bs = 4, 16
    alone = [train_fold(cfg, xd, yd, tr, va, max_epochs=epochs, batch_size=bs, seed=20 + k) for k, (tr, va) in enumerate(folds)]
    both = train_folds(cfg, [dict(x=xd, y=yd, train_idx=tr, val_idx=va, seed=20 + k) for k, (tr, va) in enumerate(folds)],
                       max_epochs=epochs, batch_size=bs)
    assert len(both) == 3
    for a, b in zip(alone, both):
        # same arithmetic, same dropout / shuffle streams; only the order of the fp32 atomics inside the weight-gradient
        # kernels differs from run to run
        np.testing.assert_allclose(b["history"]["loss"], a["history"]["loss"], rtol=2e-3)
        np.testing.assert_allclose(b["history"]["val_loss"], a["history"]["val_loss"], rtol=2e-3)
        assert np.abs(np.array(b["history"]["val_acc"]) - np.array(a["history"]["val_acc"])).max() <= 1.0 / 12 + 1e-9
        assert b["steps"] == a["steps"] == epochs * 2
        for k in a["state_dict"]:
            assert b["state_dict"][k].shape == a["state_dict"][k].shape
    # the runner: concurrent groups leave the same files as one-at-a-time
    data = {"01": (x, y)}
    jobs = cv_jobs(["01"], 3)
    d1 = run_cv_jobs(cfg, jobs, lambda s: data[s], str(tmp_path / "seq"), max_epochs=2, batch_size=bs, seed=3, n_folds=3)
    d3 = run_cv_jobs(cfg, jobs, lambda s: data[s], str(tmp_path / "par"), max_epochs=2, batch_size=bs, seed=3, n_folds=3, concurrent=3)
    assert [r["fold"] for r in d3] == [r["fold"] for r in d1] == [0, 1, 2]
    for k in range(3):
        for name in (f"fold-{k}-best.pth", f"fold-{k}_history.csv", f"fold-{k}_result.json"):
            assert (tmp_path / "par" / "sub-01" / name).exists()
        h1 = np.loadtxt(tmp_path / "seq" / "sub-01" / f"fold-{k}_history.csv", delimiter=",", skiprows=1)
        h3 = np.loadtxt(tmp_path / "par" / "sub-01" / f"fold-{k}_history.csv", delimiter=",", skiprows=1)
        np.testing.assert_allclose(h3[:, 1], h1[:, 1], rtol=2e-3)


def test_run_cv_jobs_artifacts_resume_and_finalize(tmp_path):
    from fast.train.cv import cv_jobs, finalize_cv, run_cv_jobs, shard_jobs
    cfg = make_small_cfg()
    data = {sid: _small_subject(30, seed=i) for i, sid in enumerate(["01", "02"])}
    test = {sid: _small_subject(12, seed=10 + i) for i, sid in enumerate(["01", "02"])}
    jobs = cv_jobs(["01", "02"], 3)
    # two "ranks" run one after the other on this GPU: disjoint shards, same output tree
    for r in range(2):
        done = run_cv_jobs(cfg, shard_jobs(jobs, r, 2), lambda s: data[s], str(tmp_path), max_epochs=2, batch_size=8,
                           seed=3, n_folds=3)
        assert len(done) == 3
    for sid in ("01", "02"):
        for k in range(3):
            d = tmp_path / f"sub-{sid}"
            assert (d / f"fold-{k}-best.pth").exists()
            lines = open(d / f"fold-{k}_history.csv").read().splitlines()
            assert lines[0] == "Epoch,loss,acc,val_loss,val_acc" and len(lines) == 3
            assert json.load(open(d / f"fold-{k}_result.json"))["fold"] == k
    # resume: every job is skipped (no subject is even loaded)
    again = run_cv_jobs(cfg, jobs, lambda s: (_ for _ in ()).throw(AssertionError("reloaded")), str(tmp_path), max_epochs=2,
                        batch_size=8, seed=3, n_folds=3)
    assert len(again) == 6
    rows = finalize_cv(cfg, ["01", "02"], str(tmp_path), n_folds=3, test_per_subject=test)
    assert len(rows) == 2 and all(0.0 <= r[2] <= 1.0 for r in rows)
    pred = np.loadtxt(tmp_path / "sub-01" / "test_predictions.csv", delimiter=",", dtype=int)
    assert pred.shape == (12, 2) and open(tmp_path / "sub-01" / "test_predictions.csv").readline().strip() == "# Predicted,True"
    assert np.loadtxt(tmp_path / "global_test_predictions.csv", delimiter=",").shape == (24, 2)
    # the best-subject weights reproduce the stored predictions through the public model API
    from fast.models import FAST
    m = FAST(cfg).cuda()
    m.load_state_dict(torch.load(tmp_path / "sub-01" / "best_subject.pth", weights_only=True))
    assert np.array_equal(m.predict(test["01"][0].cuda()).cpu().numpy(), pred[:, 0])


# ------------------------------------------------------------------------------------------------ 2 ranks on 2 GPUs
def _dp_worker(rank, world, port, reduce, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from fast.models import FAST
        from fast.train import DataParallelTrainer
        from fast.train.dp import shard_batch
        cfg = make_small_cfg()
        torch.manual_seed(100 + rank)                              # different initial replicas: the broadcast must fix that
        model = FAST(cfg).cuda().train()
        ft = DataParallelTrainer(model, reduce=reduce, lr=5e-4, lr_factors=[1.0, 0.5, 0.25, 1.0], use_graph=True)
        g = torch.Generator().manual_seed(77)
        losses = []
        for _ in range(4):
            x = torch.randn(8, 8, 500, generator=g)
            y = torch.randint(0, 3, (8,), generator=g).to(torch.uint8)
            sl = shard_batch(8, rank, world)
            ft.train_step(x[sl].cuda(), y[sl].cuda())
            losses.append(ft.global_loss_and_correct()[0])
        torch.cuda.synchronize()
        q.put((rank, losses, {k: v.detach().cpu() for k, v in model.state_dict().items()}))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("reduce", ["nccl", "p2p"])
def test_data_parallel_two_ranks_equal_single_gpu_on_full_batch(reduce):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import socket
    import torch.multiprocessing as mp
    from fast.models import FAST
    from fast.train import FusedTrainer
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, reduce, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = {}
    try:
        for _ in range(2):
            r, losses, sd = q.get(timeout=300)
            res[r] = (losses, sd)
    finally:
        for p in procs:
            p.join(60)
            if p.is_alive():
                p.kill()
    assert all(p.exitcode == 0 for p in procs)
    for k in res[0][1]:
        assert torch.equal(res[0][1][k], res[1][1][k]), k           # replicas bit-identical after 4 steps
    # single GPU, full batch, rank 0's initial weights
    cfg = make_small_cfg()
    torch.manual_seed(100)
    model = FAST(cfg).cuda().train()
    ft = FusedTrainer(model, lr=5e-4, lr_factors=[1.0, 0.5, 0.25, 1.0], use_graph=False)
    g = torch.Generator().manual_seed(77)
    ref_losses = []
    for _ in range(4):
        x = torch.randn(8, 8, 500, generator=g)
        y = torch.randint(0, 3, (8,), generator=g).to(torch.uint8)
        ref_losses.append(float(ft.train_step(x.cuda(), y.cuda())[0]))
    np.testing.assert_allclose(res[0][0], ref_losses, rtol=2e-5)
    p0 = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    # in_proj_bias starts at zero and its key third has zero true gradient: Adam turns rounding noise into +-lr steps there
    worst = max(rel_err(res[0][1][k], p0[k]) for k in p0 if not k.endswith("in_proj_bias"))
    assert worst < 1e-3, worst


def test_data_parallel_world_one_is_the_fused_step():
    """A 1-rank group takes the plain fused path (no exchange) and matches FusedTrainer."""
    import socket
    from fast.models import FAST
    from fast.train import DataParallelTrainer, FusedTrainer
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    try:
        cfg = make_small_cfg()
        x, y = _small_subject(8)
        outs = []
        for cls in (FusedTrainer, DataParallelTrainer):
            torch.manual_seed(5)
            m = FAST(cfg).cuda().train()
            ft = cls(m, lr=5e-4, use_graph=False, seed=1)
            for _ in range(3):
                ft.train_step(x.cuda(), y.cuda())
            outs.append(ft.flat.clone())
        # same kernels, same order; only the order of the fp32 atomicAdds of the weight-gradient flush may differ
        assert rel_err(outs[1], outs[0]) < 2e-3        # 3 Adam steps amplify that rounding noise where the true gradient is ~0
    finally:
        dist.destroy_process_group()
