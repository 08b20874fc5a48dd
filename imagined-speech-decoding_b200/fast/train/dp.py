"""Batch data-parallel training of FAST across the GPUs of one NVSwitch box (BASELINE.json configs[4]).

The reference is single-device (``devices=[args.gpu]``, scripts/train_fast.py:127-131), so this is additive.  One process
per GPU; every rank holds a replica of the 767 KB model and its AdamW state, runs the fused forward / backward on its own
slice of the batch, and the ONE exchange step of the path is the sum of the flat fp32 gradient (191,781 elements):

``reduce='p2p'``   (default on NVLink peers) the gradient buffer lives in torch symmetric memory; after a stream-ordered
                   barrier every rank's ``fastb200_adamw_flat_peers`` kernel reads all ranks' gradients straight over
                   NVLink, sums them in rank order (bit-identical on every rank) and applies AdamW in the same pass -- the
                   all-reduce never materialises, and there is no NCCL launch on the step.
``reduce='nccl'``  one ``ncclAllReduce`` on the flat buffer between the gradient graph and the update graph.

Both keep the step free of host synchronisation.  ``grad_scale = 1 / world`` makes the summed gradient the gradient of
the global-batch mean loss (equal per-rank batch sizes), i.e. exactly the single-GPU step on the concatenated batch.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from .. import _native as N
from .fused import FusedTrainer


def allreduce_flat_(grad: torch.Tensor, group=None) -> torch.Tensor:
    """SUM all-reduce of a flat gradient buffer in place (NCCL on CUDA tensors; gloo in the CPU tests)."""
    dist.all_reduce(grad, op=dist.ReduceOp.SUM, group=group)
    return grad


def shard_batch(n: int, rank: int, world: int) -> slice:
    """Equal contiguous slices of a global batch of ``n`` trials (n must divide evenly: the mean-of-means identity the
    gradient scaling relies on needs equal per-rank batch sizes)."""
    if n % world:
        raise ValueError(f"global batch {n} is not divisible by world size {world}")
    per = n // world
    return slice(rank * per, (rank + 1) * per)


class DataParallelTrainer(FusedTrainer):
    """``FusedTrainer`` whose optimiser step consumes the gradient summed over all ranks of ``group``."""

    def __init__(self, model, group=None, reduce="p2p", broadcast_from=0, **kw):
        if not dist.is_initialized():
            raise RuntimeError("DataParallelTrainer needs an initialised torch.distributed process group")
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if reduce not in ("p2p", "nccl"):
            raise ValueError(reduce)
        self.reduce = reduce
        seed = kw.pop("seed", None)
        seed = int(torch.initial_seed() if seed is None else seed)
        super().__init__(model, seed=seed + 7919 * self.rank, **kw)      # independent dropout masks per rank
        self.grad_scale = 1.0 / self.world
        # replicas start from rank `broadcast_from`'s weights
        dist.broadcast(self.flat, src=dist.get_global_rank(self.group, broadcast_from), group=self.group)
        if self._frozen is not None:
            self._snapshot.copy_(self.flat)
        self._symm = None
        if self.reduce == "p2p" and self.world > 1:
            import torch.distributed._symmetric_memory as symm_mem
            if self.world > 16:
                raise RuntimeError("p2p reduce supports at most 16 peers")
            n = self.flat.numel()
            buf = symm_mem.empty(n, dtype=torch.float32, device=self.device)
            self._symm = symm_mem.rendezvous(buf, self.group)
            buf.zero_()
            self.grad = buf                                             # the kernels now write the gradient into peer-visible memory
            self._peer_ptrs = (C.c_void_p * self.world)(*[int(p) for p in self._symm.buffer_ptrs])
            assert int(self._symm.buffer_ptrs[self.rank]) == buf.data_ptr()

    # ------------------------------------------------------------------------------------------ step
    def _exchange(self):
        allreduce_flat_(self.grad, self.group)

    def _launch_update_peers(self):
        lib, p = self.lib, N.ptr
        st = N.stream_ptr(self.device)
        self._symm.barrier(channel=0)        # every rank's gradient is complete and visible
        N.check(lib.fastb200_adamw_flat_peers(p(self.flat), self._peer_ptrs, self.world, p(self.exp_avg), p(self.exp_avg_sq),
                                              self.flat.numel(), p(self.lr_dev), self.lr_by_step, self.betas[0],
                                              self.betas[1], self.eps, self.weight_decay, p(self.step_dev), 1, st),
                "adamw_flat_peers")
        self._symm.barrier(channel=1)        # nobody re-zeroes its gradient while a peer still reads it
        if self._frozen is not None:
            torch.where(self._frozen, self._snapshot, self.flat, out=self.flat)

    def _phases(self, ws, train):
        if not train or self.world == 1:
            return super()._phases(ws, train)
        if self._symm is not None:
            # barriers are plain kernels on the stream, but a captured barrier replays fixed signal-pad state only if
            # every rank replays in lock-step; keep them outside the graphs.
            return [(lambda: self._launch_grad(ws), True), (self._launch_update_peers, False)]
        return [(lambda: self._launch_grad(ws), True), (self._exchange, False), (self._launch_update, True)]

    # ------------------------------------------------------------------------------------------ reporting
    def global_loss_and_correct(self):
        """(mean loss over the global batch, total correct) of the LAST step -- one tiny all-reduce, host readable."""
        t = torch.stack([self.loss[0].double() / self.world, self.n_correct[0].double()])
        dist.all_reduce(t, group=self.group)
        return float(t[0]), int(t[1])


def dp_selfcheck(config, reduce="p2p", steps=4, per_rank=2, group=None, seed=77):
    """Run inside an initialised process group (one rank per GPU): ``steps`` data-parallel steps on a synthetic global batch
    of ``per_rank * world`` trials, then check (1) every replica's flat parameter buffer is BIT-identical to rank 0's and
    (2) rank 0's losses / parameters equal a single-GPU ``FusedTrainer`` run on the concatenated batch.  Returns a dict
    (same on every rank) -- the evidence bench.py / smoke() print for the exchange path (``fastb200_adamw_flat_peers`` for
    ``reduce='p2p'``, one ``ncclAllReduce`` for ``'nccl'``)."""
    from ..models import FAST
    group = group if group is not None else dist.group.WORLD
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = torch.device("cuda", torch.cuda.current_device())
    n = per_rank * world
    C, T = len(config.electrodes), config.seq_len
    factors = [1.0, 0.5, 0.25, 1.0] * ((steps + 3) // 4)
    torch.manual_seed(100 + rank)                               # different initial replicas: the broadcast must fix that
    model = FAST(config).to(dev).train()
    ft = DataParallelTrainer(model, group=group, reduce=reduce, lr=5e-4, lr_factors=factors[:steps], use_graph=True, seed=1)
    g = torch.Generator().manual_seed(seed)
    batches = [(torch.randn(n, C, T, generator=g), torch.randint(0, config.n_classes, (n,), generator=g).to(torch.uint8))
               for _ in range(steps)]
    sl = shard_batch(n, rank, world)
    losses = []
    for x, y in batches:
        ft.train_step(x[sl].to(dev), y[sl].to(dev))
        losses.append(ft.global_loss_and_correct()[0])
    flats = [torch.empty_like(ft.flat) for _ in range(world)]
    dist.all_gather(flats, ft.flat.contiguous(), group=group)
    identical = all(torch.equal(flats[0], f) for f in flats[1:])
    out = torch.zeros(3, device=dev, dtype=torch.float64)
    if rank == 0:
        torch.manual_seed(100)
        ref_model = FAST(config).to(dev).train()
        ref = FusedTrainer(ref_model, lr=5e-4, lr_factors=factors[:steps], use_graph=False, seed=1)
        ref_losses = [float(ref.train_step(x.to(dev), y.to(dev))[0]) for x, y in batches]
        worst_loss = max(abs(a - b) / max(abs(b), 1e-12) for a, b in zip(losses, ref_losses))
        # Adam turns rounding noise into +-lr steps where the true gradient is ~0 (e.g. the key third of in_proj_bias): compare
        # the parameter buffers in the max norm relative to the largest parameter
        worst_param = float((ft.flat - ref.flat).abs().max() / ref.flat.abs().max())
        out = torch.tensor([worst_loss, worst_param, 1.0], device=dev, dtype=torch.float64)
    dist.broadcast(out, src=dist.get_global_rank(group, 0), group=group)
    return {"reduce": reduce if world > 1 else "none", "world": world, "steps": steps, "global_batch": n,
            "replicas_bit_identical": bool(identical), "loss_rel_err_vs_single_gpu": float(out[0]),
            "param_rel_err_vs_single_gpu": float(out[1]),
            "ok": bool(identical and float(out[0]) < 1e-4 and float(out[1]) < 2e-3)}
