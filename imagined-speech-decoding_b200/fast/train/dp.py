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
