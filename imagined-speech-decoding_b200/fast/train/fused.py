"""Device-resident fused training / evaluation step for FAST -- the throughput path behind the reference's
``training_step`` + ``AdamW.step`` + ``LambdaLR.step`` (src/fast/train/trainer.py:48-67).

One step = nine launches of ``libfastb200.so`` kernels plus two memsets on ONE stream, no host synchronisation:

    head_pack -> head_fwd -> transformer_fwd -> ce_fwd_bwd -> (zero grads) -> transformer_bwd -> head_bwd
              -> head_unpack_grad -> adamw_flat (lr and step counter read from device memory)

so the whole step is captured once per batch size into a CUDA graph and replayed.  Parameters stay the model's own
``nn.Parameter``s (views into the flat buffer), so ``state_dict()`` / checkpoints see every update.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _native as N


def lightning_lr_factors(cosine_lr_list):
    """LR factor actually applied at optimiser step 1, 2, ... by the reference's ``LambdaLR`` that indexes
    ``cosine_lr_list[global_step - 1]`` (trainer.py:50-53): the scheduler is built at ``global_step == 0`` (index -1),
    and is stepped after each optimiser step, so step k >= 2 runs at ``cosine_lr_list[k - 2]``."""
    s = np.asarray(cosine_lr_list, dtype=np.float64)
    return np.concatenate(([s[-1]], s[:-1]))


class FusedTrainer:
    """AdamW training of a ``FAST`` model with the whole step on the GPU.

    model        a ``fast.models.FAST`` on a CUDA device (production head ``Conv4Layers``)
    lr           base learning rate (reference: 5e-4)
    lr_factors   per-step multipliers in execution order (``lightning_lr_factors(cosine_scheduler(...))``), or None for
                 a constant rate
    forward_mode 'default' or 'train_head' ('train_transformer' = default with the head frozen)
    """

    def __init__(self, model, lr=5e-4, lr_factors=None, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01,
                 forward_mode='default', use_graph=True, seed=None):
        if not model.head.fused_ok(model.config.window_len, True):
            raise RuntimeError("FusedTrainer drives the fused Conv4Layers head only (dim_cnn 8/16/32, zones of <= 32 electrodes, "
                               "windows that fit the kernels); train other shapes through the nn.Module / autograd path")
        if forward_mode not in ('default', 'train_head', 'train_transformer'):
            raise NotImplementedError(forward_mode)
        self.model = model
        self.lib = N.lib()
        flat, desc = model._runtime()
        N.require_cuda(flat, "model parameters")
        self.device = flat.device
        self.flat, self.desc = flat, desc
        self._flat_version = model._fp.version
        self.forward_mode = forward_mode
        self.mode = N.MODE_TRAIN_HEAD if forward_mode == 'train_head' else N.MODE_DEFAULT
        self.freeze_head = forward_mode == 'train_transformer'
        self.betas, self.eps, self.weight_decay = betas, eps, weight_decay
        self.use_graph = use_graph
        dev = self.device
        n = flat.numel()
        self.grad = torch.zeros(n, device=dev)
        self.exp_avg = torch.zeros(n, device=dev)
        self.exp_avg_sq = torch.zeros(n, device=dev)
        self.step_dev = torch.zeros(1, device=dev, dtype=torch.int32)
        if lr_factors is None:
            self.lr_dev = torch.tensor([lr], device=dev, dtype=torch.float32)
            self.lr_by_step = 0
        else:
            self.lr_dev = torch.tensor(np.asarray(lr_factors, dtype=np.float64) * lr, device=dev, dtype=torch.float32)
            self.lr_by_step = 1
        self.n_sched = self.lr_dev.numel()
        self.steps_done = 0
        self.grad_scale = 1.0      # dlogits multiplier (a data-parallel run uses 1 / world_size: SUM all-reduce -> global mean)
        n_packed = self.lib.fastb200_head_packed_floats(desc)
        self.packed = torch.empty(n_packed, device=dev)
        self.gpacked = torch.zeros(n_packed, device=dev)
        self.loss = torch.zeros(1, device=dev)
        self.n_correct = torch.zeros(1, device=dev, dtype=torch.int32)
        self.seed = int(torch.initial_seed() if seed is None else seed) & 0x7FFFFFFFFFFFFFFF
        self._ws = {}          # (B, T, label dtype, train) -> workspace dict (+ graph)
        # mask of parameters the chosen mode does not train: their gradient stays zero, but AdamW's weight decay
        # would still move them -- the reference never hands them a .grad, so torch's AdamW skips them.
        self._frozen = None
        self._host_stage = None    # double-buffered device staging of host batches (_stage_host)
        if self.freeze_head or self.mode == N.MODE_TRAIN_HEAD:
            keep = torch.zeros(n, device=dev, dtype=torch.bool)
            for name, _, off, cnt in model._fp.items:
                if self.freeze_head:
                    trained = not name.startswith("head.")
                else:
                    trained = name.startswith(("head.", "input_layer.", "last_layer."))
                keep[off:off + cnt] = trained
            self._frozen = ~keep
            self._snapshot = flat.clone()

    # ------------------------------------------------------------------------------------------ workspaces
    def _p_drop_now(self):
        return float(self.model._p_drop) if (self.model.training and self.mode == N.MODE_DEFAULT) else 0.0

    def _rebind_if_moved(self):
        """The flat parameter buffer is re-created when the parameters move (``model.to()``,
        ``load_state_dict(assign=True)``): follow it and drop every captured graph, which holds the old addresses.  The
        AdamW moments keep their meaning (same layout)."""
        flat, desc = self.model._runtime()
        if self.model._fp.version == self._flat_version and flat.data_ptr() == self.flat.data_ptr():
            return
        if flat.device != self.device or flat.numel() != self.flat.numel():
            raise RuntimeError("the model's parameters moved to another device or changed size under a FusedTrainer")
        self.flat, self.desc, self._flat_version = flat, desc, self.model._fp.version
        if self._frozen is not None:
            self._snapshot = flat.clone()
        self._ws = {k: v for k, v in self._ws.items() if not isinstance(v, dict)}       # keep plain buffers, drop graphs

    def _workspace(self, B, T, label_dtype, train):
        from ..models import fast as FM
        # a captured graph bakes in the kernel choice AND every by-value launch argument that depends on host state: the
        # dropout probability (model.train() / .eval()) and the gradient scale
        p_drop = self._p_drop_now() if train else 0.0
        key = (B, T, label_dtype, train, FM.get_head_backward() if train else None, FM.get_head_precision(),
               p_drop, float(self.grad_scale) if train else None)
        ws = self._ws.get(key)
        if ws is not None:
            return ws
        d, dev, lib = self.desc, self.device, self.lib
        n_win = (T - d.window_len) // d.slide_step + 1
        n_terms = FM._TERMS[FM.get_head_precision()]
        ws = dict(B=B, T=T, n_win=n_win, graph=None, p_drop=p_drop, grad_scale=float(self.grad_scale), fwd_terms=0,
                  x=torch.empty(B, d.n_chan, T, device=dev),
                  y=torch.empty(B, device=dev, dtype=label_dtype),
                  feat=torch.empty(B, n_win, d.n_zones, d.dim_cnn, device=dev),
                  logits=torch.empty(B, d.n_classes, device=dev),
                  pred=torch.empty(B, device=dev, dtype=torch.int64))
        if train:
            ws["head_saved"] = torch.empty(lib.fastb200_head_saved_floats(d, B, n_win), device=dev)
            ws["tr_saved"] = torch.empty(max(1, lib.fastb200_transformer_saved_floats(d, B, n_win)), device=dev)
            ws["dlogits"] = torch.empty(B, d.n_classes, device=dev)
            ws["dfeat"] = torch.empty_like(ws["feat"])
            ws["bwd_work"] = None
        # The training forward and the backward are a pair: tensor-core forward (operand images saved) + tensor-core backward,
        # or CUDA-core forward (fp32 saved) + CUDA-core backward.  With the head frozen nothing is saved and the
        # tensor-core inference forward serves.
        tc_fwd = bool(n_terms and lib.fastb200_head_tc_supported(d))
        if train and not self.freeze_head:
            if tc_fwd and FM.head_bwd_uses_tc(d):
                ws["fwd_terms"] = n_terms
                ws["bwd_work"] = torch.empty(lib.fastb200_head_bwd_tc_work_floats(d, B, n_win), device=dev)
        elif tc_fwd:
            ws["fwd_terms"] = n_terms
        self._ws[key] = ws
        return ws

    def _packed_tc(self, n_terms):
        key = ("packed_tc", n_terms)
        if key not in self._ws:
            self._ws[key] = torch.empty(self.lib.fastb200_head_packed_tc_bytes(self.desc, n_terms), device=self.device,
                                        dtype=torch.uint8)
        return self._ws[key]

    # ------------------------------------------------------------------------------------------ launches
    def _launch_train(self, ws):
        self._launch_grad(ws)
        self._launch_update()

    def _launch_grad(self, ws):
        """Everything up to the complete flat gradient in ``self.grad`` (the buffer a data-parallel run all-reduces)."""
        lib, d, p = self.lib, self.desc, N.ptr
        st = N.stream_ptr(self.device)
        B, T, n_win = ws["B"], ws["T"], ws["n_win"]
        p_drop = ws["p_drop"]
        N.check(lib.fastb200_head_pack(d, p(self.flat), p(self.packed), st), "head_pack")
        n_terms = ws["fwd_terms"]
        if n_terms:
            # tensor-core forward; in training form it also writes y2 / y3 / y4 for the backward kernels
            pk = self._packed_tc(n_terms)
            N.check(lib.fastb200_head_pack_tc(d, p(self.flat), p(pk), n_terms, st), "head_pack_tc")
            if self.freeze_head:
                N.check(lib.fastb200_head_fwd_tc(d, p(pk), p(ws["x"]), B, T, n_win, d.slide_step, p(ws["feat"]), n_terms, st),
                        "head_fwd_tc")
            else:
                N.check(lib.fastb200_head_fwd_tc_train(d, p(pk), p(ws["x"]), B, T, n_win, d.slide_step, p(ws["feat"]),
                                                       p(ws["head_saved"]), n_terms, st), "head_fwd_tc_train")
        else:
            N.check(lib.fastb200_head_fwd(d, p(self.packed), p(ws["x"]), B, T, n_win, d.slide_step, p(ws["feat"]),
                                          None if self.freeze_head else p(ws["head_saved"]), st), "head_fwd")
        N.check(lib.fastb200_transformer_fwd(d, p(self.flat), p(ws["feat"]), B, n_win, self.mode, p_drop, self.seed, 0,
                                             p(self.step_dev), p(ws["logits"]), p(ws["tr_saved"]), st), "transformer_fwd")
        N.check(lib.fastb200_ce_fwd_bwd(p(ws["logits"]), p(ws["y"]), ws["y"].element_size(), B, d.n_classes,
                                        ws["grad_scale"], p(self.loss), p(ws["dlogits"]), p(ws["pred"]), p(self.n_correct), st),
                "ce")
        self.grad.zero_()
        N.check(lib.fastb200_transformer_bwd(d, p(self.flat), p(ws["feat"]), p(ws["tr_saved"]), p(ws["dlogits"]), B, n_win,
                                             self.mode, p_drop, self.seed, 0, p(self.step_dev), p(self.grad),
                                             None if self.freeze_head else p(ws["dfeat"]), st), "transformer_bwd")
        if not self.freeze_head:
            self.gpacked.zero_()
            if ws["bwd_work"] is not None:
                N.check(lib.fastb200_head_bwd_tc(d, p(self.packed), p(ws["x"]), p(ws["head_saved"]), p(ws["dfeat"]), B, T, n_win,
                                                 d.slide_step, p(self.gpacked), None, p(ws["bwd_work"]), st), "head_bwd_tc")
            else:
                N.check(lib.fastb200_head_bwd(d, p(self.packed), p(ws["x"]), p(ws["head_saved"]), p(ws["dfeat"]), B, T, n_win,
                                              d.slide_step, p(self.gpacked), None, st), "head_bwd")
            N.check(lib.fastb200_head_unpack_grad(d, p(self.flat), p(self.gpacked), p(self.grad), st), "head_unpack_grad")

    def _launch_update(self):
        """AdamW on the flat buffers (lr and step counter read on the device)."""
        lib, p = self.lib, N.ptr
        st = N.stream_ptr(self.device)
        N.check(lib.fastb200_adamw_flat(p(self.flat), p(self.grad), p(self.exp_avg), p(self.exp_avg_sq), self.flat.numel(),
                                        p(self.lr_dev), self.lr_by_step, self.betas[0], self.betas[1], self.eps,
                                        self.weight_decay, p(self.step_dev), 1, st), "adamw_flat")
        if self._frozen is not None:
            torch.where(self._frozen, self._snapshot, self.flat, out=self.flat)

    def _launch_eval(self, ws):
        lib, d, p = self.lib, self.desc, N.ptr
        st = N.stream_ptr(self.device)
        B, T, n_win = ws["B"], ws["T"], ws["n_win"]
        from ..models import fast as FM
        n_terms = FM._TERMS[FM.get_head_precision()]
        if n_terms and lib.fastb200_head_tc_supported(d):
            # evaluation needs no saved activations: tensor-core head (same dispatch as FAST.forward_head under no_grad)
            pk = self._packed_tc(n_terms)
            N.check(lib.fastb200_head_pack_tc(d, p(self.flat), p(pk), n_terms, st), "head_pack_tc")
            N.check(lib.fastb200_head_fwd_tc(d, p(pk), p(ws["x"]), B, T, n_win, d.slide_step, p(ws["feat"]),
                                             n_terms, st), "head_fwd_tc")
        else:
            N.check(lib.fastb200_head_pack(d, p(self.flat), p(self.packed), st), "head_pack")
            N.check(lib.fastb200_head_fwd(d, p(self.packed), p(ws["x"]), B, T, n_win, d.slide_step, p(ws["feat"]), None, st),
                    "head_fwd")
        N.check(lib.fastb200_transformer_fwd(d, p(self.flat), p(ws["feat"]), B, n_win, self.mode, 0.0, 0, 0, None,
                                             p(ws["logits"]), None, st), "transformer_fwd")
        N.check(lib.fastb200_ce_fwd_bwd(p(ws["logits"]), p(ws["y"]), ws["y"].element_size(), B, d.n_classes, 1.0,
                                        p(self.loss), None, p(ws["pred"]), p(self.n_correct), st), "ce")

    #: kernels of libfastb200.so launched per call (the memsets / frozen-parameter restore are ATen and not counted)
    LAUNCHES_TRAIN = 9      # pack, head_fwd, tr_fwd, ce, tr_bwd, head_bwd, unpack, adamw, bump_step
    LAUNCHES_EVAL = 4

    def _phases(self, ws, train):
        """[(callable, capturable in a CUDA graph)] making up one step, in order."""
        if not train:
            return [(lambda: self._launch_eval(ws), True)]
        return [(lambda: self._launch_train(ws), True)]

    def _stage_host(self, x, y):
        """Host batch -> one of two device staging pairs on a SIDE stream; the caller's stream waits for the copy only.  A call
        returns as soon as its work is enqueued, so the copy of batch k + 1 (issued by the next call) runs beside the kernels
        of batch k: end to end the step costs max(copy, kernels) instead of their sum (pin the host tensors for that)."""
        key = (tuple(x.shape), y.dtype)
        st = self._host_stage
        if st is None or st["key"] != key:
            st = dict(key=key, k=0, x=[torch.empty(x.shape, device=self.device) for _ in range(2)],
                      y=[torch.empty(y.shape, dtype=y.dtype, device=self.device) for _ in range(2)],
                      copied=[torch.cuda.Event() for _ in range(2)], freed=[torch.cuda.Event() for _ in range(2)],
                      stream=torch.cuda.Stream(self.device))
            st["stream"].wait_stream(torch.cuda.current_stream(self.device))
            self._host_stage = st
        b = st["k"] & 1
        side = st["stream"]
        with torch.cuda.stream(side):
            if st["k"] >= 2:
                side.wait_event(st["freed"][b])
            st["x"][b].copy_(x, non_blocking=True)
            st["y"][b].copy_(y, non_blocking=True)
            st["copied"][b].record(side)
        torch.cuda.current_stream(self.device).wait_event(st["copied"][b])
        st["k"] += 1
        return st["x"][b], st["y"][b], st["freed"][b]

    def _run(self, x, y, train):
        if x.dtype != torch.float32:
            x = x.float()
        if y.dtype not in (torch.uint8, torch.int64):
            y = y.long()
        self._rebind_if_moved()
        freed = None
        if not x.is_cuda:                                    # host batch (pinned for full speed): staged on a side stream
            with torch.cuda.device(self.device):
                x, y, freed = self._stage_host(x, y.cpu() if y.is_cuda else y)
        N.require_cuda(x, "x")
        ws = self._workspace(x.shape[0], x.shape[2], y.dtype, train)
        phases = self._phases(ws, train)
        with torch.cuda.device(self.device):
            ws["x"].copy_(x, non_blocking=True)
            ws["y"].copy_(y.to(self.device, non_blocking=True), non_blocking=True)
            if freed is not None:
                freed.record(torch.cuda.current_stream(self.device))
            if not self.use_graph:
                for fn, _ in phases:
                    fn()
            elif ws["graph"] is None:
                # one eager run (sets kernel attributes, warms allocators), then capture; capture itself does not
                # execute, so the eager run IS this step and the graphs serve the following ones.  Phases flagged
                # non-capturable (a collective between two graphs) are kept as plain calls.
                for fn, _ in phases:
                    fn()
                replay = []
                side = torch.cuda.Stream(self.device)
                for fn, capturable in phases:
                    if not capturable:
                        replay.append(fn)
                        continue
                    g = torch.cuda.CUDAGraph()
                    side.wait_stream(torch.cuda.current_stream(self.device))
                    c0 = self.lib.fastb200_launch_count()
                    with torch.cuda.stream(side):
                        with torch.cuda.graph(g, stream=side):
                            fn()
                    torch.cuda.current_stream(self.device).wait_stream(side)
                    n_launch = self.lib.fastb200_launch_count() - c0       # kernels recorded into this graph

                    def replay_fn(g=g, n_launch=n_launch):
                        g.replay()
                        N.note_replayed(n_launch)
                    replay.append(replay_fn)
                ws["graph"] = replay
            else:
                for fn in ws["graph"]:
                    fn()
        return ws

    # ------------------------------------------------------------------------------------------ public API
    def train_step(self, x, y):
        """One optimiser step on batch (x (B,C,T) float32, y (B,) uint8/int64; CUDA tensors, or HOST tensors -- pinned for full
        speed -- which are uploaded on a side stream so that the next call's copy overlaps this call's kernels).  Returns (loss, n_correct):
        1-element device tensors that the NEXT call overwrites -- read them (``.item()``) only when needed."""
        if self.lr_by_step and self.steps_done >= self.n_sched:
            raise RuntimeError(f"LR schedule has {self.n_sched} entries; step {self.steps_done + 1} is past its end")
        self._run(x, y, True)
        self.steps_done += 1
        return self.loss, self.n_correct

    @torch.no_grad()
    def eval_step(self, x, y):
        """Forward + CE + argmax on one batch (no dropout).  Returns (loss, n_correct, pred) device tensors."""
        ws = self._run(x, y, False)
        return self.loss, self.n_correct, ws["pred"]

    @torch.no_grad()
    def predict(self, x):
        """argmax class per trial, int64 CUDA tensor (the inner op of ``inference_on_loader``, trainer.py:89)."""
        y = torch.zeros(x.shape[0], device=self.device, dtype=torch.uint8)
        return self._run(x, y, False)["pred"]
