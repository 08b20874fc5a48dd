"""Lightning-facing training module for FAST (drop-in for the reference's ``src/fast/train/trainer.py``).

``EEG_Encoder_Module`` keeps the reference's constructor, attributes and step methods (trainer.py:30-79): the model is
the CUDA-kernel ``FAST``; the loss is ``nn.CrossEntropyLoss`` on its logits; the optimiser is ``AdamW(lr=5e-4)`` with
the per-step ``LambdaLR`` that indexes ``cosine_lr_list[global_step - 1]``.  Lightning / torchmetrics are optional on
this image: with them the class is a ``LightningModule`` using torchmetrics; without them it is a plain ``nn.Module``
with the same methods, a dict-backed ``log`` and small on-device metric accumulators.

For throughput use ``fast.train.FusedTrainer`` (device-resident data, one CUDA graph per step) -- same arithmetic.
"""
import numpy as np
import torch
import torch.nn as nn
import torch.optim as optim

from ..models import FAST

try:
    import lightning as pl
    _Base = pl.LightningModule
except Exception:                                   # pragma: no cover - depends on the environment
    pl = None
    _Base = nn.Module

try:
    import torchmetrics
except Exception:                                   # pragma: no cover
    torchmetrics = None


def cosine_scheduler(base_value, final_value, epochs, niter_per_ep, warmup_epochs=0, start_warmup_value=0):
    """One LR factor per optimiser step: linear warm-up over ``warmup_epochs`` then a half cosine from ``base_value``
    down to ``final_value`` (trainer.py:15-27).  Raises AssertionError when the warm-up is longer than the run."""
    total = epochs * niter_per_ep
    warm = warmup_epochs * niter_per_ep
    ramp = np.linspace(start_warmup_value, base_value, warm) if warmup_epochs > 0 else np.array([])
    n = total - warm
    k = np.arange(n)
    decay = final_value + 0.5 * (base_value - final_value) * (1 + np.cos(np.pi * k / len(k)))
    schedule = np.concatenate((ramp, decay))
    assert len(schedule) == total
    return schedule


class _ConfusionMetric(nn.Module):
    """Multiclass accuracy / macro-F1 from a running confusion matrix (stands in for torchmetrics when absent).
    Calling it returns the metric of the batch and accumulates, like ``torchmetrics.Metric.forward``."""

    def __init__(self, num_classes, kind):
        super().__init__()
        self.num_classes, self.kind = num_classes, kind
        self.register_buffer("confusion", torch.zeros(num_classes, num_classes, dtype=torch.long), persistent=False)

    def _value(self, cm):
        cm = cm.double()
        tp = cm.diag()
        if self.kind == 'accuracy':
            return (tp.sum() / cm.sum().clamp_min(1)).float()
        denom = cm.sum(0) + cm.sum(1)                       # 2TP + FP + FN
        f1 = torch.where(denom > 0, 2 * tp / denom.clamp_min(1), torch.zeros_like(tp))
        seen = denom > 0                                    # classes absent from both target and prediction are skipped
        return (f1.sum() / seen.sum().clamp_min(1)).float()

    def forward(self, logits, target):
        pred = logits.argmax(dim=1)
        idx = target.long() * self.num_classes + pred
        cm = torch.bincount(idx, minlength=self.num_classes ** 2).view(self.num_classes, self.num_classes)
        self.confusion += cm
        return self._value(cm)

    def compute(self):
        return self._value(self.confusion)

    def reset(self):
        self.confusion.zero_()


def _metric(kind, num_classes):
    if torchmetrics is not None:
        if kind == 'accuracy':
            return torchmetrics.Accuracy('multiclass', num_classes=num_classes)
        return torchmetrics.F1Score('multiclass', num_classes=num_classes, average='macro')
    return _ConfusionMetric(num_classes, kind)


class EEG_Encoder_Module(_Base):
    """Wrapper that trains ``FAST`` (trainer.py:30-79)."""

    def __init__(self, config, max_epochs, niter_per_ep):
        super().__init__()
        self.config = config
        self.model = FAST(config)
        self.loss = nn.CrossEntropyLoss()
        self.cosine_lr_list = cosine_scheduler(1, 0.1, max_epochs, niter_per_ep, warmup_epochs=10)
        self.train_acc = _metric('accuracy', config.n_classes)
        self.train_f1 = _metric('f1', config.n_classes)
        self.val_acc = _metric('accuracy', config.n_classes)
        self.val_f1 = _metric('f1', config.n_classes)
        if pl is None:
            self.global_step = 0
            self.logged = {}

    if pl is None:
        def log(self, name, value, **_):
            self.logged[name] = value

    def configure_optimizers(self):
        self.optimizer = optim.AdamW(self.parameters(), lr=0.0005)
        self.scheduler = optim.lr_scheduler.LambdaLR(self.optimizer, lambda epoch: self.cosine_lr_list[self.global_step - 1])
        return [self.optimizer], [{'scheduler': self.scheduler, 'interval': 'step'}]

    def _shared_step(self, batch):
        x, y = batch
        logits = self.model(x)
        return logits, y, self.loss(logits, y)

    def training_step(self, batch, batch_idx):
        logits, y, loss = self._shared_step(batch)
        self.log('train_loss', loss, on_step=False, on_epoch=True, prog_bar=True)
        self.log('train_acc', self.train_acc(logits, y), on_step=False, on_epoch=True, prog_bar=True)
        self.log('train_f1', self.train_f1(logits, y), on_step=False, on_epoch=True, prog_bar=True)
        return loss

    def validation_step(self, batch, batch_idx):
        logits, y, loss = self._shared_step(batch)
        self.val_f1(logits, y)
        self.log('val_loss', loss, on_step=False, on_epoch=True, prog_bar=True)
        self.log('val_acc', self.val_acc(logits, y), on_step=False, on_epoch=True, prog_bar=True)
        return loss


def inference_on_loader(model, loader):
    """Predictions and labels over a DataLoader as NumPy arrays (trainer.py:82-93)."""
    model.eval()
    model.cuda()
    Pred, Real = [], []
    with torch.no_grad():
        for x, y in loader:
            Pred.append(torch.argmax(model(x.cuda(non_blocking=True)), dim=1).cpu())
            Real.append(y)
    return torch.cat(Pred).numpy(), torch.cat(Real).numpy()
