"""Per-epoch history capture (reference src/fast/train/callbacks.py:8-28)."""
try:                                           # Lightning is optional on this image
    from lightning.pytorch.callbacks import Callback
except Exception:                              # pragma: no cover - depends on the environment
    class Callback:                            # the hook protocol is duck-typed; a bare base keeps the class usable
        pass


class HistoryCallback(Callback):
    """Copies train/val loss and accuracy out of ``trainer.callback_metrics`` at the end of every epoch."""

    _TRAIN = (('loss', 'train_loss'), ('acc', 'train_acc'))
    _VAL = (('val_loss', 'val_loss'), ('val_acc', 'val_acc'))

    def __init__(self):
        self.history = {'loss': [], 'acc': [], 'val_loss': [], 'val_acc': []}

    def _grab(self, trainer, pairs):
        for dst, src in pairs:
            value = trainer.callback_metrics.get(src)
            if value is not None:
                self.history[dst].append(value.item())

    def on_train_epoch_end(self, trainer, pl_module):
        self._grab(trainer, self._TRAIN)

    def on_validation_epoch_end(self, trainer, pl_module):
        self._grab(trainer, self._VAL)
