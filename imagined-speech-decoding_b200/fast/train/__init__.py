"""Training utilities for FAST (drop-in for the reference's ``fast.train``, src/fast/train/__init__.py:5-13)."""
from .trainer import EEG_Encoder_Module, cosine_scheduler, inference_on_loader
from .callbacks import HistoryCallback
from .fused import FusedTrainer

__all__ = [
    "EEG_Encoder_Module",
    "cosine_scheduler",
    "inference_on_loader",
    "HistoryCallback",
    "FusedTrainer",      # additive: device-resident, CUDA-graphed training step
]
