"""Training utilities for FAST (drop-in for the reference's ``fast.train``, src/fast/train/__init__.py:5-13)."""
from .trainer import EEG_Encoder_Module, cosine_scheduler, inference_on_loader
from .callbacks import HistoryCallback
from .fused import FusedTrainer
from .dp import DataParallelTrainer
from . import cv

__all__ = [
    "EEG_Encoder_Module",
    "cosine_scheduler",
    "inference_on_loader",
    "HistoryCallback",
    "FusedTrainer",      # additive: device-resident, CUDA-graphed training step
    "DataParallelTrainer",   # additive: batch data-parallel step, flat-gradient all-reduce fused into AdamW over NVLink
    "cv",                # additive: subject x fold jobs sharded over GPUs (train_fold / run_cv_jobs / finalize_cv)
]
