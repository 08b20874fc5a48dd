"""FAST: Functional Areas Spatio-Temporal Transformer for EEG classification -- B200-native hot path.

Drop-in for the reference package ``fast`` (kidusabe1/Imagined-Speech-Decoding ``src/fast/__init__.py:7-10``): put the
directory that holds this package on ``PYTHONPATH`` ahead of the reference's ``src`` and the reference's
``scripts/train_fast.py`` / ``scripts/preprocess.py`` pick up the CUDA implementation unchanged.
"""
__version__ = "1.0.0"

from .models.fast import FAST
from .utils import seed_all, green, yellow, Tick, Tock

__all__ = ["FAST", "seed_all", "green", "yellow", "Tick", "Tock"]
