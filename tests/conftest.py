"""Shared fixtures.  ``-m "not gpu"`` runs on a CPU-only box; ``-m gpu`` needs a B200 and the built libfastb200.so."""
import os
import sys
import types

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "imagined-speech-decoding_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a) and libfastb200.so")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


class Cfg(types.SimpleNamespace):
    """Attribute bag with the 12 fields FAST.__init__ reads (reference fast.py:219-230)."""


def make_prod_cfg(**over):
    from fast.data import Electrodes, Zones
    kw = dict(electrodes=Electrodes, zone_dict=Zones, dim_cnn=32, dim_token=32, seq_len=800, window_len=250,
              slide_step=125, head="Conv4Layers", n_classes=5, num_layers=4, num_heads=8, dropout=0.1)
    kw.update(over)
    return Cfg(**kw)


def make_small_cfg(**over):
    electrodes = ['Fp1', 'Fp2', 'F3', 'F4', 'C3', 'C4', 'O1', 'O2']
    zones = {'Frontal': ['Fp1', 'Fp2', 'F3', 'F4'], 'Central': ['C3', 'C4'], 'Occipital': ['O1', 'O2']}
    kw = dict(electrodes=electrodes, zone_dict=zones, dim_cnn=16, dim_token=16, seq_len=500, window_len=250,
              slide_step=125, head="Conv4Layers", n_classes=3, num_layers=1, num_heads=4, dropout=0.0)
    kw.update(over)
    return Cfg(**kw)


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "fast_golden.npz"))


@pytest.fixture(scope="session")
def filter_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "filter_golden.npz"))


@pytest.fixture(scope="session")
def prod_cfg():
    return make_prod_cfg()


@pytest.fixture(scope="session")
def small_cfg():
    return make_small_cfg()


def small_state_dict(golden):
    return {k[len("small/sd/"):]: torch.from_numpy(golden[k]) for k in golden.files if k.startswith("small/sd/")}


def probe(n, seed=12345):
    return np.random.default_rng(seed).standard_normal(n)


def fingerprint(t):
    """Same 4-number fingerprint tests/golden/make_golden.py stores for big tensors."""
    a = t.detach().double().cpu().numpy().ravel()
    return np.array([a.sum(), np.abs(a).sum(), float(a @ probe(a.size)), np.abs(a).max()])


def rel_err(a, b):
    """max|a-b| / max|b|  -- the parity norm of SURVEY.md section 8d."""
    def t(v):
        return v.detach().double().cpu() if isinstance(v, torch.Tensor) else torch.from_numpy(np.array(v, dtype=np.float64))
    a, b = t(a), t(b)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
