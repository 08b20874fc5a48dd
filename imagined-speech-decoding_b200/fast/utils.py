"""Host helpers the reference exports from ``fast.utils`` (src/fast/utils.py:17-114): colours, timers, seeding."""
import datetime
import os
import random
import string
import time

import numpy as np
import torch

_CODES = dict(bold=1, dim=2, italicized=3, underline=4, blink=5, inverse=7, gray=90, red=91, green=92, yellow=93,
              blue=94, magenta=95, cyan=96, white=97)


def _paint(code):
    def f(x):
        return f"\033[{code}m{x}\033[0m"
    return f


for _n, _c in _CODES.items():
    globals()[_n] = _paint(_c)
del _n, _c


def convert_to_number(value):
    """'12' -> 12, '1.5' -> 1.5, anything else unchanged (utils.py:36-40)."""
    try:
        return int(value) if value.isdigit() else float(value)
    except ValueError:
        return value


def find_available_path(folder_list):
    for folder in folder_list:
        if os.path.exists(folder):
            return folder
    raise FileNotFoundError('None of the given path exists' + str(folder_list))


def now(fmt="%Y-%m-%d_%H:%M:%S"):
    return str(datetime.datetime.now().strftime(fmt))


def random_string(length=10):
    return ''.join(random.choice(string.ascii_letters + string.digits) for _ in range(length))


class _Timer:
    def __enter__(self):
        self.t_start = time.time()
        self._begin()
        return self

    def __exit__(self, exc_type, exc_val, exc_tb):
        self.t_end = time.time()
        self.delta = self.t_end - self.t_start
        self.fps = 1 / self.delta if self.delta > 0 else float('inf')
        self._finish()

    def _begin(self):
        pass


class Tick(_Timer):
    """``with Tick('name'):`` prints the name, then the elapsed seconds (utils.py:62-78)."""

    def __init__(self, name='', silent=False):
        self.name, self.silent = name, silent

    def _begin(self):
        if not self.silent:
            print('%s ' % (self.name), end='', flush=True)

    def _finish(self):
        if not self.silent:
            print(yellow('[%.3fs]' % (self.delta)), flush=True)       # noqa: F821


class Tock(_Timer):
    """Inline timer (utils.py:81-98)."""

    def __init__(self, name=None, report_time=True):
        self.name = '' if name is None else name + ':'
        self.report_time = report_time

    def _finish(self):
        if self.report_time:
            print(yellow(self.name) + cyan('%.3fs' % (self.delta)), end=' ', flush=True)   # noqa: F821
        else:
            print(yellow('.'), end='', flush=True)                    # noqa: F821


def seed_all(seed):
    """Seed python / numpy / torch (CPU + every CUDA device) and pin the library determinism knobs (utils.py:104-114)."""
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed(seed)
        torch.cuda.manual_seed_all(seed)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    torch.set_float32_matmul_precision('medium')
