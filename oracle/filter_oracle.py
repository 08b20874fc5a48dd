"""CPU oracle for the band-pass / resample sub-path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

PARITY UNPINNED BY THE REFERENCE: kidusabe1/Imagined-Speech-Decoding has no Butterworth /
sosfiltfilt / resample code (its only band-pass is MNE's FIR in
notebooks/svm_baseline.ipynb:238-239).  BASELINE.json's north_star (c) specifies the op, so
the arithmetic is pinned against the third-party library the reference declares
(``scipy>=1.10.0``, pyproject.toml:18; installed here: scipy 1.18.1).  This file restates
scipy's published algorithm in plain NumPy float64:

* ``sosfiltfilt``  -- scipy/signal/_signaltools.py:5091-5190 (odd extension ``_arraytools.py:57-75``,
  ``sosfilt_zi`` :4464-4490, ``lfilter_zi`` :4330-4400, DF-II-transposed section loop ``_sosfilt.pyx``)
* ``resample_poly``-- scipy/signal/_signaltools.py:3884-4100 + ``upfirdn`` (zero-stuffed FIR, keep every
  ``down``-th sample), default Kaiser(5.0) design, ``padtype='constant'``.

``tests/test_oracle.py`` checks it against scipy itself and against
``tests/golden/filter_golden.npz`` (scipy outputs generated in the build container).
"""
from __future__ import annotations

import numpy as np


def sos_steady_state(sos: np.ndarray) -> np.ndarray:
    """Per-section DF-II-transposed state for a unit step input, (n_sections, 2)."""
    sos = np.asarray(sos, dtype=np.float64)
    zi = np.empty((sos.shape[0], 2))
    scale = 1.0
    for s in range(sos.shape[0]):
        b = sos[s, :3] / sos[s, 3]
        a = sos[s, 3:] / sos[s, 3]
        gain = b.sum() / a.sum()
        d = b - gain * a
        # reversed cumulative sum of d, dropping the first entry
        zi[s, 0] = scale * (d[1] + d[2])
        zi[s, 1] = scale * d[2]
        scale *= gain
    return zi


def filtfilt_padlen(sos: np.ndarray) -> int:
    sos = np.asarray(sos)
    ntaps = 2 * sos.shape[0] + 1
    ntaps -= min(int((sos[:, 2] == 0).sum()), int((sos[:, 5] == 0).sum()))
    return 3 * ntaps


def _cascade(sos: np.ndarray, x: np.ndarray, z: np.ndarray) -> np.ndarray:
    """Run the cascade along the last axis.  x (..., n); z (n_sections, ..., 2) modified in place."""
    y = np.array(x, dtype=np.float64, copy=True)
    n = y.shape[-1]
    for t in range(n):
        v = y[..., t]
        for s in range(sos.shape[0]):
            b0, b1, b2, _, a1, a2 = sos[s]
            out = b0 * v + z[s, ..., 0]
            z[s, ..., 0] = b1 * v - a1 * out + z[s, ..., 1]
            z[s, ..., 1] = b2 * v - a2 * out
            v = out
        y[..., t] = v
    return y


def sosfiltfilt(sos: np.ndarray, x: np.ndarray, padlen: int | None = None) -> np.ndarray:
    """Zero-phase forward-backward cascade along the last axis, odd extension, float64 out."""
    sos = np.asarray(sos, dtype=np.float64)
    sos = sos / sos[:, 3:4]
    x = np.asarray(x)
    edge = filtfilt_padlen(sos) if padlen is None else padlen
    if x.shape[-1] <= edge:
        raise ValueError("input shorter than padlen")
    if edge > 0:
        # NOTE: scipy builds the odd extension in the INPUT dtype (float32 data -> one float32
        # rounding per extended sample) and only then promotes to float64 inside sosfilt.
        left = 2 * x[..., :1] - x[..., edge:0:-1]
        right = 2 * x[..., -1:] - x[..., -2:-(edge + 2):-1]
        ext = np.concatenate([left, x, right], axis=-1)
    else:
        ext = x
    ext = ext.astype(np.float64)
    zi = sos_steady_state(sos)                                   # (S,2)
    shape = (sos.shape[0],) + ext.shape[:-1] + (2,)
    z = np.broadcast_to(zi.reshape((sos.shape[0],) + (1,) * (ext.ndim - 1) + (2,)), shape) * ext[..., :1][None]
    y = _cascade(sos, ext, np.array(z))
    y = y[..., ::-1]
    z = np.broadcast_to(zi.reshape((sos.shape[0],) + (1,) * (ext.ndim - 1) + (2,)), shape) * y[..., :1][None]
    y = _cascade(sos, y, np.array(z))[..., ::-1]
    return y[..., edge:-edge] if edge > 0 else y


def kaiser_lowpass(up: int, down: int, beta: float = 5.0) -> np.ndarray:
    """scipy.signal.firwin(2*half+1, 1/max(up,down), window=('kaiser', beta)) * up, float64."""
    max_rate = max(up, down)
    half = 10 * max_rate
    n = 2 * half + 1
    fc = 1.0 / max_rate
    m = np.arange(n) - half
    h = fc * np.sinc(fc * m)
    h = h * np.kaiser(n, beta)
    h = h / h.sum()                     # unity gain at DC
    return h * up


def resample_poly(x: np.ndarray, up: int, down: int, h: np.ndarray | None = None) -> np.ndarray:
    """Rational-rate polyphase resampling along the last axis (zero padding at the edges)."""
    g = np.gcd(up, down)
    up, down = up // g, down // g
    x = np.asarray(x, dtype=np.float64)
    if up == down == 1:
        return x.copy()
    n_in = x.shape[-1]
    n_out = -(-n_in * up // down)
    if h is None:
        h = kaiser_lowpass(up, down)
    half = (len(h) - 1) // 2
    n_pre_pad = down - half % down
    n_pre_remove = (half + n_pre_pad) // down
    hp = np.concatenate([np.zeros(n_pre_pad), h])
    # y_full[m] = sum_i x[i] * hp[m*down - i*up]; taps past the end of hp read as zero, which is what
    # scipy's n_post_pad loop guarantees too.
    y = np.zeros(x.shape[:-1] + (n_out,))
    L = len(hp)
    for j in range(n_out):
        pos = (j + n_pre_remove) * down
        i_hi = min(pos // up, n_in - 1)
        i_lo = max(0, -(-(pos - L + 1) // up))
        for i in range(i_lo, i_hi + 1):
            y[..., j] += x[..., i] * hp[pos - i * up]
    return y
