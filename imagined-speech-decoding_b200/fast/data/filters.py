"""Batched zero-phase band-pass and rational resampling of EEG trials on the GPU (BASELINE.json north_star (c)).

The reference repository has no filtering code; these functions are additive to ``fast.data``
and follow ``scipy.signal.sosfiltfilt`` / ``scipy.signal.resample_poly`` semantics along the last axis.  The filter
*design* (a handful of coefficients) is host arithmetic; every sample is touched only by the CUDA kernels
``fastb200_sosfiltfilt`` / ``fastb200_resample_poly`` / ``fastb200_sosfiltfilt_resample`` (the chain in one kernel, the
filtered float64 row never leaving shared memory) of include/fastb200.h.  There is no CPU fallback.
"""
import ctypes as C
import math

import numpy as np
import torch

from .. import _native as N


def bandpass_sos(order=4, low=4.0, high=40.0, fs=250.0):
    """Butterworth band-pass as second-order sections, (order, 6) float64 --
    ``scipy.signal.butter(order, [low, high], 'bandpass', fs=fs, output='sos')``."""
    from scipy import signal
    return np.ascontiguousarray(signal.butter(order, [low, high], 'bandpass', fs=fs, output='sos'), dtype=np.float64)


def default_padlen(sos):
    """scipy.signal.sosfiltfilt's default edge length: 3 * (2*n_sections + 1 - min(#zero b2, #zero a2))."""
    sos = np.asarray(sos)
    ntaps = 2 * sos.shape[0] + 1
    ntaps -= min(int((sos[:, 2] == 0).sum()), int((sos[:, 5] == 0).sum()))
    return 3 * ntaps


def _as_rows(x):
    if not isinstance(x, torch.Tensor):
        raise TypeError("expected a CUDA tensor")
    N.require_cuda(x, "x")
    if x.dtype not in (torch.float32, torch.float64):
        x = x.float()
    return x.contiguous()


def _out_dtype(x, out_dtype):
    if out_dtype is None:
        return x.dtype
    if out_dtype not in (torch.float32, torch.float64):
        raise TypeError("out_dtype must be torch.float32 or torch.float64")
    return out_dtype


def sosfiltfilt(sos, x, padlen=None, out_dtype=None):
    """Zero-phase forward-backward IIR cascade along the last axis of a CUDA tensor (odd extension of ``padlen``
    samples, steady-state initial conditions), arithmetic in float64.  ``out_dtype`` defaults to ``x.dtype``
    (scipy returns float64; pass ``torch.float64`` for that)."""
    sos = np.ascontiguousarray(sos, dtype=np.float64)
    if sos.ndim != 2 or sos.shape[1] != 6:
        raise ValueError("sos must be shape (n_sections, 6)")
    x = _as_rows(x)
    T = x.shape[-1]
    edge = default_padlen(sos) if padlen is None else int(padlen)
    if T <= edge:
        raise ValueError(f"The length of the input vector x must be greater than padlen, which is {edge}.")
    y = torch.empty(x.shape, device=x.device, dtype=_out_dtype(x, out_dtype))
    rows = x.numel() // T if T else 0
    with torch.cuda.device(x.device):
        N.check(N.lib().fastb200_sosfiltfilt(N.ptr(x), x.element_size(), N.ptr(y), y.element_size(), rows, T,
                                             sos.ctypes.data_as(C.POINTER(C.c_double)), sos.shape[0], edge,
                                             N.stream_ptr(x.device)), "sosfiltfilt")
    return y


def _kaiser_lowpass(up, down, beta=5.0):
    """The default anti-aliasing FIR of scipy.signal.resample_poly: firwin(20*max(up,down)+1, 1/max(up,down),
    window=('kaiser', 5.0)), scaled by ``up``."""
    max_rate = max(up, down)
    half = 10 * max_rate
    m = np.arange(-half, half + 1, dtype=np.float64)
    fc = 1.0 / max_rate
    h = fc * np.sinc(fc * m) * np.kaiser(2 * half + 1, beta)
    return h / h.sum() * up


_FIR_CACHE = {}
_STREAMS = {}


def _fir_plan(device, up, down, window):
    """(h_host with scipy's leading zero pad, h_dev, n_pre_remove) for the reduced ratio; cached per device."""
    key = (device, up, down, None if window is None else np.asarray(window, dtype=np.float64).tobytes())
    plan = _FIR_CACHE.get(key)
    if plan is None:
        h = _kaiser_lowpass(up, down) if window is None else np.asarray(window, dtype=np.float64) * up
        half = (len(h) - 1) // 2
        n_pre_pad = down - half % down
        n_pre_remove = (half + n_pre_pad) // down
        h_host = np.ascontiguousarray(np.concatenate([np.zeros(n_pre_pad), h]))
        plan = (h_host, torch.from_numpy(h_host).to(device), n_pre_remove)
        _FIR_CACHE[key] = plan
    return plan


def _reduce_ratio(up, down):
    up, down = int(up), int(down)
    if up < 1 or down < 1:
        raise ValueError("up and down must be >= 1")
    g = math.gcd(up, down)
    return up // g, down // g


def resample_poly(x, up, down, window=None, out_dtype=None):
    """Polyphase resampling by ``up/down`` along the last axis of a CUDA tensor with zero-padded edges
    (``scipy.signal.resample_poly(x, up, down, axis=-1)`` defaults).  ``window``: optional 1-D FIR (already designed
    for the reduced up/down, unity DC gain before the ``up`` scaling)."""
    up, down = _reduce_ratio(up, down)
    x = _as_rows(x)
    odt = _out_dtype(x, out_dtype)
    if up == down == 1:
        return x.to(odt, copy=True)
    T_in = x.shape[-1]
    T_out = -(-T_in * up // down)
    h_host, h_dev, n_pre_remove = _fir_plan(x.device, up, down, window)
    y = torch.empty(*x.shape[:-1], T_out, device=x.device, dtype=odt)
    rows = x.numel() // T_in if T_in else 0
    with torch.cuda.device(x.device):
        N.check(N.lib().fastb200_resample_poly(N.ptr(x), x.element_size(), N.ptr(y), y.element_size(), rows, T_in, T_out,
                                               N.ptr(h_dev), h_dev.numel(), up, down, n_pre_remove,
                                               h_host.ctypes.data_as(C.POINTER(C.c_double)),
                                               N.stream_ptr(x.device)), "resample_poly")
    return y


def sosfiltfilt_resample(sos, x, up, down, padlen=None, window=None, out_dtype=None):
    """``resample_poly(sosfiltfilt(sos, x), up, down)`` with a float64 intermediate.  Where the fused kernel supports
    the shape (4/5 with the default FIR) the filtered row never leaves shared memory; other shapes run the two
    kernels back to back.  Results are those of the two-call chain either way."""
    sos = np.ascontiguousarray(sos, dtype=np.float64)
    if sos.ndim != 2 or sos.shape[1] != 6:
        raise ValueError("sos must be shape (n_sections, 6)")
    up, down = _reduce_ratio(up, down)
    x = _as_rows(x)
    odt = _out_dtype(x, out_dtype)
    T = x.shape[-1]
    edge = default_padlen(sos) if padlen is None else int(padlen)
    if T <= edge:
        raise ValueError(f"The length of the input vector x must be greater than padlen, which is {edge}.")
    T_out = -(-T * up // down)
    L = N.lib()
    if not (up == down == 1):
        h_host, _, n_pre_remove = _fir_plan(x.device, up, down, window)
        hp = h_host.ctypes.data_as(C.POINTER(C.c_double))
        if sos.shape[0] <= 4 and L.fastb200_sosfiltfilt_resample_supported(T, sos.shape[0], edge, T_out, hp, len(h_host), up, down,
                                                                          n_pre_remove):
            y = torch.empty(*x.shape[:-1], T_out, device=x.device, dtype=odt)
            rows = x.numel() // T
            with torch.cuda.device(x.device):
                N.check(L.fastb200_sosfiltfilt_resample(N.ptr(x), x.element_size(), N.ptr(y), y.element_size(), rows, T,
                                                        sos.ctypes.data_as(C.POINTER(C.c_double)), sos.shape[0], edge, T_out,
                                                        hp, len(h_host), up, down, n_pre_remove, N.stream_ptr(x.device)),
                        "sosfiltfilt_resample")
            return y
    return resample_poly(sosfiltfilt(sos, x, padlen=edge, out_dtype=torch.float64), up, down, window=window, out_dtype=odt)


def bandpass_resample(x, fs=250.0, low=4.0, high=40.0, order=4, up=4, down=5, out_dtype=torch.float32):
    """The preprocessing chain of BASELINE.json config 3: order-``order`` Butterworth ``low``-``high`` Hz zero-phase
    band-pass, then ``up/down`` polyphase resampling (250 -> 200 Hz by default).  The intermediate stays float64
    (in shared memory where the fused kernel applies)."""
    return sosfiltfilt_resample(bandpass_sos(order, low, high, fs), x, up, down, out_dtype=out_dtype)


def bandpass_resample_host(blocks, fs=250.0, low=4.0, high=40.0, order=4, up=4, down=5, device=None, depth=2):
    """Stream host blocks through ``bandpass_resample``: ``blocks`` is an iterable of host arrays / tensors (n, C, T)
    float32 (pinned memory makes the copies asynchronous); block k+1 is copied in and block k-1 copied out while block k
    is filtered, on ``depth`` CUDA streams (PCIe is full duplex, and the copies, not the kernel, are what a host-to-host
    preprocessing run waits for).  Yields pinned float32 host tensors (n, C, T_out) in input order."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    sos = bandpass_sos(order, low, high, fs)
    # the side streams are kept per device: the caching allocator pools blocks per stream, so fresh streams would cudaMalloc
    # their staging tensors on every call
    pool = _STREAMS.setdefault(dev, [])
    while len(pool) < max(1, int(depth)):
        pool.append(torch.cuda.Stream(dev))
    streams = pool[:max(1, int(depth))]
    cur = torch.cuda.current_stream(dev)
    for s in streams:
        s.wait_stream(cur)
    inflight = []                                                    # (event, pinned output, tensors kept alive)
    for k, blk in enumerate(blocks):
        s = streams[k % len(streams)]
        xh = torch.as_tensor(blk)
        if xh.dtype != torch.float32:
            xh = xh.float()
        with torch.cuda.stream(s):
            xd = xh.to(dev, non_blocking=True)
            y = sosfiltfilt_resample(sos, xd, up, down, out_dtype=torch.float32)
            out = torch.empty(y.shape, dtype=torch.float32, pin_memory=True)
            out.copy_(y, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(s)
        inflight.append((ev, out, (xh, xd, y)))
        if len(inflight) >= len(streams):
            ev0, out0, _ = inflight.pop(0)
            ev0.synchronize()
            yield out0
    for ev0, out0, _ in inflight:
        ev0.synchronize()
        yield out0
