"""Signal-level operations: framing, spectra, deltas, CMVN.

Mirror of speech_feature_extraction/speechpy/processing.py (reference line numbers
are cited per function).  numpy arrays in, numpy arrays out; the arithmetic runs in
libspeechpy_b200.so on the GPU.  Integer frame geometry is computed on the host with
the reference's own expressions so that frame counts are bit-exact.
"""
import decimal
import math

import numpy as np

from .. import _lib

__all__ = ["round_half_up", "preemphasis", "stack_frames", "fft_spectrum", "power_spectrum",
           "log_power_spectrum", "derivative_extraction", "cmvn", "cmvnw"]


def round_half_up(number):
    """1.4 -> 1, 1.5 -> 2, 1.6 -> 2 (processing.py:38-42; unused by the reference)."""
    return int(decimal.Decimal(number).quantize(decimal.Decimal('1'), rounding=decimal.ROUND_HALF_UP))


_PRE_DTYPES = {np.dtype(np.float32): _lib.F32, np.dtype(np.float64): _lib.F64,
               np.dtype(np.int16): _lib.I16, np.dtype(np.int32): _lib.I32}


def preemphasis(signal, shift=1, cof=0.98):
    """signal - cof * np.roll(signal, shift), circular (processing.py:45-58), on the device
    (``spb_preemphasis``).  Result dtype as numpy gives it: float32 for a float32 signal,
    float64 for float64 and integer signals; any shape (np.roll flattens when axis is None)."""
    signal = np.asarray(signal)
    x = np.ascontiguousarray(signal)
    if x.dtype not in _PRE_DTYPES:
        x = x.astype(np.float32 if x.dtype == np.float16 else np.float64)
    out = np.empty(x.shape, dtype=np.float32 if x.dtype == np.float32 else np.float64)
    if x.size:
        ctx = _lib.context()
        _lib.check(ctx.lib.spb_preemphasis(ctx.handle, _lib.ptr(x), _PRE_DTYPES[x.dtype], x.size, int(shift),
                                           float(cof), _lib.ptr(out)), ctx.handle)
    return out


def _geometry(length_signal, sampling_frequency, frame_length, frame_stride):
    # processing.py:93-98 -- numpy half-even rounding of the float products
    frame_sample_length = int(np.round(sampling_frequency * frame_length))
    stride = float(np.round(sampling_frequency * frame_stride))
    return frame_sample_length, stride


def stack_frames(sig, sampling_frequency, frame_length=0.020, frame_stride=0.020,
                 filter=lambda x: np.ones((x,)), zero_padding=True):
    """Frame a 1-D signal into (num_frames, frame_len) float64 rows times a window.

    processing.py:61-139.  ``zero_padding=True`` gives ceil((N-L)/S) frames,
    ``False`` floor((N-L)/S); a negative count raises ValueError like numpy does in
    the reference.  ``filter`` is a callable L -> window, evaluated on the host.
    """
    s = "Signal dimention should be of the format of (N,) but it is %s instead"
    assert sig.ndim == 1, s % str(sig.shape)
    n = sig.shape[0]
    L, S = _geometry(n, sampling_frequency, frame_length, frame_stride)
    if S == 0:
        raise ZeroDivisionError("float division by zero")
    lib = _lib.load()
    nf = _lib.c_i64()
    _lib.check(lib.spb_num_frames(n, L, int(S), 1 if zero_padding else 0, _lib.C.byref(nf)))
    F = nf.value
    window = np.ascontiguousarray(np.asarray(filter(L), dtype=np.float64).reshape(-1))
    if window.shape[0] != L:
        raise ValueError("operands could not be broadcast together with shapes (%d,%d) (%d,%d)"
                         % (F, L, F, window.shape[0]))
    out = np.empty((F, L), dtype=np.float64)
    if F == 0 or L == 0:
        return out
    x = np.ascontiguousarray(sig, dtype=np.float64)
    win_ptr = None if np.all(window == 1.0) else _lib.ptr(window)
    ctx = _lib.context()
    _lib.check(lib.spb_stack_frames(ctx.handle, _lib.ptr(x), n, L, int(S), win_ptr, F, _lib.ptr(out)),
               ctx.handle)
    return out


def _spectrum(frames, fft_points, mode):
    frames = np.asarray(frames)
    lead = frames.shape[:-1]
    L = frames.shape[-1]
    x = np.ascontiguousarray(frames, dtype=np.float64).reshape(-1, L)
    F = x.shape[0]
    K = fft_points // 2 + 1
    out = np.empty((F, K), dtype=np.float64)
    if F:
        ctx = _lib.context()
        _lib.check(ctx.lib.spb_power_spectrum(ctx.handle, _lib.ptr(x), F, L, int(fft_points), mode,
                                               _lib.ptr(out)), ctx.handle)
    return out.reshape(lead + (K,))


def fft_spectrum(frames, fft_points=512):
    """|rfft(frames, n=fft_points)| along the last axis (processing.py:142-159)."""
    return _spectrum(frames, fft_points, 1)


def power_spectrum(frames, fft_points=512):
    """|rfft|^2 / fft_points (processing.py:162-174)."""
    return _spectrum(frames, fft_points, 0)


def log_power_spectrum(frames, fft_points=512, normalize=True):
    """10 log10(max(P, 1e-20)), minus the global maximum when ``normalize``
    (processing.py:177-198); power spectrum and float64 log epilogue both on the device."""
    return _spectrum(frames, fft_points, 3 if normalize else 2)


def derivative_extraction(feat, DeltaWindows):
    """The reference's delta operator, exactly as written (processing.py:201-236):
    sum_r r * feat[:, min(c + r, cols - 1)] / (2 sum_r r^2) along the FEATURE axis (the
    subtraction on processing.py:232 is a separate no-op statement)."""
    rows, cols = feat.shape
    x = np.ascontiguousarray(feat, dtype=np.float64)
    out = np.empty((rows, cols), dtype=np.float64)
    if rows and cols:
        ctx = _lib.context()
        _lib.check(ctx.lib.spb_derivative(ctx.handle, _lib.ptr(x), rows, cols, int(DeltaWindows),
                                          _lib.ptr(out)), ctx.handle)
    if feat.dtype == np.float32:
        return out.astype(np.float32)
    return out


def cmvn(vec, variance_normalization=False):
    """Global mean (and variance) normalisation, float64 (processing.py:239-271)."""
    rows, cols = vec.shape
    x = np.ascontiguousarray(vec, dtype=np.float64)
    out = np.empty((rows, cols), dtype=np.float64)
    if rows and cols:
        ctx = _lib.context()
        _lib.check(ctx.lib.spb_cmvn(ctx.handle, _lib.ptr(x), rows, cols, 1 if variance_normalization else 0,
                                    _lib.ptr(out)), ctx.handle)
    return out


def cmvnw(vec, win_size=301, variance_normalization=False):
    """Sliding-window mean (and variance) normalisation; float32 output
    (processing.py:274-327)."""
    rows, cols = vec.shape
    assert isinstance(win_size, int), "Size must be of type 'int'!"
    assert win_size % 2 == 1, "Windows size must be odd!"
    x = np.ascontiguousarray(vec, dtype=np.float64)
    out = np.empty((rows, cols), dtype=np.float32)
    if rows and cols:
        ctx = _lib.context()
        _lib.check(ctx.lib.spb_cmvnw(ctx.handle, _lib.ptr(x), rows, cols, win_size,
                                     1 if variance_normalization else 0, _lib.ptr(out)), ctx.handle)
    return out
