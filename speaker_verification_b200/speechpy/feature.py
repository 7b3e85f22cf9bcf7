"""Mel filterbank features: filterbanks, mfe, lmfe, mfcc, delta cube.

Mirror of speech_feature_extraction/speechpy/feature.py.  The filterbank (integer FFT-bin
edges + triangle weights) is built on the host in float64 with the reference's
expression order and cached per parameter tuple; everything per-signal runs in
libspeechpy_b200.so.
"""
from __future__ import division

import functools

import numpy as np

from .. import _lib
from . import functions, processing

__all__ = ["filterbanks", "mfcc", "mfe", "lmfe", "extract_derivative_feature"]


def _edges(num_filter, coefficients, sampling_freq, low_freq, high_freq):
    # feature.py:65-82 -- same numpy calls in the same order (bit-exact integer edges)
    mels = np.linspace(functions.frequency_to_mel(low_freq),
                       functions.frequency_to_mel(high_freq), num_filter + 2)
    hertz = functions.mel_to_frequency(mels)
    return (np.floor((coefficients + 1) * hertz / sampling_freq)).astype(int)


@functools.lru_cache(maxsize=64)
def _bank_cached(num_filter, coefficients, sampling_freq, low_freq, high_freq):
    edges = _edges(num_filter, coefficients, sampling_freq, low_freq, high_freq)
    bank = np.zeros([num_filter, coefficients])
    for i in range(num_filter):
        left, middle, right = int(edges[i]), int(edges[i + 1]), int(edges[i + 2])
        z = np.linspace(left, right, num=right - left + 1)
        bank[i, left:right + 1] = functions.triangle(z, left=left, middle=middle, right=right)
    bank.setflags(write=False)
    return bank


def filterbanks(num_filter, coefficients, sampling_freq, low_freq=None, high_freq=None):
    """Dense (num_filter, coefficients) mel filterbank, float64 (feature.py:33-99).

    Quirks kept: ``low_freq or 300`` (0/None -> 300 Hz, feature.py:56) and the
    ``coefficients + 1`` multiplier in the bin mapping (feature.py:77-82).
    """
    high_freq = high_freq or sampling_freq / 2
    low_freq = low_freq or 300
    s = "High frequency cannot be greater than half of the sampling frequency!"
    assert high_freq <= sampling_freq / 2, s
    assert low_freq >= 0, "low frequency cannot be less than zero!"
    return _bank_cached(int(num_filter), int(coefficients), sampling_freq, low_freq, high_freq).copy()


def _prepare(signal):
    """signal.astype(float) semantics (feature.py:182) with the narrowest upload dtype."""
    signal = np.asarray(signal)
    s = "Signal dimention should be of the format of (N,) but it is %s instead"
    assert signal.ndim == 1, s % str(signal.shape)
    if signal.dtype == np.float32:
        return np.ascontiguousarray(signal), _lib.F32
    if signal.dtype == np.int16:
        return np.ascontiguousarray(signal), _lib.I16
    if signal.dtype == np.int32:
        return np.ascontiguousarray(signal), _lib.I32
    return np.ascontiguousarray(signal.astype(float)), _lib.F64


def _mfe_like(which, signal, sampling_frequency, frame_length, frame_stride, num_filters,
              fft_length, low_frequency, high_frequency, num_cepstral=0, dc_elimination=True):
    x, dtype = _prepare(signal)
    n = x.shape[0]
    L, S = processing._geometry(n, sampling_frequency, frame_length, frame_stride)
    if S == 0:
        raise ZeroDivisionError("float division by zero")
    lib = _lib.load()
    nf = _lib.c_i64()
    _lib.check(lib.spb_num_frames(n, L, int(S), 0, _lib.C.byref(nf)))  # zero_padding=False (feature.py:193)
    F = nf.value
    high_frequency = high_frequency or sampling_frequency / 2
    coefficients = int(fft_length) // 2 + 1
    bank = filterbanks(num_filters, coefficients, sampling_frequency, low_frequency, high_frequency)
    C = min(int(num_cepstral), int(num_filters))
    ctx = _lib.context()
    plan = ctx.plan(L, int(S), int(fft_length), int(num_filters), C, 1 if dc_elimination else 0, bank)
    if which == "mfcc":
        out = np.empty((F, C), dtype=np.float64)
        if F:
            _lib.check(lib.spb_mfcc(ctx.handle, plan, _lib.ptr(x), dtype, n, _lib.ptr(out)), ctx.handle)
        return out
    feats = np.empty((F, int(num_filters)), dtype=np.float64)
    if which == "lmfe":
        if F:
            _lib.check(lib.spb_lmfe(ctx.handle, plan, _lib.ptr(x), dtype, n, _lib.ptr(feats)), ctx.handle)
        return feats
    energy = np.empty((F,), dtype=np.float64)
    if F:
        _lib.check(lib.spb_mfe(ctx.handle, plan, _lib.ptr(x), dtype, n, _lib.ptr(feats), _lib.ptr(energy)),
                   ctx.handle)
    return feats, energy


def mfcc(signal, sampling_frequency, frame_length=0.020, frame_stride=0.01, num_cepstral=13,
         num_filters=40, fft_length=512, low_frequency=0, high_frequency=None, dc_elimination=True):
    """MFCC (feature.py:102-153): DCT-II (ortho) of ln(mfe), first ``num_cepstral``
    coefficients, c0 replaced by ln(frame energy) when ``dc_elimination``."""
    out = _mfe_like("mfcc", signal, sampling_frequency, frame_length, frame_stride, num_filters,
                    fft_length, low_frequency, high_frequency, num_cepstral, dc_elimination)
    if len(out) == 0:
        return np.empty((0, num_cepstral))
    return out


def mfe(signal, sampling_frequency, frame_length=0.020, frame_stride=0.01, num_filters=40,
        fft_length=512, low_frequency=0, high_frequency=None):
    """Mel filterbank energies and per-frame energy (feature.py:156-219)."""
    return _mfe_like("mfe", signal, sampling_frequency, frame_length, frame_stride, num_filters,
                     fft_length, low_frequency, high_frequency)


def lmfe(signal, sampling_frequency, frame_length=0.020, frame_stride=0.01, num_filters=40,
         fft_length=512, low_frequency=0, high_frequency=None):
    """Natural log of the mel filterbank energies (feature.py:222-258)."""
    return _mfe_like("lmfe", signal, sampling_frequency, frame_length, frame_stride, num_filters,
                     fft_length, low_frequency, high_frequency)


def extract_derivative_feature(feature):
    """(N, M) -> (N, M, 3): static, first and second "derivative" (feature.py:261-282)."""
    rows, cols = feature.shape
    x = np.ascontiguousarray(feature, dtype=np.float64)
    out = np.empty((rows, cols, 3), dtype=np.float64)
    if rows and cols:
        ctx = _lib.context()
        _lib.check(ctx.lib.spb_delta_cube(ctx.handle, _lib.ptr(x), rows, cols, _lib.ptr(out)), ctx.handle)
    if feature.dtype == np.float32:
        return out.astype(np.float32)
    return out
