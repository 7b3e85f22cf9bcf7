"""CPU oracle for the speechpy feature path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

This module is a float64 numpy restatement of the algorithm implemented by the
reference's bundled speechpy (``speech_feature_extraction/speechpy/*.py``;
abbreviated ``sfe/`` below).  It exists only so that the CUDA path can be
checked on machines where ``/root/reference`` is absent (the GPU box).  Only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it; nothing under
``speaker_verification_b200/`` does.

Parity status: PINNED against the reference itself.  ``oracle/make_golden.py``
imports the unmodified reference from ``/root/reference`` (with the one-line
``np.lib.pad = np.pad`` shim numpy>=2 needs, sfe/speechpy/processing.py:223,301,314),
checks every function below against it on seeded inputs (bit-exact for frame
counts/indices/filterbank edges, <=1e-12 elsewhere) and writes the golden
vectors in ``tests/golden/``; ``tests/test_oracle.py`` re-checks the oracle
against those vectors without needing the reference.  The reference's own
tests pin only shapes, API presence and the cmvn mean/std property
(sfe/tests/test_speechpy.py:24-106); those are reproduced in
``tests/test_gpu_features.py::test_reference_test_suite_semantics`` and, verbatim, by ``test_reference_test_file_runs_unchanged_against_the_dropin``.

The loop structure of the reference (one numpy reduction per frame in cmvnw,
filterbank rebuilt on every mfe call, index-matrix gather in stack_frames) is
kept on purpose: this file is also the "port" timed as the CPU baseline, and a
restructured algorithm would not be the reference's cost profile.
"""
from __future__ import annotations

import math

import numpy as np
from scipy.fftpack import dct as _dct

FLOAT_EPS = float(np.finfo(float).eps)  # sfe/speechpy/functions.py:62
CMVN_EPS = 2.0 ** -30  # sfe/speechpy/processing.py:253,292


# ----------------------------------------------------------------------------
# functions.py
# ----------------------------------------------------------------------------
def frequency_to_mel(f):
    """Hz -> mel, 1127*ln(1+f/700).  sfe/speechpy/functions.py:26-32."""
    return 1127 * np.log(1 + f / 700.0)


def mel_to_frequency(mel):
    """mel -> Hz, 700*(exp(mel/1127)-1).  sfe/speechpy/functions.py:35-41."""
    return 700 * (np.exp(mel / 1127.0) - 1)


def triangle(x, left, middle, right):
    """Triangular window on sample points x.  sfe/speechpy/functions.py:44-52.

    Rising edge on (left, middle], falling edge on [middle, right); the falling
    assignment is applied second so x == middle evaluates the falling formula.
    """
    x = np.asarray(x, dtype=float)
    y = np.zeros(x.shape)
    up = (x > left) & (x <= middle)
    y[up] = (x[up] - left) / (middle - left)
    down = (x >= middle) & (x < right)
    y[down] = (right - x[down]) / (right - middle)
    return y


def zero_handling(x):
    """Exact zeros -> float64 eps.  sfe/speechpy/functions.py:55-62."""
    return np.where(x == 0, FLOAT_EPS, x)


# ----------------------------------------------------------------------------
# processing.py
# ----------------------------------------------------------------------------
def preemphasis(signal, shift=1, cof=0.98):
    """x - cof*roll(x, shift) (circular).  sfe/speechpy/processing.py:45-58."""
    return signal - cof * np.roll(signal, shift)


def frame_geometry(num_samples, sampling_frequency, frame_length, frame_stride,
                   zero_padding):
    """(L, S, F) exactly as sfe/speechpy/processing.py:93-116 derives them.

    L = int(round(fs*frame_length)); S = float(round(fs*frame_stride)) (numpy
    half-even rounding); F = ceil((N-L)/S) with zero padding, floor((N-L)/S)
    without.  F may be negative (N < L); the caller then fails like the
    reference does.
    """
    L = int(np.round(sampling_frequency * frame_length))
    S = float(np.round(sampling_frequency * frame_stride))
    if zero_padding:
        F = int(math.ceil((num_samples - L) / S))
    else:
        F = int(math.floor((num_samples - L) / S))
    return L, S, F


def frame_indices(L, S, F):
    """int32 (F, L) gather matrix idx[f, j] = j + f*S.  processing.py:123-131."""
    cols = np.arange(0, L)[None, :]
    rows = np.arange(0, F * S, S)[:, None]
    return np.array(cols + rows, dtype=np.int32).reshape(F, L)


def stack_frames(sig, sampling_frequency, frame_length=0.020, frame_stride=0.020,
                 filter=lambda x: np.ones((x,)), zero_padding=True):
    """Overlapping frames times a window.  sfe/speechpy/processing.py:61-139."""
    assert sig.ndim == 1, (
        "Signal dimention should be of the format of (N,) but it is %s instead"
        % str(sig.shape))
    n = sig.shape[0]
    L, S, F = frame_geometry(n, sampling_frequency, frame_length, frame_stride,
                             zero_padding)
    if zero_padding:
        # processing.py:108-110 -- np.zeros raises ValueError on a negative count
        total = int(F * S + L)
        padded = np.concatenate((sig, np.zeros((total - n,))))
    else:
        # processing.py:119-120
        total = int((F - 1) * S + L)
        padded = sig[0:total]
    # np.tile with a negative repeat count raises ValueError (processing.py:123)
    if F < 0:
        raise ValueError("negative dimensions are not allowed")
    idx = frame_indices(L, S, F)
    frames = padded[idx]
    win = np.tile(filter(L), (F, 1))
    return frames * win


def fft_spectrum(frames, fft_points=512):
    """|rfft(frames, n=fft_points)|.  sfe/speechpy/processing.py:142-159."""
    return np.absolute(np.fft.rfft(frames, n=fft_points, axis=-1, norm=None))


def power_spectrum(frames, fft_points=512):
    """|X|^2 / NFFT.  sfe/speechpy/processing.py:162-174."""
    return 1.0 / fft_points * np.square(fft_spectrum(frames, fft_points))


def log_power_spectrum(frames, fft_points=512, normalize=True):
    """10 log10(max(P, 1e-20)) [- global max].  processing.py:177-198."""
    p = power_spectrum(frames, fft_points)
    p[p <= 1e-20] = 1e-20
    lp = 10 * np.log10(p)
    return lp - np.max(lp) if normalize else lp


def derivative_extraction(feat, DeltaWindows):
    """'Delta' features exactly as the reference computes them.

    sfe/speechpy/processing.py:201-236.  The subtraction on processing.py:232 is
    a separate expression statement and has no effect, so the result is
    sum_r r*x[:, min(c+r, cols-1)] / (2*sum_r r^2) along the FEATURE axis
    (edge padding of axis 1, processing.py:223).
    """
    rows, cols = feat.shape
    acc = np.zeros(feat.shape, dtype=feat.dtype)
    scale = 0
    padded = np.pad(feat, ((0, 0), (DeltaWindows, DeltaWindows)), 'edge')
    for r in range(1, DeltaWindows + 1):
        acc += r * padded[:, DeltaWindows + r:DeltaWindows + r + cols]
        scale += 2 * r ** 2
    return acc / scale


def cmvn(vec, variance_normalization=False):
    """Global mean (and variance) normalisation.  processing.py:239-271."""
    rows, cols = vec.shape
    centred = vec - np.tile(np.mean(vec, axis=0), (rows, 1))
    if not variance_normalization:
        return centred
    sd = np.tile(np.std(centred, axis=0), (rows, 1))
    return centred / (sd + CMVN_EPS)


def cmvnw(vec, win_size=301, variance_normalization=False):
    """Sliding-window mean (and variance) normalisation.  processing.py:274-327.

    One np.mean / np.std call per row over a symmetric-padded copy; the
    mean-subtracted array is float32 (processing.py:302) and so is the output.
    """
    rows, cols = vec.shape
    assert isinstance(win_size, int), "Size must be of type 'int'!"
    assert win_size % 2 == 1, "Windows size must be odd!"
    half = int((win_size - 1) / 2)
    padded = np.pad(vec, ((half, half), (0, 0)), 'symmetric')
    centred = np.zeros(np.shape(vec), dtype=np.float32)
    for i in range(rows):
        centred[i, :] = vec[i, :] - np.mean(padded[i:i + win_size, :], axis=0)
    if not variance_normalization:
        return centred
    out = np.zeros(np.shape(vec), dtype=np.float32)
    padded2 = np.pad(centred, ((half, half), (0, 0)), 'symmetric')
    for i in range(rows):
        sd = np.std(padded2[i:i + win_size, :], axis=0)
        out[i, :] = centred[i, :] / (sd + CMVN_EPS)
    return out


# ----------------------------------------------------------------------------
# feature.py
# ----------------------------------------------------------------------------
def filterbank_edges(num_filter, coefficients, sampling_freq, low_freq=None,
                     high_freq=None):
    """Integer FFT-bin edges of the mel filterbank.  sfe/speechpy/feature.py:55-82.

    Note ``low_freq or 300`` (0/None -> 300 Hz, feature.py:56) and the
    ``coefficients + 1`` multiplier (feature.py:77-82).
    """
    high_freq = high_freq or sampling_freq / 2
    low_freq = low_freq or 300
    assert high_freq <= sampling_freq / 2, (
        "High frequency cannot be greater than half of the sampling frequency!")
    assert low_freq >= 0, "low frequency cannot be less than zero!"
    mels = np.linspace(frequency_to_mel(low_freq), frequency_to_mel(high_freq),
                       num_filter + 2)
    hertz = mel_to_frequency(mels)
    return (np.floor((coefficients + 1) * hertz / sampling_freq)).astype(int)


def filterbanks(num_filter, coefficients, sampling_freq, low_freq=None,
                high_freq=None):
    """Dense (num_filter, coefficients) triangular bank.  feature.py:33-99."""
    edges = filterbank_edges(num_filter, coefficients, sampling_freq, low_freq,
                             high_freq)
    bank = np.zeros([num_filter, coefficients])
    for i in range(num_filter):
        lo, mid, hi = int(edges[i]), int(edges[i + 1]), int(edges[i + 2])
        pts = np.linspace(lo, hi, num=hi - lo + 1)
        bank[i, lo:hi + 1] = triangle(pts, left=lo, middle=mid, right=hi)
    return bank


def mfe(signal, sampling_frequency, frame_length=0.020, frame_stride=0.01,
        num_filters=40, fft_length=512, low_frequency=0, high_frequency=None):
    """Mel filterbank energies + per-frame energy.  feature.py:156-219."""
    signal = signal.astype(float)
    frames = stack_frames(signal, sampling_frequency=sampling_frequency,
                          frame_length=frame_length, frame_stride=frame_stride,
                          filter=lambda x: np.ones((x,)), zero_padding=False)
    high_frequency = high_frequency or sampling_frequency / 2
    spec = power_spectrum(frames, fft_length)
    coefficients = spec.shape[1]
    energies = zero_handling(np.sum(spec, 1))
    bank = filterbanks(num_filters, coefficients, sampling_frequency,
                       low_frequency, high_frequency)
    feats = zero_handling(np.dot(spec, bank.T))
    return feats, energies


def lmfe(signal, sampling_frequency, frame_length=0.020, frame_stride=0.01,
         num_filters=40, fft_length=512, low_frequency=0, high_frequency=None):
    """ln(mfe).  feature.py:222-258."""
    feats, _ = mfe(signal, sampling_frequency=sampling_frequency,
                   frame_length=frame_length, frame_stride=frame_stride,
                   num_filters=num_filters, fft_length=fft_length,
                   low_frequency=low_frequency, high_frequency=high_frequency)
    return np.log(feats)


def mfcc(signal, sampling_frequency, frame_length=0.020, frame_stride=0.01,
         num_cepstral=13, num_filters=40, fft_length=512, low_frequency=0,
         high_frequency=None, dc_elimination=True):
    """DCT-II (ortho) of ln(mfe); c0 <- ln(frame energy).  feature.py:102-153."""
    feats, energies = mfe(signal, sampling_frequency=sampling_frequency,
                          frame_length=frame_length, frame_stride=frame_stride,
                          num_filters=num_filters, fft_length=fft_length,
                          low_frequency=low_frequency,
                          high_frequency=high_frequency)
    if len(feats) == 0:
        return np.empty((0, num_cepstral))
    ceps = _dct(np.log(feats), type=2, axis=-1, norm='ortho')[:, :num_cepstral]
    if dc_elimination:
        ceps[:, 0] = np.log(energies)
    return ceps


def extract_derivative_feature(feature):
    """(F, M) -> (F, M, 3) = [x, D(x), D(D(x))].  feature.py:261-282."""
    d1 = derivative_extraction(feature, DeltaWindows=2)
    d2 = derivative_extraction(d1, DeltaWindows=2)
    return np.concatenate(
        (feature[:, :, None], d1[:, :, None], d2[:, :, None]), axis=2)


# ----------------------------------------------------------------------------
# Convenience for tests at sizes where the per-row loop is too slow.  NOT the
# timed baseline.  Validated against cmvnw() above in tests/test_oracle.py.
# ----------------------------------------------------------------------------
def cmvnw_vectorised(vec, win_size=301, variance_normalization=False):
    """Same arithmetic as cmvnw() with the row loop replaced by a strided view.

    np.mean / np.std over axis 1 of the (rows, win, cols) view perform the
    reductions in the same dtype as the reference (float64 for the mean stage,
    float32 for the std stage); results agree with the loop to ~1e-6.
    """
    from numpy.lib.stride_tricks import sliding_window_view
    half = int((win_size - 1) / 2)
    padded = np.pad(vec, ((half, half), (0, 0)), 'symmetric')
    centred = np.empty(vec.shape, dtype=np.float32)
    step = 2048
    for s in range(0, vec.shape[0], step):
        e = min(vec.shape[0], s + step)
        w = sliding_window_view(padded[s:e + win_size - 1], win_size, axis=0)
        centred[s:e] = vec[s:e] - np.mean(w, axis=-1)
    if not variance_normalization:
        return centred
    out = np.empty(vec.shape, dtype=np.float32)
    padded2 = np.pad(centred, ((half, half), (0, 0)), 'symmetric')
    for s in range(0, vec.shape[0], step):
        e = min(vec.shape[0], s + step)
        w = sliding_window_view(padded2[s:e + win_size - 1], win_size, axis=0)
        sd = np.std(w.astype(np.float64), axis=-1)
        out[s:e] = centred[s:e] / (sd + CMVN_EPS)
    return out


# ----------------------------------------------------------------------------
# the DataLoader transforms that follow lmfe (utils.py:325-397)
# ----------------------------------------------------------------------------
def cmvn_transform(feature, derivative, normalize):
    """``utils.CMVN.__call__`` (utils.py:383-397) with c.DERIVATIVE / c.NORMALIZE as arguments."""
    if derivative:
        feature = extract_derivative_feature(feature)
        if normalize:
            for k in range(3):
                feature[:, :, k] = cmvn(feature[:, :, k], variance_normalization=True)
    elif normalize:
        feature = cmvn(feature, variance_normalization=True)
    return feature


def feature_cube(feature, idx, num_frames):
    """``utils.FeatureCube.__call__`` (utils.py:354-373, feature (F, M)) and
    ``FeatureCube3C.__call__`` (utils.py:335-350, feature (F, M, 3)) with the random crop
    starts ``idx`` given: float32 (1, crops, frames, M) / (3, crops, frames, M)."""
    idx = np.asarray(idx)
    if feature.ndim == 3:
        f = feature.transpose(2, 0, 1)
        cube = np.zeros((f.shape[0], len(idx), num_frames, f.shape[2]), dtype=np.float32)
        for num, index in enumerate(idx):
            cube[:, num, :, :] = f[:, index:index + num_frames, :]
        return cube
    cube = np.zeros((len(idx), num_frames, feature.shape[1]), dtype=np.float32)
    for num, index in enumerate(idx):
        cube[num, :, :] = feature[index:index + num_frames, :]
    return cube[None, :, :, :]
