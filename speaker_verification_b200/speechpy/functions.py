"""Scalar helpers of the mel filterbank (host side, float64).

Mirror of speech_feature_extraction/speechpy/functions.py.  These stay on the host
on purpose: the filterbank's integer bin edges depend on float64 ``log`` / ``exp``
results that land within one ulp of an integer, so they are
evaluated with exactly the reference's numpy expressions and handed to the device
as data.  ``zero_handling`` is provided for API compatibility; inside the CUDA path
the same exact-zero test is applied in the kernels.
"""
import numpy as np

__all__ = ["frequency_to_mel", "mel_to_frequency", "triangle", "zero_handling"]


def frequency_to_mel(f):
    """Hz -> mel (functions.py:26-32)."""
    return 1127 * np.log(1 + f / 700.)


def mel_to_frequency(mel):
    """mel -> Hz (functions.py:35-41)."""
    return 700 * (np.exp(mel / 1127.0) - 1)


def triangle(x, left, middle, right):
    """Triangular window sampled at x (functions.py:44-52): rising on (left, middle],
    falling on [middle, right), zero elsewhere; the falling branch wins at x == middle."""
    x = np.asarray(x)
    out = np.zeros(x.shape)
    rising = np.logical_and(x > left, x <= middle)
    out[rising] = (x[rising] - left) / (middle - left)
    falling = np.logical_and(x >= middle, x < right)
    out[falling] = (right - x[falling]) / (right - middle)
    return out


def zero_handling(x):
    """Replace exact zeros by float64 eps (functions.py:55-62)."""
    return np.where(x == 0, np.finfo(float).eps, x)
