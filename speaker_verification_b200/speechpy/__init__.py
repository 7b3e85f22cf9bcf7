"""B200-native mirror of the reference's bundled speechpy package.

Same module names, function names, positional order, defaults, return dtypes and
error behaviour as ``speech_feature_extraction/speechpy/{processing,feature,functions}.py``
of MingmChen/Speaker_Verification; every numeric function calls
``libspeechpy_b200.so`` through ctypes (numpy arrays in and out).
"""
from . import feature, functions, processing  # noqa: F401

__all__ = ["processing", "feature", "functions"]
