"""speaker_verification_b200 -- B200-native drop-in for the speechpy front end and the
cosine / EER scoring of MingmChen/Speaker_Verification.

Layout
  csrc/            CUDA kernels (sm_100a) and the C-ABI of libspeechpy_b200.so
  _lib.py          ctypes binding (no CPU fallback)
  speechpy/        mirror of the reference's speechpy modules (processing, feature, functions)
  dropin/          import shims: ``speechpy`` and ``speech_feature_extraction.speechpy``
  batch.py         ragged-batch API, length bucketing, multi-GPU sharding
  scoring.py       cosine score matrix, trial scoring, EER/AUC (evaluation.py mirror)
"""
import os

__version__ = "0.1.0"


def dropin_path() -> str:
    """Directory to put on sys.path so that ``import speechpy`` and
    ``import speech_feature_extraction.speechpy.feature`` resolve to this package."""
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "dropin")


def install_dropin() -> None:
    import sys
    p = dropin_path()
    if p not in sys.path:
        sys.path.insert(0, p)
