"""from speechpy import processing drop-in (speech_feature_extraction/tests/test_speechpy.py:8-10)."""
from speaker_verification_b200.speechpy.processing import *  # noqa: F401,F403
from speaker_verification_b200.speechpy.processing import __all__  # noqa: F401
