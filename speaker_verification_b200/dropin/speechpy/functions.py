"""from speechpy import functions drop-in (speech_feature_extraction/tests/test_speechpy.py:8-10)."""
from speaker_verification_b200.speechpy.functions import *  # noqa: F401,F403
from speaker_verification_b200.speechpy.functions import __all__  # noqa: F401
