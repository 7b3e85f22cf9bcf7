"""speech_feature_extraction.speechpy.feature drop-in (load_data.py:5, utils.py:6)."""
from speaker_verification_b200.speechpy.feature import *  # noqa: F401,F403
from speaker_verification_b200.speechpy.feature import __all__  # noqa: F401
